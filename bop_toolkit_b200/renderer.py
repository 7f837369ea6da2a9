"""Renderer interface of the reference (bop_toolkit_lib/renderer.py:11-104) with the CUDA depth
rasteriser as back-end: create_renderer(width, height, 'cuda', mode='depth')."""
import numpy as np

from . import batch
from . import inout


class Renderer(object):
  """Abstract renderer (bop_toolkit_lib/renderer.py:11-74)."""

  def __init__(self, width, height):
    self.width = width
    self.height = height
    self.light_cam_pos = (0, 0, 0)
    self.light_color = (1.0, 1.0, 1.0)
    self.light_ambient_weight = 0.5
    self.light_diffuse_weight = 1.0
    self.light_specular_weight = 0.0
    self.light_specular_shininess = 0.0

  def set_light_cam_pos(self, light_cam_pos):
    self.light_cam_pos = light_cam_pos

  def set_light_ambient_weight(self, light_ambient_weight):
    self.light_ambient_weight = light_ambient_weight

  def add_object(self, obj_id, model_path, **kwargs):
    raise NotImplementedError

  def remove_object(self, obj_id):
    raise NotImplementedError

  def render_object(self, obj_id, R, t, fx, fy, cx, cy):
    raise NotImplementedError


class RendererCuda(Renderer):
  """Depth-only renderer on the hand-written sm_100a rasteriser (K1).

  Pixel (r, c) holds the nearest-surface camera Z (mm) along the ray through (c+0.5, r+0.5),
  0 = background -- the semantics of RendererVispy/RendererPython in 'depth' mode
  (renderer_vispy.py:433-466,509-538).  Unlike RendererVispy it is not a
  process-wide singleton: every instance owns its models.
  """

  def __init__(self, width, height, mode='depth', shading='phong', bg_color=(0.0, 0.0, 0.0, 0.0),
               device=None):
    super(RendererCuda, self).__init__(width, height)
    if mode != 'depth':
      raise ValueError("RendererCuda supports mode='depth' only (the VSD path; RGB shading is "
                       "out of scope).")
    self.mode = mode
    self.model_set = batch.ModelSet(device)
    self._removed = set()

  def add_object(self, obj_id, model_path=None, **kwargs):
    """Loads a PLY (`model_path`) or takes in-memory `pts=`/`faces=` (+ optional `diameter=`,
    `syms=`) and makes the model resident on the GPU."""
    if 'pts' in kwargs:
      pts, faces = kwargs['pts'], kwargs['faces']
    else:
      m = inout.load_ply(model_path)
      pts, faces = m['pts'], m['faces']
    self.model_set.add(obj_id, pts, faces, kwargs.get('syms'), kwargs.get('diameter', 0.0))
    self._removed.discard(obj_id)

  def remove_object(self, obj_id):
    if obj_id not in self.model_set or obj_id in self._removed:
      raise KeyError(obj_id)
    self._removed.add(obj_id)

  def render_object(self, obj_id, R, t, fx, fy, cx, cy):
    if obj_id not in self.model_set or obj_id in self._removed:
      raise KeyError(obj_id)
    pose = batch.pack_poses(np.asarray(R, np.float64).reshape(1, 3, 3),
                            np.asarray(t, np.float64).reshape(1, 3))
    cam = np.array([[fx, fy, cx, cy]], dtype=np.float64)
    d = batch.render_batch(self.model_set, self.model_set.index_of(obj_id), pose, cam,
                           self.height, self.width)
    return {'depth': d[0].cpu().numpy()}

  def render_objects(self, obj_ids, Rs, ts, cams):
    """Batched variant: [B,H,W] float32 device tensor."""
    poses = batch.pack_poses(Rs, ts)
    return batch.render_batch(self.model_set, self.model_set.indices(list(obj_ids)), poses,
                              np.asarray(cams, np.float64), self.height, self.width)


def create_renderer(width, height, renderer_type='cuda', mode='depth', shading='phong',
                    bg_color=(0.0, 0.0, 0.0, 0.0)):
  """Factory with the reference's signature (renderer.py:77-104).  'cuda' is the only back-end
  of this package; the reference's 'vispy' / 'python' / 'cpp' types need a graphics stack or the
  un-vendored bop_renderer and are not provided."""
  if renderer_type == 'cuda':
    return RendererCuda(width, height, mode, shading, bg_color)
  raise ValueError('Unknown renderer type.')

"""Drop-in replacement for bop_toolkit_lib.pose_error (vsd, mssd, mspd, add, adi, re, te, proj,
cus, cou_mask, cou_bb, cou_bb_proj).

Same positional signatures, numpy in / Python float (or list of floats) out, same exceptions
(bop_toolkit_lib/pose_error.py:17-232).  Each call is a batch-of-1 wrapper over the batched
CUDA path (batch.py -> libbop_b200.so); there is no CPU fallback.  Models passed as `pts`
arrays are uploaded once and cached by content fingerprint.
"""
import hashlib

import numpy as np
import torch

from . import _lib
from . import batch

_DEVICE = None
_model_cache = {}      # fingerprint -> (ModelSet, row)
_MAX_CACHE = 64


def set_device(device):
  global _DEVICE
  _DEVICE = torch.device(device)
  _model_cache.clear()


def _device():
  return _DEVICE if _DEVICE is not None else torch.device('cuda:0')


def _fingerprint(arr):
  a = np.ascontiguousarray(arr)
  return (a.shape, a.dtype.str, hashlib.blake2b(a.view(np.uint8).reshape(-1), digest_size=16).digest())


def _model_for(pts, syms=None):
  """Device model for an (pts, syms) pair as passed to pose_error.* ."""
  pts = np.asarray(pts)
  assert pts.ndim == 2 and pts.shape[1] == 3   # misc.py:102,274
  key = [_fingerprint(pts)]
  sym_arr = None
  if syms is not None:
    sym_arr = batch.pack_syms(syms)
    key.append(_fingerprint(sym_arr))
  key = tuple(key)
  hit = _model_cache.get(key)
  if hit is None:
    if len(_model_cache) >= _MAX_CACHE:
      _model_cache.pop(next(iter(_model_cache)))
    ms = batch.ModelSet(_device())
    ms.add(0, pts.astype(np.float64), None, sym_arr, 0.0)
    hit = _model_cache[key] = ms
  return hit


def _pose(R, t):
  R = np.asarray(R, dtype=np.float64)
  t = np.asarray(t, dtype=np.float64)
  return batch.pack_poses(R.reshape(1, 3, 3), t.reshape(1, 3))


def _scalar(t):
  return float(t.cpu()[0])


def mssd(R_est, t_est, R_gt, t_gt, pts, syms):
  """Maximum Symmetry-Aware Surface Distance (pose_error.py:96-118)."""
  ms = _model_for(pts, syms)
  return _scalar(batch.mssd_batch(ms, 0, _pose(R_est, t_est), _pose(R_gt, t_gt)))


def mspd(R_est, t_est, R_gt, t_gt, K, pts, syms):
  """Maximum Symmetry-Aware Projection Distance (pose_error.py:121-144)."""
  ms = _model_for(pts, syms)
  K = np.asarray(K, dtype=np.float64).reshape(1, 9)
  return _scalar(batch.mspd_batch(ms, 0, _pose(R_est, t_est), _pose(R_gt, t_gt), K))


def add(R_est, t_est, R_gt, t_gt, pts):
  """Average Distance of Model Points (pose_error.py:147-161)."""
  ms = _model_for(pts)
  return _scalar(batch.add_batch(ms, 0, _pose(R_est, t_est), _pose(R_gt, t_gt)))


def adi(R_est, t_est, R_gt, t_gt, pts):
  """Average Distance of Model Points, indistinguishable views (pose_error.py:164-184)."""
  ms = _model_for(pts)
  return _scalar(batch.adi_batch(ms, 0, _pose(R_est, t_est), _pose(R_gt, t_gt)))


def proj(R_est, t_est, R_gt, t_gt, K, pts):
  """Average projection distance (pose_error.py:217-232)."""
  ms = _model_for(pts)
  K = np.asarray(K, dtype=np.float64).reshape(1, 9)
  return _scalar(batch.proj_batch(ms, 0, _pose(R_est, t_est), _pose(R_gt, t_gt), K))


def re(R_est, R_gt):
  """Rotational error in degrees (pose_error.py:187-202)."""
  R_est, R_gt = np.asarray(R_est), np.asarray(R_gt)
  assert (R_est.shape == R_gt.shape == (3, 3))
  z = np.zeros(3)
  r, _ = batch.re_te_batch(_pose(R_est, z), _pose(R_gt, z), _device())
  return _scalar(r)


def te(t_est, t_gt):
  """Translational error (pose_error.py:205-214)."""
  t_est, t_gt = np.asarray(t_est), np.asarray(t_gt)
  assert (t_est.size == t_gt.size == 3)
  I = np.eye(3)
  _, t = batch.re_te_batch(_pose(I, t_est), _pose(I, t_gt), _device())
  return _scalar(t)


def vsd(R_est, t_est, R_gt, t_gt, depth_test, K, delta, taus,
        normalized_by_diameter, diameter, renderer, obj_id, cost_type='step'):
  """Visible Surface Discrepancy (pose_error.py:17-93).

  With the package's own RendererCuda the two renders and the pixel pass run fused on the GPU
  (nothing but `depth_test` is read from HBM).  With any other Renderer object (duck-typed, as
  in the reference) its render_object() is called and the fused K2 pixel pass consumes the
  returned depth images.
  """
  if cost_type not in _lib.COST_TYPES:
    raise ValueError('Unknown pixel matching cost.')
  K = np.asarray(K, dtype=np.float64)
  taus = list(taus)
  if len(taus) == 0:
    return []
  from .renderer import RendererCuda
  if isinstance(renderer, RendererCuda):
    ms = renderer.model_set
    row = ms.index_of(obj_id)
    # `diameter` is an argument of vsd, not a property of the renderer's model.
    err = batch.vsd_batch(ms, row, _pose(R_est, t_est), _pose(R_gt, t_gt),
                          np.asarray(depth_test), None, K.reshape(1, 9), delta, taus,
                          normalized_by_diameter, cost_type, 'bop19', False,
                          diameter_override=float(diameter))
  else:
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    depth_est = renderer.render_object(obj_id, R_est, t_est, fx, fy, cx, cy)['depth']
    depth_gt = renderer.render_object(obj_id, R_gt, t_gt, fx, fy, cx, cy)['depth']
    err = batch.vsd_from_depth_batch(
      np.asarray(depth_est, np.float32), np.asarray(depth_gt, np.float32),
      np.asarray(depth_test), None, K.reshape(1, 9), float(diameter), delta, taus,
      normalized_by_diameter, cost_type, 'bop19', device=_device())
  return [float(e) for e in err.cpu()[0]]


def _mask_stats(R_est, t_est, R_gt, t_gt, K, renderer, obj_id):
  """Silhouette statistics of the two poses (the shared front half of cus / cou_bb_proj)."""
  K = np.asarray(K, dtype=np.float64)
  from .renderer import RendererCuda
  if isinstance(renderer, RendererCuda):
    ms = renderer.model_set
    return batch.mask_overlap_batch(ms, ms.index_of(obj_id), _pose(R_est, t_est), _pose(R_gt, t_gt),
                                    K.reshape(1, 9), renderer.height, renderer.width)
  fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
  depth_est = renderer.render_object(obj_id, R_est, t_est, fx, fy, cx, cy)['depth']
  depth_gt = renderer.render_object(obj_id, R_gt, t_gt, fx, fy, cx, cy)['depth']
  return batch.mask_overlap_from_depth_batch(np.asarray(depth_est, np.float32) > 0,
                                             np.asarray(depth_gt, np.float32) > 0, _device())


def cus(R_est, t_est, R_gt, t_gt, K, renderer, obj_id):
  """Complement over Union of projected 2D masks (pose_error.py:256-286)."""
  return _scalar(_mask_stats(R_est, t_est, R_gt, t_gt, K, renderer, obj_id)['cus'])


def cou_mask(mask_est, mask_gt):
  """Complement over Union of 2D binary masks (pose_error.py:235-253)."""
  st = batch.mask_overlap_from_depth_batch(np.asarray(mask_est).astype(bool),
                                           np.asarray(mask_gt).astype(bool), _device())
  return _scalar(st['cus'])


def cou_bb(bb_est, bb_gt):
  """Complement over Union of 2D bounding boxes (x, y, w, h) (pose_error.py:289-297 with
  misc.iou, misc.py:236-263).  Four scalars per box: plain host arithmetic."""
  ax0, ay0, ax1, ay1 = bb_est[0], bb_est[1], bb_est[0] + bb_est[2], bb_est[1] + bb_est[3]
  bx0, by0, bx1, by1 = bb_gt[0], bb_gt[1], bb_gt[0] + bb_gt[2], bb_gt[1] + bb_gt[3]
  w_inter = min(ax1, bx1) - max(ax0, bx0)
  h_inter = min(ay1, by1) - max(ay0, by0)
  iou = 0.0
  if w_inter > 0 and h_inter > 0:
    area_inter = w_inter * h_inter
    iou = area_inter / float(bb_est[2] * bb_est[3] + bb_gt[2] * bb_gt[3] - area_inter)
  return 1.0 - iou


def cou_bb_proj(R_est, t_est, R_gt, t_gt, K, renderer, obj_id):
  """Complement over Union of projected 2D bounding boxes (pose_error.py:300-330).  Like the
  reference (`xs.min()` of an empty array), raises ValueError when a silhouette is empty."""
  st = _mask_stats(R_est, t_est, R_gt, t_gt, K, renderer, obj_id)
  counts = st['counts'].cpu()[0]
  if int(counts[2]) == 0 or int(counts[3]) == 0:
    raise ValueError('zero-size array to reduction operation minimum which has no identity')
  return _scalar(st['cou_bb'])

"""Batched pose-error API: whole shards of pose pairs per call, on one B200.

This is the fast path the drop-in facade (pose_error.py) and the driver (eval_calc_errors.py)
sit on.  torch is used for device memory, streams and pinned staging only; all arithmetic runs
in the hand-written sm_100a kernels behind the C ABI (include/bop_b200.h).
"""
import contextlib
import ctypes
import os
import threading

import numpy as np
import torch

from . import _lib


def _dev(device):
  return torch.device(device if device is not None else 'cuda:0')


def _stream_ptr():
  return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def to_device(x, dtype, device, shape=None):
  """numpy / torch / list -> contiguous device tensor of `dtype` (non-blocking when pinned)."""
  if isinstance(x, torch.Tensor):
    t = x
  else:
    arr = np.ascontiguousarray(x)
    if not arr.flags.writeable:      # e.g. arrays out of np.load / broadcast views
      arr = arr.copy()
    t = torch.from_numpy(arr)
  if t.dtype != dtype:
    t = t.to(dtype)
  if t.device != device:
    t = t.to(device, non_blocking=True)
  t = t.contiguous()
  if shape is not None:
    t = t.reshape(shape)
  return t


def pack_poses(Rs, ts):
  """(R 3x3, t 3x1|3,) pairs -> [P,12] float64 array (R row-major then t)."""
  Rs = np.asarray(Rs, dtype=np.float64).reshape(-1, 9)
  ts = np.asarray(ts, dtype=np.float64).reshape(-1, 3)
  return np.ascontiguousarray(np.hstack([Rs, ts]))


def pack_syms(syms):
  """List of {'R','t'} dicts (misc.get_symmetry_transformations) -> [S,12] float64."""
  out = np.empty((len(syms), 12), dtype=np.float64)
  for i, s in enumerate(syms):
    out[i, :9] = np.asarray(s['R'], dtype=np.float64).reshape(9)
    out[i, 9:] = np.asarray(s['t'], dtype=np.float64).reshape(3)
  return out


def closed_and_oriented(faces, n_pts):
  """True iff every directed edge of the triangle list occurs exactly once and its reverse occurs
  too: a closed surface whose faces are wound consistently (necessary for back-face culling to
  leave the depth image unchanged)."""
  f = np.asarray(faces, dtype=np.int64).reshape(-1, 3)
  if f.shape[0] == 0:
    return False
  a = np.concatenate([f[:, 0], f[:, 1], f[:, 2]])
  b = np.concatenate([f[:, 1], f[:, 2], f[:, 0]])
  if np.any(a == b):
    return False
  fwd = a * int(n_pts) + b
  if np.unique(fwd).size != fwd.size:
    return False
  return bool(np.isin(b * int(n_pts) + a, fwd).all())


def _cull_probe_poses(diameter, n=24, seed=12345):
  """Poses around the object for the culling verification: random rotations at distances from
  1.2 to 8 diameters, on and off the optical axis."""
  rng = np.random.default_rng(seed)
  q = rng.normal(size=(n, 4))
  q /= np.linalg.norm(q, axis=1, keepdims=True)
  w, x, y, z = q.T
  R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], axis=1)
  d = max(float(diameter), 1e-3)
  tz = d * rng.uniform(1.2, 8.0, size=n)
  t = np.stack([tz * rng.uniform(-0.3, 0.3, size=n), tz * rng.uniform(-0.25, 0.25, size=n), tz], axis=1)
  return np.ascontiguousarray(np.hstack([R, t]))


class ModelSet(object):
  """Object models resident in HBM: the device-side counterpart of the reference's
  `models[obj_id]` (inout.load_ply), `models_sym[obj_id]` (misc.get_symmetry_transformations),
  `models_info[obj_id]['diameter']` and the renderer's vertex/index buffers
  (scripts/eval_calc_errors.py:157-191).  Replicated per GPU (a 5k-vertex model is ~0.3 MB)."""

  def __init__(self, device=None):
    _lib.require_cuda()
    self.device = _dev(device)
    self._rows = []      # ModelDev structs
    self._keep = []      # tensors kept alive
    self._index = {}     # obj_id -> row
    self._table = None
    self.max_pts = 1
    self.max_geom = 1      # max(n_pts, ceil(n_faces / 2)): sizes the rasteriser's geometry scratch
    self.max_syms = 1
    self.info = {}       # obj_id -> dict(n_pts, n_faces, n_syms, diameter)

  def add(self, obj_id, pts, faces=None, syms=None, diameter=0.0, cull=None):
    """pts: Nx3 float64 as pose_error.* receives them; faces: Fx3 ints (needed for VSD /
    rendering); syms: list of {'R','t'} or [S,12]; diameter: models_info 'diameter'.
    cull: False (the default, unless the environment variable BOP_B200_CULL=auto is set) never
    culls -- the reference's GL state (renderer_vispy.py:470), which every drop-in driver uses.
    'auto' is an opt-in: it enables back-face culling in the rasteriser only when the mesh is
    closed and consistently oriented (closed_and_oriented) AND renders of probe poses with
    culling have the same silhouettes and depths within 1e-5 relative of renders without
    (_verify_cull).  Culling halves the triangle work; on such a mesh the nearest surface along
    every ray is front-facing, so the depth image can only differ by float rounding where a front
    and a back face nearly coincide at a silhouette -- inside the 1e-3 depth tolerance, but not
    bit-identical, hence not the default."""
    if cull is None:
      cull = 'auto' if os.environ.get('BOP_B200_CULL', '').lower() == 'auto' else False
    pts = np.ascontiguousarray(np.asarray(pts, dtype=np.float64).reshape(-1, 3))
    n = pts.shape[0]
    if n == 0:
      raise ValueError('model without points')
    p32 = pts.astype(np.float32)
    f32_exact = bool(np.array_equal(p32.astype(np.float64), pts))
    with torch.cuda.device(self.device):
      if f32_exact:
        d_pts = torch.from_numpy(p32).to(self.device)
      else:
        d_pts = torch.from_numpy(pts).to(self.device)
      v4 = np.zeros((n, 4), dtype=np.float32)
      v4[:, :3] = p32
      d_v4 = torch.from_numpy(v4).to(self.device)
      row = _lib.ModelDev()
      row.pts = d_pts.data_ptr()
      row.verts4 = d_v4.data_ptr()
      row.pts_f64 = 0 if f32_exact else 1
      row.n_pts = n
      row.diameter = float(diameter)
      # camera centres within this distance of the model origin may be inside the mesh: no culling
      row.cull_min_dist = float(np.sqrt((p32.astype(np.float64) ** 2).sum(axis=1).max())) * 1.001
      keep = [d_pts, d_v4]
      nf = 0
      if faces is not None and len(faces):
        faces = np.asarray(faces).astype(np.int64).reshape(-1, 3)
        if faces.min() < 0 or faces.max() >= n:
          raise ValueError('face index out of range')
        nf = faces.shape[0]
        f4 = np.zeros((nf, 4), dtype=np.int32)
        f4[:, :3] = faces
        d_f4 = torch.from_numpy(f4).to(self.device)
        row.faces4 = d_f4.data_ptr()
        row.n_faces = nf
        keep.append(d_f4)
      ns = 0
      if syms is not None and len(syms):
        s = pack_syms(syms) if isinstance(syms, (list, tuple)) else \
          np.ascontiguousarray(np.asarray(syms, dtype=np.float64).reshape(-1, 12))
        ns = s.shape[0]
        d_s = torch.from_numpy(s).to(self.device)
        row.syms = d_s.data_ptr()
        row.n_syms = ns
        keep.append(d_s)
    if obj_id in self._index:
      self._rows[self._index[obj_id]] = row
    else:
      self._index[obj_id] = len(self._rows)
      self._rows.append(row)
    self._keep.append(keep)
    self.max_pts = max(self.max_pts, n)
    self.max_geom = max(self.max_geom, n, (nf + 1) // 2)
    self.max_syms = max(self.max_syms, ns)
    self.info[obj_id] = dict(n_pts=n, n_faces=nf, n_syms=ns, diameter=float(diameter), cull=0)
    self._table = None
    if cull == 'auto' and nf > 0 and closed_and_oriented(faces, n):
      self._verify_cull(obj_id, float(diameter) if diameter else float(np.linalg.norm(np.ptp(pts, axis=0))))
    return self._index[obj_id]

  def _verify_cull(self, obj_id, size):
    """Renders probe poses without culling and with either sign; keeps a sign only if every
    silhouette is identical and every depth agrees to 1e-5 relative.  (Depths are not always
    bit-identical: where a front and a back face nearly coincide at a silhouette pixel, the
    unculled z-test keeps whichever of the two float roundings is smaller.)"""
    row = self._rows[self._index[obj_id]]
    poses = _cull_probe_poses(size)
    cam = np.array([572.4114, 573.57043, 325.2611, 242.04899])
    ref = render_batch(self, self._index[obj_id], poses, cam, 480, 640)
    if not bool((ref > 0).any()):
      return
    for sign in (1, -1):
      row.cull = sign
      self._table = None
      got = render_batch(self, self._index[obj_id], poses, cam, 480, 640)
      if torch.equal(got > 0, ref > 0) and bool(((got - ref).abs() <= 1e-5 * ref).all()):
        self.info[obj_id]['cull'] = sign
        return
    row.cull = 0
    self._table = None

  def index_of(self, obj_id):
    return self._index[obj_id]

  def indices(self, obj_ids):
    return np.fromiter((self._index[o] for o in obj_ids), dtype=np.int32, count=len(obj_ids))

  def __contains__(self, obj_id):
    return obj_id in self._index

  @property
  def table(self):
    """Device copy of the bop_model_dev table."""
    if self._table is None:
      buf = b''.join(bytes(r) for r in self._rows)
      host = torch.frombuffer(bytearray(buf), dtype=torch.uint8)
      self._table = host.to(self.device)
    return self._table


class Workspace(object):
  """Grow-only device scratch buffer (256-byte aligned by the caching allocator), ONE per device:
  every batched call on the device uses it, so calls are ordered behind each other -- a call
  issued on another stream than the previous one first waits for that stream (the library keeps
  its work counters in the buffer; overlapping calls would race)."""

  def __init__(self, device):
    self.device = device
    self.buf = None
    self.stream = None

  def get(self, nbytes):
    cur = torch.cuda.current_stream(self.device)
    if self.stream is not None and self.stream != cur and not torch.cuda.is_current_stream_capturing():
      cur.wait_stream(self.stream)
    if not torch.cuda.is_current_stream_capturing():
      self.stream = cur
    if self.buf is None or self.buf.numel() < nbytes:
      self.buf = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
    return self.buf


_workspaces = {}
_private = threading.local()


@contextlib.contextmanager
def private_workspace(ws):
  """Batched calls inside the block (this thread) use `ws`, a Workspace the caller owns, instead
  of the device's shared one: what calls that are meant to run concurrently on several streams
  need -- one Workspace per stream."""
  prev = getattr(_private, 'ws', None)
  _private.ws = ws
  try:
    yield ws
  finally:
    _private.ws = prev


def _workspace(device, nbytes):
  ws = getattr(_private, 'ws', None)
  if ws is not None:
    if ws.device != device:
      raise ValueError('private workspace lives on %s, the call runs on %s' % (ws.device, device))
    return ws.get(nbytes)
  key = (device.type, device.index)
  ws = _workspaces.get(key)
  if ws is None:
    ws = _workspaces[key] = Workspace(device)
  return ws.get(nbytes)


def _obj_idx(models, obj, n):
  if isinstance(obj, (int, np.integer)):
    return torch.full((n,), int(obj), dtype=torch.int32, device=models.device)
  return to_device(obj, torch.int32, models.device, (n,))


def _poses(models, pe, pg):
  pe = to_device(pe, torch.float64, models.device)
  pg = to_device(pg, torch.float64, models.device)
  if pe.numel() % 12 or pe.shape != pg.shape:
    raise ValueError('poses must be [P,12] (R row-major, then t)')
  return pe, pg, pe.numel() // 12


def re_te_batch(pose_est, pose_gt, device=None):
  """pose_error.re / te for P pairs -> (re[P] deg, te[P] mm) device tensors."""
  _lib.require_cuda()
  dev = _dev(device)
  pe = to_device(pose_est, torch.float64, dev)
  pg = to_device(pose_gt, torch.float64, dev)
  n = pe.numel() // 12
  re = torch.empty(n, dtype=torch.float64, device=dev)
  te = torch.empty(n, dtype=torch.float64, device=dev)
  with torch.cuda.device(dev):
    _lib.check(_lib.lib().bop_re_te_dev(pe.data_ptr(), pg.data_ptr(), n, re.data_ptr(),
                                        te.data_ptr(), _stream_ptr()))
  return re, te


def _points_call(fn_name, models, obj, pose_est, pose_gt, K, max_n, gate):
  pe, pg, n = _poses(models, pose_est, pose_gt)
  if n and fn_name in ('bop_mssd_dev', 'bop_mspd_dev'):
    _require_syms(models, obj)
  oi = _obj_idx(models, obj, n)
  out = torch.empty(n, dtype=torch.float64, device=models.device)
  fn = getattr(_lib.lib(), fn_name)
  with torch.cuda.device(models.device):
    if K is not None:
      Kd = to_device(K, torch.float64, models.device)
      if Kd.numel() == 9:
        Kd = Kd.reshape(1, 9).expand(n, 9).contiguous()
      if Kd.numel() != 9 * n:
        raise ValueError('K must be [P,9] or a single 3x3')
      rc = fn(models.table.data_ptr(), oi.data_ptr(), pe.data_ptr(), pg.data_ptr(), Kd.data_ptr(),
              n, int(max_n), out.data_ptr(), _stream_ptr())
    else:
      rc = fn(models.table.data_ptr(), oi.data_ptr(), pe.data_ptr(), pg.data_ptr(), n,
              int(max_n), int(bool(gate)), out.data_ptr(), _stream_ptr())
  _lib.check(rc)
  return out


def _require_syms(models, obj):
  """pose_error.mssd / mspd take min() over the symmetry set (pose_error.py:112-118,137-144): an
  empty set raises ValueError there, and must not silently score +inf here."""
  if all(i['n_syms'] > 0 for i in models.info.values()):
    return          # the usual case: nothing to look up (and no device -> host copy of `obj`)
  if isinstance(obj, (int, np.integer)):
    rows = [int(obj)]
  else:
    rows = np.unique(obj.cpu().numpy() if isinstance(obj, torch.Tensor) else np.asarray(obj)).tolist()
  by_row = {r: o for o, r in models._index.items()}
  for r in rows:
    if r not in by_row:
      raise ValueError('model row %d is not in the ModelSet' % r)
    if models.info[by_row[r]]['n_syms'] == 0:
      raise ValueError('min() arg is an empty sequence: model %r has no symmetry transformations' % (by_row[r],))


def mssd_batch(models, obj, pose_est, pose_gt, gate=False):
  """pose_error.mssd for P pairs; gate=True applies the driver's sphere gate (-> inf)."""
  return _points_call('bop_mssd_dev', models, obj, pose_est, pose_gt, None, models.max_syms, gate)


def mspd_batch(models, obj, pose_est, pose_gt, K):
  """pose_error.mspd for P pairs; K: [P,9] per pair or one 3x3."""
  return _points_call('bop_mspd_dev', models, obj, pose_est, pose_gt, K, models.max_syms, False)


def add_batch(models, obj, pose_est, pose_gt, gate=False):
  return _points_call('bop_add_dev', models, obj, pose_est, pose_gt, None, models.max_pts, gate)


def adi_batch(models, obj, pose_est, pose_gt, gate=False):
  return _points_call('bop_adi_dev', models, obj, pose_est, pose_gt, None, models.max_pts, gate)


def proj_batch(models, obj, pose_est, pose_gt, K):
  return _points_call('bop_proj_dev', models, obj, pose_est, pose_gt, K, models.max_pts, False)


def make_vsd_params(height, width, delta, taus, normalized_by_diameter=True, cost_type='step',
                    visib_mode='bop19', gate=False, depth_dtype=_lib.DEPTH_F32):
  if cost_type not in _lib.COST_TYPES:
    raise ValueError('Unknown pixel matching cost.')  # pose_error.py:88
  if visib_mode not in _lib.VISIB_MODES:
    raise ValueError('Unknown visibility mode.')      # visibility.py:40
  taus = [float(t) for t in taus]
  if not 1 <= len(taus) <= _lib.BOP_MAX_TAUS:
    raise ValueError('between 1 and %d taus per launch' % _lib.BOP_MAX_TAUS)
  p = _lib.VsdParams()
  for i, t in enumerate(taus):
    p.taus[i] = t
  p.delta = float(delta)
  p.n_taus = len(taus)
  p.normalized_by_diameter = int(bool(normalized_by_diameter))
  p.cost_type = _lib.COST_TYPES[cost_type]
  p.visib_mode = _lib.VISIB_MODES[visib_mode]
  p.height, p.width = int(height), int(width)
  p.gate = int(bool(gate))
  p.depth_dtype = depth_dtype
  return p


def _depth_input(depth_test, device, zero_copy=False):
  """-> (tensor whose data_ptr() the kernel reads, dtype code).  zero_copy: a PINNED host tensor
  is handed to the kernel as is (unified addressing) and read across PCIe on demand."""
  if zero_copy and isinstance(depth_test, torch.Tensor) and not depth_test.is_cuda:
    if not depth_test.is_pinned() or not depth_test.is_contiguous():
      raise ValueError('zero-copy depth needs a contiguous pinned host tensor')
    if depth_test.dtype in (torch.uint16, torch.int16):
      dd, code = depth_test.view(torch.int16), _lib.DEPTH_U16
    elif depth_test.dtype == torch.float32:
      dd, code = depth_test, _lib.DEPTH_F32
    else:
      raise ValueError('zero-copy depth must be float32 or uint16')
    return (dd.unsqueeze(0) if dd.dim() == 2 else dd), code
  if isinstance(depth_test, np.ndarray):
    depth_test = torch.from_numpy(np.ascontiguousarray(depth_test))
  if depth_test.dtype == torch.uint16 or depth_test.dtype == torch.int16:
    dd, code = depth_test.view(torch.int16), _lib.DEPTH_U16
  else:
    dd, code = depth_test.to(torch.float32), _lib.DEPTH_F32
  if dd.dim() == 2:
    dd = dd.unsqueeze(0)
  if dd.device != device:
    dd = dd.to(device, non_blocking=True)
  return dd.contiguous(), code


def _tau_chunks(taus):
  taus = list(taus)
  return [taus[i:i + _lib.BOP_MAX_TAUS] for i in range(0, len(taus), _lib.BOP_MAX_TAUS)]


MAX_GROUP = 32   # kMaxGroup in csrc/vsd.cu


def group_pairs_by_gt(obj, pose_gt, im_idx, diameter=None):
  """Pairs that share (object, image, GT pose[, diameter]) -- the k estimates
  scripts/eval_calc_errors.py:281-289 pairs with one GT pose.  Host arrays in, -> (order, group_off):
  `order` sorts the pairs so that groups are consecutive (stable: original order inside a group),
  `group_off` (int32 [n_groups + 1]) cuts the sorted pairs into groups of at most MAX_GROUP, the
  layout vsd_batch(group_off=...) / bop_vsd_grouped_dev take."""
  obj = np.asarray(obj).reshape(-1)
  n = obj.shape[0]
  pg = np.ascontiguousarray(np.asarray(pose_gt, dtype=np.float64).reshape(n, 12))
  cols = [pg.view(np.int64).reshape(n, 12), obj.astype(np.int64).reshape(n, 1),
          np.asarray(im_idx).astype(np.int64).reshape(n, 1)]
  if diameter is not None:
    cols.append(np.ascontiguousarray(np.broadcast_to(np.asarray(diameter, np.float64), (n,))).view(np.int64).reshape(n, 1))
  key = np.ascontiguousarray(np.hstack(cols))
  _, inv = np.unique(key.view([('', key.dtype)] * key.shape[1]).reshape(n), return_inverse=True)
  inv = inv.reshape(-1)
  order = np.argsort(inv, kind='stable')
  sizes = np.bincount(inv)
  off = [0]
  for k in sizes:          # groups above MAX_GROUP are cut
    k = int(k)
    while k > 0:
      step = min(k, MAX_GROUP)
      off.append(off[-1] + step)
      k -= step
  return order, np.asarray(off, dtype=np.int32)


def vsd_batch(models, obj, pose_est, pose_gt, depth_test, im_idx, K, delta, taus,
              normalized_by_diameter=True, cost_type='step', visib_mode='bop19', gate=False,
              depth_scale=None, return_counts=False, diameter_override=None, zero_copy=False,
              group_off=None, raster_stream=None):
  """pose_error.vsd for P pairs with the fused raster + pixel-pass kernel.

  raster_stream (a torch stream): the geometry pass stays on the current stream, the raster +
  pixel pass and the results go to raster_stream (bop_vsd_split_dev): for callers that give the
  two passes an SM partition each (sm_partition.py) so that the PCIe-bound pixel pass of
  host-resident depth overlaps the next batch's geometry pass.  The returned tensors are ready
  on raster_stream; reuse the workspace (private_workspace) only behind it.

  group_off (host int32 [G + 1], from group_pairs_by_gt): consecutive pairs of a group share
  object, image, GT pose and diameter; the GT pose's geometry and raster work is then done once
  per group (bit-identical results).

  depth_test: [n_im,H,W] float32 mm (or uint16 raw + depth_scale[n_im]); im_idx: [P] image of
  each pair (None = pair i uses image i); K: [n_im,9].  Returns errors [P,len(taus)] float64
  (device tensor) and, if return_counts, int64 [P,2+T] = union, inter, step-cost counts.
  """
  pe, pg, n = _poses(models, pose_est, pose_gt)
  dev = models.device
  oi = _obj_idx(models, obj, n)
  dd, code = _depth_input(depth_test, dev, zero_copy)
  n_im, H, W = dd.shape
  Kd = to_device(K, torch.float64, dev).reshape(-1, 9)
  if Kd.shape[0] != n_im:
    if Kd.shape[0] == 1:
      Kd = Kd.expand(n_im, 9).contiguous()
    else:
      raise ValueError('K must be [n_im,9]')
  if im_idx is None:
    if n_im == 1:
      ii = torch.zeros(n, dtype=torch.int32, device=dev)
    else:
      ii = torch.arange(n, dtype=torch.int32, device=dev)
  else:
    if not isinstance(im_idx, torch.Tensor) or not im_idx.is_cuda:   # host indices: range check for free
      chk = np.asarray(im_idx.numpy() if isinstance(im_idx, torch.Tensor) else im_idx)
      if chk.size and (chk.min() < 0 or chk.max() >= n_im):
        raise ValueError('im_idx out of range [0, %d)' % n_im)
    ii = to_device(im_idx, torch.int32, dev, (n,))
  sc = None
  if code == _lib.DEPTH_U16:
    if depth_scale is None:
      depth_scale = np.ones(n_im, dtype=np.float32)
    sc = to_device(depth_scale, torch.float32, dev, (n_im,))
  diam = None
  if diameter_override is not None:
    if isinstance(diameter_override, torch.Tensor):
      diam = to_device(diameter_override, torch.float64, dev, (n,))
    else:
      diam = to_device(np.ascontiguousarray(np.broadcast_to(
        np.asarray(diameter_override, dtype=np.float64), (n,))), torch.float64, dev, (n,))
  outs, cnts = [], []
  wsb = _lib.lib().bop_vsd_workspace_bytes(n, H, W, models.max_geom)
  ws = _workspace(dev, wsb)
  with torch.cuda.device(dev):
    for chunk in _tau_chunks(taus):
      prm = make_vsd_params(H, W, delta, chunk, normalized_by_diameter, cost_type, visib_mode,
                            gate, code)
      err = torch.empty((n, len(chunk)), dtype=torch.float64, device=dev)
      cnt = torch.empty((n, 2 + len(chunk)), dtype=torch.int64, device=dev) if return_counts else None
      if raster_stream is not None:
        if group_off is not None:
          raise ValueError('raster_stream and group_off cannot be combined')
        if outs:   # the previous tau chunk's raster pass still reads the workspace
          torch.cuda.current_stream().wait_stream(raster_stream)
        for t_ in (err, cnt, dd, sc):
          if t_ is not None and t_.is_cuda:
            t_.record_stream(raster_stream)
        rc = _lib.lib().bop_vsd_split_dev(
          models.table.data_ptr(), oi.data_ptr(), pe.data_ptr(), pg.data_ptr(), dd.data_ptr(),
          sc.data_ptr() if sc is not None else None, ii.data_ptr(), Kd.data_ptr(),
          diam.data_ptr() if diam is not None else None, n, models.max_geom,
          ctypes.byref(prm), err.data_ptr(), cnt.data_ptr() if cnt is not None else None,
          ws.data_ptr(), wsb, _stream_ptr(), ctypes.c_void_p(raster_stream.cuda_stream))
      elif group_off is not None:
        goff = np.ascontiguousarray(group_off, dtype=np.int32)
        rc = _lib.lib().bop_vsd_grouped_dev(
          models.table.data_ptr(), oi.data_ptr(), pe.data_ptr(), pg.data_ptr(), dd.data_ptr(),
          sc.data_ptr() if sc is not None else None, ii.data_ptr(), Kd.data_ptr(),
          diam.data_ptr() if diam is not None else None, n, goff.ctypes.data, int(goff.shape[0]) - 1,
          models.max_geom, ctypes.byref(prm), err.data_ptr(), cnt.data_ptr() if cnt is not None else None,
          ws.data_ptr(), wsb, _stream_ptr())
      else:
        rc = _lib.lib().bop_vsd_dev(
          models.table.data_ptr(), oi.data_ptr(), pe.data_ptr(), pg.data_ptr(), dd.data_ptr(),
          sc.data_ptr() if sc is not None else None, ii.data_ptr(), Kd.data_ptr(),
          diam.data_ptr() if diam is not None else None, n, models.max_geom,
          ctypes.byref(prm), err.data_ptr(), cnt.data_ptr() if cnt is not None else None,
          ws.data_ptr(), wsb, _stream_ptr())
      _lib.check(rc)
      outs.append(err)
      cnts.append(cnt)
  # several tau chunks are joined on the stream their values are ready on
  with torch.cuda.stream(raster_stream) if raster_stream is not None and len(outs) > 1 else contextlib.nullcontext():
    err = outs[0] if len(outs) == 1 else torch.cat(outs, dim=1)
    if return_counts:
      cnt = cnts[0] if len(cnts) == 1 else torch.cat([cnts[0]] + [c[:, 2:] for c in cnts[1:]], dim=1)
      return err, cnt
  return err


def vsd_from_depth_batch(depth_est, depth_gt, depth_test, im_idx, K, diameters, delta, taus,
                         normalized_by_diameter=True, cost_type='step', visib_mode='bop19',
                         depth_scale=None, return_counts=False, device=None):
  """K2 alone: pose_error.vsd lines 48-91 on pre-rendered depth images [P,H,W] float32."""
  _lib.require_cuda()
  dev = _dev(device if device is not None else
             (depth_est.device if isinstance(depth_est, torch.Tensor) and depth_est.is_cuda else None))
  de = to_device(depth_est, torch.float32, dev)
  dg = to_device(depth_gt, torch.float32, dev)
  if de.dim() == 2:
    de, dg = de.unsqueeze(0), dg.unsqueeze(0)
  n, H, W = de.shape
  dd, code = _depth_input(depth_test, dev)
  n_im = dd.shape[0]
  Kd = to_device(K, torch.float64, dev).reshape(-1, 9)
  if Kd.shape[0] == 1 and n_im > 1:
    Kd = Kd.expand(n_im, 9).contiguous()
  if im_idx is None:
    ii = torch.zeros(n, dtype=torch.int32, device=dev) if n_im == 1 else \
      torch.arange(n, dtype=torch.int32, device=dev)
  else:
    ii = to_device(im_idx, torch.int32, dev, (n,))
  diam = to_device(np.broadcast_to(np.asarray(diameters, dtype=np.float64), (n,)) if not
                   isinstance(diameters, torch.Tensor) else diameters, torch.float64, dev, (n,))
  sc = None
  if code == _lib.DEPTH_U16:
    sc = to_device(depth_scale if depth_scale is not None else np.ones(n_im, np.float32),
                   torch.float32, dev, (n_im,))
  wsb = _lib.lib().bop_vsd_workspace_bytes(n, H, W, 1)
  ws = _workspace(dev, wsb)
  outs, cnts = [], []
  with torch.cuda.device(dev):
    for chunk in _tau_chunks(taus):
      prm = make_vsd_params(H, W, delta, chunk, normalized_by_diameter, cost_type, visib_mode,
                            False, code)
      err = torch.empty((n, len(chunk)), dtype=torch.float64, device=dev)
      cnt = torch.empty((n, 2 + len(chunk)), dtype=torch.int64, device=dev) if return_counts else None
      rc = _lib.lib().bop_vsd_from_depth_dev(
        dd.data_ptr(), sc.data_ptr() if sc is not None else None, ii.data_ptr(), Kd.data_ptr(),
        de.data_ptr(), dg.data_ptr(), diam.data_ptr(), n, ctypes.byref(prm), err.data_ptr(),
        cnt.data_ptr() if cnt is not None else None, ws.data_ptr(), wsb, _stream_ptr())
      _lib.check(rc)
      outs.append(err)
      cnts.append(cnt)
  err = outs[0] if len(outs) == 1 else torch.cat(outs, dim=1)
  if return_counts:
    cnt = cnts[0] if len(cnts) == 1 else torch.cat([cnts[0]] + [c[:, 2:] for c in cnts[1:]], dim=1)
    return err, cnt
  return err


def render_batch(models, obj, poses, cams, height, width):
  """Renderer.render_object ('depth' mode) for B poses -> [B,H,W] float32 device tensor.
  cams: [B,4] or [4] = fx, fy, cx, cy."""
  dev = models.device
  p = to_device(poses, torch.float64, dev)
  n = p.numel() // 12
  oi = _obj_idx(models, obj, n)
  c = to_device(cams, torch.float64, dev)
  if c.numel() == 4:
    c = c.reshape(1, 4).expand(n, 4).contiguous()
  out = torch.empty((n, int(height), int(width)), dtype=torch.float32, device=dev)
  wsb = _lib.lib().bop_render_workspace_bytes(n, int(height), int(width), models.max_geom)
  ws = _workspace(dev, wsb)
  with torch.cuda.device(dev):
    rc = _lib.lib().bop_render_depth_dev(models.table.data_ptr(), oi.data_ptr(), p.data_ptr(),
                                         c.data_ptr(), n, models.max_geom, int(height), int(width),
                                         out.data_ptr(),
                                         ws.data_ptr(), wsb, _stream_ptr())
  _lib.check(rc)
  return out


def _cams(K, n, dev):
  """K: one 3x3 / [P,9] / [P,3,3] intrinsics -> [P,4] fx, fy, cx, cy float64 on `dev`."""
  Kd = to_device(K, torch.float64, dev).reshape(-1, 9)
  if Kd.shape[0] == 1 and n != 1:
    Kd = Kd.expand(n, 9)
  if Kd.shape[0] != n:
    raise ValueError('K must be [P,9] or a single 3x3')
  return Kd[:, [0, 4, 2, 5]].contiguous()


def mask_overlap_batch(models, obj, pose_est, pose_gt, K, height, width, gate=False):
  """pose_error.cus and pose_error.cou_bb_proj for P pairs (both silhouettes rasterised in
  shared memory; only statistics leave the SM).  Returns a dict of device tensors:
  'cus' [P], 'cou_bb' [P] (NaN where a silhouette is empty), 'counts' [P,4] int64 = union,
  inter, est, gt pixels, 'bbox' [P,8] int32 = (x, y, w, h) of est then gt."""
  pe, pg, n = _poses(models, pose_est, pose_gt)
  dev = models.device
  oi = _obj_idx(models, obj, n)
  cams = _cams(K, n, dev)
  out = {'cus': torch.empty(n, dtype=torch.float64, device=dev),
         'cou_bb': torch.empty(n, dtype=torch.float64, device=dev),
         'counts': torch.empty((n, 4), dtype=torch.int64, device=dev),
         'bbox': torch.empty((n, 8), dtype=torch.int32, device=dev)}
  wsb = _lib.lib().bop_vsd_workspace_bytes(n, int(height), int(width), models.max_geom)
  ws = _workspace(dev, wsb)
  with torch.cuda.device(dev):
    rc = _lib.lib().bop_mask_overlap_dev(
      models.table.data_ptr(), oi.data_ptr(), pe.data_ptr(), pg.data_ptr(), cams.data_ptr(), n,
      models.max_geom, int(height), int(width), int(bool(gate)), out['cus'].data_ptr(),
      out['cou_bb'].data_ptr(), out['counts'].data_ptr(), out['bbox'].data_ptr(), ws.data_ptr(),
      wsb, _stream_ptr())
  _lib.check(rc)
  return out


def cus_batch(models, obj, pose_est, pose_gt, K, height, width, gate=False):
  """pose_error.cus for P pairs -> [P] float64 device tensor."""
  return mask_overlap_batch(models, obj, pose_est, pose_gt, K, height, width, gate)['cus']


def cou_bb_proj_batch(models, obj, pose_est, pose_gt, K, height, width):
  """pose_error.cou_bb_proj for P pairs -> [P] float64 device tensor (NaN: empty silhouette)."""
  return mask_overlap_batch(models, obj, pose_est, pose_gt, K, height, width, False)['cou_bb']


def mask_overlap_from_depth_batch(depth_est, depth_gt, device=None):
  """Same statistics as mask_overlap_batch for pre-rendered depth images / binary masks
  ([P,H,W] or [H,W]; non-zero = inside)."""
  _lib.require_cuda()
  dev = _dev(device if device is not None else
             (depth_est.device if isinstance(depth_est, torch.Tensor) and depth_est.is_cuda else None))
  de = to_device(depth_est, torch.float32, dev)
  dg = to_device(depth_gt, torch.float32, dev)
  if de.shape != dg.shape:
    raise ValueError('masks must have the same shape')
  if de.dim() == 2:
    de, dg = de.unsqueeze(0), dg.unsqueeze(0)
  n, H, W = de.shape
  out = {'cus': torch.empty(n, dtype=torch.float64, device=dev),
         'cou_bb': torch.empty(n, dtype=torch.float64, device=dev),
         'counts': torch.empty((n, 4), dtype=torch.int64, device=dev),
         'bbox': torch.empty((n, 8), dtype=torch.int32, device=dev)}
  wsb = _lib.lib().bop_mask_stats_workspace_bytes(n)
  ws = _workspace(dev, wsb)
  with torch.cuda.device(dev):
    rc = _lib.lib().bop_mask_overlap_from_depth_dev(
      de.data_ptr(), dg.data_ptr(), n, H, W, out['cus'].data_ptr(), out['cou_bb'].data_ptr(),
      out['counts'].data_ptr(), out['bbox'].data_ptr(), ws.data_ptr(), wsb, _stream_ptr())
  _lib.check(rc)
  return out


def gt_info_batch(models, obj, pose_gt, depth_test, im_idx, K, delta=15, depth_scale=None,
                  margin=None, want_masks=False, zero_copy=False):
  """The per-instance body of scripts/calc_gt_info.py / calc_gt_masks.py for B GT instances.
  depth_test: [n_im,H,W] float32 mm or uint16 raw + depth_scale[n_im]; im_idx: [B]; K: [n_im,9].
  margin: (mx, my) canvas margin around the image; None = (W, H), the 3x canvas of calc_gt_info;
  (0, 0) = calc_gt_masks.  Returns a dict of device tensors: 'counts' [B,3] int64 =
  px_count_all / valid / visib, 'visib_fract' [B], 'bbox' [B,8] int32 = bbox_obj, bbox_visib
  (x, y, w, h), and with want_masks 'mask' / 'mask_visib' [B,H,W] uint8 (0 / 255)."""
  dev = models.device
  p = to_device(pose_gt, torch.float64, dev)
  n = p.numel() // 12
  oi = _obj_idx(models, obj, n)
  dd, code = _depth_input(depth_test, dev, zero_copy)
  n_im, H, W = dd.shape
  Kd = to_device(K, torch.float64, dev).reshape(-1, 9)
  if Kd.shape[0] == 1 and n_im > 1:
    Kd = Kd.expand(n_im, 9).contiguous()
  ii = to_device(im_idx, torch.int32, dev, (n,)) if im_idx is not None else \
    (torch.zeros(n, dtype=torch.int32, device=dev) if n_im == 1 else torch.arange(n, dtype=torch.int32, device=dev))
  sc = None
  if code == _lib.DEPTH_U16:
    sc = to_device(depth_scale if depth_scale is not None else np.ones(n_im, np.float32),
                   torch.float32, dev, (n_im,))
  mx, my = (W, H) if margin is None else (int(margin[0]), int(margin[1]))
  out = {'counts': torch.empty((n, 3), dtype=torch.int64, device=dev),
         'visib_fract': torch.empty(n, dtype=torch.float64, device=dev),
         'bbox': torch.empty((n, 8), dtype=torch.int32, device=dev)}
  if want_masks:
    out['mask'] = torch.empty((n, H, W), dtype=torch.uint8, device=dev)
    out['mask_visib'] = torch.empty((n, H, W), dtype=torch.uint8, device=dev)
  wsb = _lib.lib().bop_gt_info_workspace_bytes(n, H, W, mx, my, models.max_geom)
  ws = _workspace(dev, wsb)
  with torch.cuda.device(dev):
    rc = _lib.lib().bop_gt_info_dev(
      models.table.data_ptr(), oi.data_ptr(), p.data_ptr(), dd.data_ptr(),
      sc.data_ptr() if sc is not None else None, ii.data_ptr(), Kd.data_ptr(), n, models.max_geom,
      H, W, code, float(delta), mx, my, out['counts'].data_ptr(), out['visib_fract'].data_ptr(),
      out['bbox'].data_ptr(), out['mask'].data_ptr() if want_masks else None,
      out['mask_visib'].data_ptr() if want_masks else None, ws.data_ptr(), wsb, _stream_ptr())
  _lib.check(rc)
  return out


# ------------------------------------------------------------------------ depth PNGs on the GPU
_png_staging = {}


def png_stage(paths, threads=8, slot=0):
  """Host half of the depth-PNG path: files -> zlib streams in a grow-only pinned buffer (one per
  `slot`, so a caller can stage the next chunk on a thread while the GPU decodes this one; the
  ctypes call releases the GIL).  Returns an opaque tuple for `decode_depth_staged`."""
  L = _lib.lib()
  n = len(paths)
  arr = (ctypes.c_char_p * n)(*[os.fsencode(p) for p in paths])
  offs = np.zeros(n, dtype=np.int64)
  nbytes = np.zeros(n, dtype=np.int32)
  infos = (_lib.PngInfo * n)()
  need = ctypes.c_size_t()
  for _ in range(2):
    buf = _png_staging.get(slot)
    cap = buf.numel() if buf is not None else 0
    rc = L.bop_png_stage_files_host(arr, n, buf.data_ptr() if buf is not None else None, cap,
                                    offs.ctypes.data, nbytes.ctypes.data, infos, int(threads),
                                    ctypes.byref(need))
    if rc != _lib.PNG_STAGING_TOO_SMALL:
      _lib.check(rc)
      return list(paths), buf, int(need.value), offs, nbytes, infos
    _png_staging.pop(slot, None)
    _png_staging[slot] = torch.empty(max(int(need.value * 1.25), 1 << 24), dtype=torch.uint8,
                                      pin_memory=torch.cuda.is_available())
  raise _lib.BopCudaError('bop_png_stage_files_host: staging buffer still too small')


def decode_depth_pngs(paths, height, width, device=None, threads=8):
  """`inout.load_depth` for a whole chunk of 16-bit depth PNGs, decoded ON THE GPU
  (inout.py:39-46 as called by eval_calc_errors.py:236-243): host threads only read the files and
  strip the PNG container (bop_png_stage_files_host), the zlib streams cross PCIe compressed and
  one warp per image inflates + unfilters them (bop_png_decode_dev).  Returns an int16 CUDA tensor
  [n,H,W] holding the raw uint16 samples (what the VSD entry points take with depth_scale).
  Files that are not 16-bit greyscale non-interlaced PNGs (none in the BOP datasets) are decoded
  on the host by OpenCV and uploaded; they must still be 16-bit."""
  _lib.require_cuda()
  if len(paths) == 0:
    return torch.empty((0, height, width), dtype=torch.int16, device=_dev(device))
  return decode_depth_staged(png_stage(paths, threads), height, width, device)


def decode_depth_staged(staged, height, width, device=None):
  """Device half of `decode_depth_pngs` on the result of `png_stage`."""
  _lib.require_cuda()
  dev = _dev(device)
  paths, buf, used, offs, nbytes, infos = staged
  n = len(paths)
  out = torch.empty((n, height, width), dtype=torch.int16, device=dev)
  foreign = []
  for i in range(n):
    st = infos[i].status
    if st == _lib.PNG_UNSUPPORTED:
      foreign.append(i)
    elif st != 0:
      raise IOError('cannot read depth image {} (status {})'.format(paths[i], st))
    if (infos[i].height, infos[i].width) != (height, width):
      raise ValueError('depth image {} is {}x{}, expected {}x{}'.format(
        paths[i], infos[i].width, infos[i].height, width, height))
  L = _lib.lib()
  with torch.cuda.device(dev):
    comp = buf[:max(used, 16)].to(dev, non_blocking=True)
    offs_d = torch.from_numpy(offs).to(dev, non_blocking=True)
    nb_d = torch.from_numpy(nbytes).to(dev, non_blocking=True)
    status = torch.zeros(n, dtype=torch.int32, device=dev)
    wsb = L.bop_png_decode_workspace_bytes(n, height, width)
    ws = _workspace(dev, wsb)
    _lib.check(L.bop_png_decode_dev(comp.data_ptr(), offs_d.data_ptr(), nb_d.data_ptr(), n, height,
                                    width, out.data_ptr(), status.data_ptr(), ws.data_ptr(), wsb,
                                    _stream_ptr()))
    st = status.cpu().numpy()       # synchronises: the staging buffer may be refilled afterwards
  for i in foreign:
    from . import inout
    im = inout.load_depth_raw(paths[i])
    if im.dtype != np.uint16:
      raise ValueError('depth image {} is not 16-bit'.format(paths[i]))
    out[i] = torch.from_numpy(im.view(np.int16)).to(dev)
    st[i] = 0
  bad = np.nonzero(st)[0]
  if bad.size:
    raise IOError('corrupt depth PNG {} (decoder status {})'.format(paths[bad[0]], int(st[bad[0]])))
  return out


def depth_bytes_fetched(device, depth_dtype_code):
  """Bytes of depth_test the last vsd_batch pass on `device` fetched (synchronises)."""
  ws = _workspaces[(device.type, device.index)].buf
  out = ctypes.c_uint64(0)
  _lib.check(_lib.lib().bop_vsd_depth_bytes_fetched(ws.data_ptr(), depth_dtype_code, ctypes.byref(out)))
  return int(out.value)


def kernel_launch_count():
  return int(_lib.lib().bop_kernel_launch_count())


def debug_triangle_list(models, n_pairs, height, width, pair, pose):
  """Test hook: the geometry pass's triangle list of (pair, pose) from the last vsd_batch call with
  these sizes on models.device -> int32 [n, 6] snapped window positions (1/256 px)."""
  dev = models.device
  ws = _workspaces[(dev.type, dev.index)].buf
  cap = 2 * ((models.max_geom + 7) & ~7)
  xy = np.zeros((cap, 6), dtype=np.int32)
  cnt = ctypes.c_int(0)
  _lib.check(_lib.lib().bop_vsd_debug_triangle_list(
    ws.data_ptr(), int(n_pairs), int(height), int(width), models.max_geom, int(pair), int(pose),
    xy.ctypes.data, cap, ctypes.byref(cnt)))
  return xy[:min(cap, cnt.value)]

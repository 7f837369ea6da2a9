"""Calculates errors of 6D object pose estimates -- batched, B200-native drop-in for
scripts/eval_calc_errors.py of the reference.

Same command-line flags and defaults (eval_calc_errors.py:21-129), same input files (BOP19
results CSV, targets JSON, scene_camera / scene_gt, depth PNGs, eval models) and the same output
files (`<eval_path>/<result_name>/<error_sign>/errors_<scene_id:06d>.json`, one file per tau for
VSD, eval_calc_errors.py:391-418).  What changes is the execution order: instead of one
pose_error.* call per (estimate, GT) pair inside four nested Python loops (:216-389), every pair
of a scene is flattened into arrays and evaluated by one batched CUDA call per error type; the
driver's sphere gates (:296-339) run on the device.

Under torch.distributed (one process per GPU, launched by torchrun) the images of each scene
are split over the ranks, every rank evaluates its own pairs and the error values are
all-gathered; rank 0 writes the files.
"""
import argparse
import os
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import dataset_params, inout, misc

VSD_DELTAS = {'hb': 15, 'icbin': 15, 'icmi': 15, 'itodd': 5, 'lm': 15, 'lmo': 15, 'ruapc': 15,
              'tless': 15, 'tudl': 15, 'tyol': 15, 'ycbv': 15, 'hope': 15}

ERROR_TYPES = ('vsd', 'mssd', 'mspd', 'ad', 'add', 'adi', 'cus', 'proj', 'rete', 're', 'te')


def default_params():
  """The `p` dict of eval_calc_errors.py:21-84 (paths default to the BOP_PATH convention of
  bop_toolkit_lib/config.py)."""
  bop = os.environ.get('BOP_PATH', '/path/to/bop/datasets')
  return {
    'n_top': 1,
    'error_type': 'vsd',
    'vsd_deltas': dict(VSD_DELTAS),
    'vsd_taus': list(np.arange(0.05, 0.51, 0.05)),
    'vsd_normalized_by_diameter': True,
    'max_sym_disc_step': 0.01,
    'skip_missing': True,
    'renderer_type': 'cuda',
    'result_filenames': ['/path/to/csv/with/results'],
    'results_path': os.environ.get('BOP_RESULTS_PATH', '/path/to/folder/with/results'),
    'eval_path': os.environ.get('BOP_EVAL_PATH', '/path/to/eval/folder'),
    'datasets_path': bop,
    'targets_filename': 'test_targets_bop19.json',
    'out_errors_tpath': os.path.join('{eval_path}', '{result_name}', '{error_sign}',
                                     'errors_{scene_id:06d}.json'),
  }


def parse_args(argv=None):
  """Same flags as eval_calc_errors.py:93-129, including the `bool(str)` quirk of
  --vsd_normalized_by_diameter / --skip_missing (any non-empty string is True)."""
  p = default_params()
  deltas = ','.join('{}:{}'.format(k, v) for k, v in p['vsd_deltas'].items())
  ap = argparse.ArgumentParser()
  ap.add_argument('--n_top', default=p['n_top'])
  ap.add_argument('--error_type', default=p['error_type'])
  ap.add_argument('--vsd_deltas', default=deltas)
  ap.add_argument('--vsd_taus', default=','.join(map(str, p['vsd_taus'])))
  ap.add_argument('--vsd_normalized_by_diameter', default=p['vsd_normalized_by_diameter'])
  ap.add_argument('--max_sym_disc_step', default=p['max_sym_disc_step'])
  ap.add_argument('--skip_missing', default=p['skip_missing'])
  ap.add_argument('--renderer_type', default=p['renderer_type'])
  ap.add_argument('--result_filenames', default=','.join(p['result_filenames']),
                  help='Comma-separated names of files with results.')
  ap.add_argument('--results_path', default=p['results_path'])
  ap.add_argument('--eval_path', default=p['eval_path'])
  ap.add_argument('--datasets_path', default=p['datasets_path'])
  ap.add_argument('--targets_filename', default=p['targets_filename'])
  ap.add_argument('--out_errors_tpath', default=p['out_errors_tpath'])
  a = ap.parse_args(argv)
  p['n_top'] = int(a.n_top)
  p['error_type'] = str(a.error_type)
  p['vsd_deltas'] = {str(e.split(':')[0]): float(e.split(':')[1]) for e in a.vsd_deltas.split(',')}
  p['vsd_taus'] = list(map(float, a.vsd_taus.split(',')))
  p['vsd_normalized_by_diameter'] = bool(a.vsd_normalized_by_diameter)
  p['max_sym_disc_step'] = float(a.max_sym_disc_step)
  p['skip_missing'] = bool(a.skip_missing)
  p['renderer_type'] = str(a.renderer_type)
  p['result_filenames'] = a.result_filenames.split(',')
  for k in ('results_path', 'eval_path', 'datasets_path', 'targets_filename', 'out_errors_tpath'):
    p[k] = str(getattr(a, k))
  return p


def parse_result_filename(result_filename):
  """`<method>_<dataset>-<split>[-<split_type>].csv` (eval_calc_errors.py:144-150)."""
  result_name = os.path.splitext(os.path.basename(result_filename))[0]
  info = result_name.split('_')
  method = str(info[0])
  ds = info[1].split('-')
  return result_name, method, str(ds[0]), str(ds[1]), (str(ds[2]) if len(ds) > 2 else None)


# --------------------------------------------------------------------------- flattening
class ScenePairs(object):
  """All (estimate, GT) pairs of one scene as flat arrays + the record skeletons they fill.

  records[i] = {'im_id', 'obj_id', 'est_id', 'score', 'errors': {gt_id: None}} in exactly the
  order the reference appends them (eval_calc_errors.py:224-389): targets order of images and
  objects, estimates by descending score (stable), GT instances by ascending gt_id.
  """

  def __init__(self):
    self.records = []
    self.rec = []      # record index per pair
    self.gt_id = []
    self.obj_id = []
    self.im_pos = []   # index into self.im_ids
    self.pose_est = []
    self.pose_gt = []
    self.im_ids = []   # images that own at least one pair, in first-use order
    self.n_ests = 0


def select_top_estimates(obj_ests, n_top_curr):
  """(est_id, est) of the n best-scored estimates; est_id = position in CSV order; stable sort
  by descending score (eval_calc_errors.py:271-276)."""
  ranked = sorted(enumerate(obj_ests), key=lambda x: x[1]['score'], reverse=True)
  return ranked[slice(0, n_top_curr)]


def flatten_scene(p, scene_id, scene_targets, scene_ests, scene_gt):
  sp = ScenePairs()
  im_pos_of = {}
  est_R, est_t, gt_R, gt_t = [], [], [], []
  for im_id, im_targets in scene_targets.items():
    gts = scene_gt[im_id]
    for obj_id, target in im_targets.items():
      if p['n_top'] == 0:
        n_top_curr = None
      elif p['n_top'] == -1:
        n_top_curr = target['inst_count']
      else:
        n_top_curr = p['n_top']
      obj_ests = scene_ests.get(im_id, {}).get(obj_id, [])
      if not p['skip_missing'] and len(obj_ests) < n_top_curr:
        raise ValueError(
          'Not enough estimates for scene: {}, im: {}, obj: {} (provided: {}, expected: {})'.format(
            scene_id, im_id, obj_id, len(obj_ests), n_top_curr))
      top = select_top_estimates(obj_ests, n_top_curr)
      sp.n_ests += len(top)
      gt_ids = [g for g, gt in enumerate(gts) if gt['obj_id'] == obj_id]
      for est_id, est in top:
        ri = len(sp.records)
        sp.records.append({'im_id': im_id, 'obj_id': obj_id, 'est_id': est_id,
                           'score': est['score'], 'errors': {g: None for g in gt_ids}})
        if not gt_ids:
          continue
        if im_id not in im_pos_of:
          im_pos_of[im_id] = len(sp.im_ids)
          sp.im_ids.append(im_id)
        pos = im_pos_of[im_id]
        for g in gt_ids:
          gt = gts[g]
          sp.rec.append(ri)
          sp.gt_id.append(g)
          sp.obj_id.append(obj_id)
          sp.im_pos.append(pos)
          est_R.append(est['R'])      # stacked once per scene below: one numpy call per pair
          est_t.append(est['t'])      # costs more than the kernels spend on it
          gt_R.append(gt['cam_R_m2c'])
          gt_t.append(gt['cam_t_m2c'])
  n = len(sp.rec)
  sp.rec = np.asarray(sp.rec, dtype=np.int64)
  sp.gt_id = np.asarray(sp.gt_id, dtype=np.int64)
  sp.obj_id = np.asarray(sp.obj_id, dtype=np.int64)
  sp.im_pos = np.asarray(sp.im_pos, dtype=np.int32)

  def stack(Rs, ts):
    out = np.empty((n, 12), dtype=np.float64)
    if n:
      try:
        out[:, :9] = np.asarray(Rs, dtype=np.float64).reshape(n, 9)
        out[:, 9:] = np.asarray(ts, dtype=np.float64).reshape(n, 3)
      except ValueError:       # ragged input: convert entry by entry (raises where the reference would)
        for k in range(n):
          out[k, :9] = np.asarray(Rs[k], np.float64).reshape(9)
          out[k, 9:] = np.asarray(ts[k], np.float64).reshape(3)
    return out
  sp.pose_est = stack(est_R, est_t)
  sp.pose_gt = stack(gt_R, gt_t)
  return sp


# ------------------------------------------------------------------------------ host I/O
_MEMO = {}
_FLAT = {}


def flatten_scene_cached(p, scene_id, scene_targets, scene_ests, scene_gt, files):
  """flatten_scene, remembered per (input files, scene, n_top, skip_missing): the pairs of a scene
  do not depend on the error type, and a BOP19 evaluation flattens every scene three times.  The
  flat arrays are shared (read-only); the record skeletons are fresh on every call because
  fill_records writes the errors into them and the caller keeps them."""
  key = tuple((f, os.stat(f).st_mtime_ns, os.stat(f).st_size) for f in files) + \
    (scene_id, p['n_top'], bool(p['skip_missing']))
  base = _FLAT.get(key)
  if base is None:
    if len(_FLAT) >= 256:
      _FLAT.clear()
    base = _FLAT[key] = flatten_scene(p, scene_id, scene_targets, scene_ests, scene_gt)
  sp = ScenePairs()
  sp.__dict__.update(base.__dict__)
  sp.records = [{'im_id': r['im_id'], 'obj_id': r['obj_id'], 'est_id': r['est_id'], 'score': r['score'],
                 'errors': dict.fromkeys(r['errors'])} for r in base.records]
  return sp


def _memo(fn, path, *args):
  """Parsed files keyed by (path, mtime, size): eval_bop19_pose runs three error types over the
  same result CSV and the same PLY models in one process.  Callers treat the result as read-only."""
  st = os.stat(path)
  key = (fn.__name__, path, st.st_mtime_ns, st.st_size, args)
  if key not in _MEMO:
    if len(_MEMO) >= 128:
      _MEMO.clear()
    _MEMO[key] = fn(path, *args)
  return _MEMO[key]


def _io_threads():
  try:
    return max(1, min(32, len(os.sched_getaffinity(0))))
  except AttributeError:
    return max(1, min(32, os.cpu_count() or 8))


_PINNED = {}


def _pinned_stack(dtype, n, height, width):
  """[n,H,W] view of a grow-only pinned host buffer (cudaHostAlloc runs at a few GB/s, which is
  seconds per dataset when every chunk allocates its own).  The caller finishes with a chunk --
  the driver reads the errors back, which synchronises -- before asking for the next one."""
  import torch
  need = n * height * width
  if not torch.cuda.is_available():
    return torch.empty((n, height, width), dtype=dtype)
  buf = _PINNED.get(dtype)
  if buf is None or buf.numel() < need:
    _PINNED.pop(dtype, None)
    buf = _PINNED[dtype] = torch.empty(max(need, 64 * height * width), dtype=dtype, pin_memory=True)
  return buf[:need].view(n, height, width)


# One image keeps the GPU decoder busy for tens of milliseconds (a DEFLATE stream is serial); below
# this many images per call the host threads are quicker.  The VSD driver is not affected: it
# groups images across scenes (GpuDepthSource).
GPU_DECODE_MIN_IMAGES = 192


def gpu_depth_decode(paths):
  """True when the depth images of a chunk go through the GPU PNG decoder (all '.png', and
  BOP_B200_DEPTH_DECODE is not 'host')."""
  return os.environ.get('BOP_B200_DEPTH_DECODE', 'gpu') != 'host' and len(paths) > 0 and \
    all(q.lower().endswith('.png') for q in paths)


def load_depth_stack(paths, scales, height, width, threads=None, device=None):
  """The depth images of a chunk as ONE tensor [n,H,W] + scales (float32 [n] or None).
  With `device` and PNG files: inflated and unfiltered on the GPU (batch.decode_depth_pngs), a
  CUDA int16 tensor of raw samples.  Otherwise decoded by OpenCV on host threads into ONE pinned
  host tensor the fused kernel reads in place.  uint16 PNG samples stay uint16 (the kernel
  multiplies by float32(depth_scale) exactly as `depth_im *= depth_scale` does,
  eval_calc_errors.py:239-243); any other sample type is converted to float32 millimetres."""
  import torch
  n = len(paths)
  threads = threads or _io_threads()      # PNG inflate releases the GIL: one decoder per host core
  if device is not None and n >= GPU_DECODE_MIN_IMAGES and gpu_depth_decode(paths):
    from . import batch
    return batch.decode_depth_pngs(paths, height, width, device, threads), \
      np.asarray(scales, dtype=np.float32)
  with ThreadPoolExecutor(max_workers=max(1, min(threads, n))) as ex:
    ims = list(ex.map(inout.load_depth_raw, paths))
  for im, path in zip(ims, paths):
    if im.shape != (height, width):
      raise ValueError('depth image {} is {}x{}, expected {}x{}'.format(
        path, im.shape[1], im.shape[0], width, height))
  if all(im.dtype == np.uint16 for im in ims):
    out = _pinned_stack(torch.int16, n, height, width)
    view = out.numpy().view(np.uint16)
    for i, im in enumerate(ims):
      view[i] = im
    return out, np.asarray(scales, dtype=np.float32)
  out = _pinned_stack(torch.float32, n, height, width)
  view = out.numpy()
  for i, im in enumerate(ims):
    view[i] = im.astype(np.float32)
    view[i] *= np.float32(scales[i])
  return out, None


class GpuDepthSource(object):
  """The depth PNGs of a whole run, decoded on the GPU in groups that cut ACROSS scenes.

  One DEFLATE stream is serial, so a single image takes the decoder tens of milliseconds however
  few images are in flight; its throughput comes from thousands of streams decoding side by side
  (one warp each).  The driver therefore registers the images of every scene up front and this
  class decodes them `group` at a time (a scene of 100 images shares its launch with its
  neighbours, a scene of 5,000 spans two), staging group g+1 on a host thread while group g is
  in use.  Access is expected in registration order; groups behind the current one are freed."""

  def __init__(self, batch, device, height, width, group=None, budget_bytes=None):
    # A group holds, per image, H*(2W+1) bytes of filtered scanlines + 2*H*W bytes of samples on
    # the device (and up to two groups are live while a request spans a boundary): size it by
    # bytes, not by count -- 4,096 images of 640x480 are 5 GB, of 1920x1080 they would be 34 GB.
    if group is None:
      budget = int(budget_bytes if budget_bytes is not None else
                   float(os.environ.get('BOP_B200_DEPTH_GROUP_GB', '5')) * (1 << 30))
      group = int(min(4096, max(64, budget // (height * (4 * width + 1)))))
    self.batch, self.device, self.h, self.w, self.group = batch, device, height, width, group
    self.paths, self.index = [], {}
    self.live, self.futs = {}, {}
    self.pool = None

  def register(self, scene_id, im_ids, paths):
    for im_id, path in zip(im_ids, paths):
      if (scene_id, im_id) not in self.index:
        self.index[(scene_id, im_id)] = len(self.paths)
        self.paths.append(path)

  def _submit(self, g):
    lo = g * self.group
    if lo >= len(self.paths) or g in self.futs or g in self.live:
      return
    if self.pool is None:
      self.pool = ThreadPoolExecutor(max_workers=1)
    # the group's extent is fixed at submission: images registered later start a new read
    self.futs[g] = self.pool.submit(self.batch.png_stage, self.paths[lo:lo + self.group],
                                    _io_threads(), g & 1)

  def _ensure(self, g, n_needed):
    t = self.live.get(g)
    if t is not None and t.shape[0] >= n_needed:
      return t
    self.live.pop(g, None)
    if g in self.futs and len(self.futs[g].result()[0]) < n_needed:
      self.futs.pop(g)                       # staged before later images were registered
    self._submit(g)
    staged = self.futs.pop(g).result()
    self._submit(g + 1)
    t = self.live[g] = self.batch.decode_depth_staged(staged, self.h, self.w, self.device)
    return t

  def get(self, scene_id, im_ids):
    """-> CUDA int16 tensor [len(im_ids),H,W] of raw samples (a view when the images are
    consecutive inside one group)."""
    import torch
    pos = np.asarray([self.index[(scene_id, i)] for i in im_ids], dtype=np.int64)
    if pos.size == 0:
      return torch.empty((0, self.h, self.w), dtype=torch.int16, device=self.device)
    grp = pos // self.group
    for g in [k for k in self.live if k < grp.min()]:
      del self.live[g]
    parts = []
    for g in np.unique(grp):
      local = pos[grp == g] - g * self.group
      t = self._ensure(int(g), int(local.max()) + 1)
      if np.array_equal(local, np.arange(local[0], local[0] + local.size)):
        parts.append(t[int(local[0]):int(local[0]) + local.size])
      else:
        parts.append(t.index_select(0, torch.from_numpy(local).to(t.device)))
    if len(parts) > 1 and not np.all(np.diff(grp) >= 0):
      raise ValueError('depth images must be requested in registration order')
    return parts[0] if len(parts) == 1 else torch.cat(parts, dim=0)

  def close(self):
    self.live.clear()
    for f in self.futs.values():
      f.cancel()
    self.futs.clear()
    if self.pool is not None:
      self.pool.shutdown(wait=True)
      self.pool = None


# ---------------------------------------------------------------------------- compute
class ErrorEngine(object):
  """Models resident on one GPU + the batched error calls for one result file."""

  def __init__(self, p, dataset, dp_model, dp_split, needed_obj_ids, device=None):
    import torch
    from . import batch
    self.p, self.dataset, self.dp_model, self.dp_split = p, dataset, dp_model, dp_split
    self.torch, self.batch = torch, batch
    et = p['error_type']
    if et not in ERROR_TYPES:
      raise ValueError('Unknown pose error function.')
    self.models_info = None
    if et in ('ad', 'add', 'adi', 'vsd', 'mssd', 'mspd', 'cus'):
      self.models_info = inout.load_json(dp_model['models_info_path'], keys_to_int=True)
    self.ms = None
    if et not in ('rete', 're', 'te'):
      self.ms = batch.ModelSet(device)
      need_faces = et in ('vsd', 'cus')
      for obj_id in dp_model['obj_ids']:
        if obj_id not in needed_obj_ids:
          continue
        model = _memo(inout.load_ply, dp_model['model_tpath'].format(obj_id=obj_id))
        syms = None
        if et in ('mssd', 'mspd'):
          syms = misc.get_symmetry_transformations(self.models_info[obj_id], p['max_sym_disc_step'])
        diameter = self.models_info[obj_id]['diameter'] if self.models_info is not None else 0.0
        self.ms.add(obj_id, model['pts'], model.get('faces') if need_faces else None, syms, diameter)
    self.device = self.ms.device if self.ms is not None else torch.device(device or 'cuda:0')
    self.depth_source = None

  def _depth_paths(self, scene_id, im_ids):
    return [self.dp_split['depth_tpath'].format(scene_id=scene_id, im_id=i) for i in im_ids]

  def register_scene(self, scene_id, sp, sel):
    """VSD only: announces the depth images the pairs `sel` of a scene will read, so that the
    GPU PNG decoder can work on groups of images that cut across scenes (GpuDepthSource)."""
    if self.p['error_type'] != 'vsd' or len(sel) == 0:
      return
    im_ids = [sp.im_ids[i] for i in np.unique(sp.im_pos[sel])]
    paths = self._depth_paths(scene_id, im_ids)
    if not gpu_depth_decode(paths):
      return
    if self.depth_source is None:
      w, h = self.dp_split['im_size']
      self.depth_source = GpuDepthSource(self.batch, self.device, h, w)
    self.depth_source.register(scene_id, im_ids, paths)

  def close(self):
    if self.depth_source is not None:
      self.depth_source.close()
      self.depth_source = None

  def _K_rows(self, scene_camera, im_ids):
    return np.stack([np.asarray(scene_camera[i]['cam_K'], np.float64).reshape(9) for i in im_ids]) \
      if len(im_ids) else np.zeros((0, 9))

  def compute(self, scene_id, sp, scene_camera, sel):
    """Errors of the pairs `sel` (index array into sp's pair arrays) -> float64 [len(sel), T]."""
    p, et, batch, torch = self.p, self.p['error_type'], self.batch, self.torch
    n = len(sel)
    T = len(p['vsd_taus']) if et == 'vsd' else (2 if et == 'rete' else 1)
    if n == 0:
      return np.zeros((0, T))
    pe, pg = sp.pose_est[sel], sp.pose_gt[sel]
    if et in ('rete', 're', 'te'):
      re, te = batch.re_te_batch(pe, pg, self.device)
      cols = {'rete': [re, te], 're': [re], 'te': [te]}[et]
      return torch.stack(cols, dim=1).cpu().numpy()
    oi = self.ms.indices(sp.obj_id[sel])
    if et == 'vsd':
      return self._vsd(scene_id, sp, scene_camera, sel, oi, pe, pg)
    if et in ('mspd', 'proj'):
      K = self._K_rows(scene_camera, sp.im_ids)[sp.im_pos[sel]]
      fn = batch.mspd_batch if et == 'mspd' else batch.proj_batch
      return fn(self.ms, oi, pe, pg, K).cpu().numpy().reshape(n, 1)
    if et == 'mssd':
      return batch.mssd_batch(self.ms, oi, pe, pg, gate=True).cpu().numpy().reshape(n, 1)
    if et == 'add':
      return batch.add_batch(self.ms, oi, pe, pg, gate=True).cpu().numpy().reshape(n, 1)
    if et == 'adi':
      return batch.adi_batch(self.ms, oi, pe, pg, gate=True).cpu().numpy().reshape(n, 1)
    if et == 'ad':   # ADI for objects with ambiguous views, ADD otherwise (:341-347)
      sym_ids = self.dp_model['symmetric_obj_ids']
      is_sym = np.isin(sp.obj_id[sel], np.asarray(sym_ids, dtype=np.int64))
      out = np.empty((n, 1))
      for mask, fn in ((is_sym, batch.adi_batch), (~is_sym, batch.add_batch)):
        if mask.any():
          out[mask, 0] = fn(self.ms, oi[mask], pe[mask], pg[mask], gate=True).cpu().numpy()
      return out
    if et == 'cus':
      K = self._K_rows(scene_camera, sp.im_ids)[sp.im_pos[sel]]
      w, h = self.dp_split['im_size']
      return batch.cus_batch(self.ms, oi, pe, pg, K, h, w, gate=True).cpu().numpy().reshape(n, 1)
    raise ValueError('Unknown pose error function.')

  def _vsd(self, scene_id, sp, scene_camera, sel, oi, pe, pg, images_per_chunk=2048):
    p, batch = self.p, self.batch
    w, h = self.dp_split['im_size']
    T = len(p['vsd_taus'])
    out = np.empty((len(sel), T))
    im_pos = sp.im_pos[sel]
    used = np.unique(im_pos)                      # images this rank touches, ascending position
    # 16-bit PNGs (every BOP dataset) are inflated and unfiltered on the GPU; the host threads only
    # read the files.  BOP_B200_DEPTH_DECODE=host keeps the OpenCV decode + zero-copy pinned path.
    self.register_scene(scene_id, sp, sel)        # no-op if calc_errors announced the scene already
    for c0 in range(0, len(used), images_per_chunk):
      chunk = used[c0:c0 + images_per_chunk]
      im_ids = [sp.im_ids[i] for i in chunk]
      paths = self._depth_paths(scene_id, im_ids)
      depth = None
      if self.depth_source is not None and gpu_depth_decode(paths):
        try:
          depth = self.depth_source.get(scene_id, im_ids)
          sc = np.asarray([scene_camera[i]['depth_scale'] for i in im_ids], dtype=np.float32)
        except self.torch.cuda.OutOfMemoryError:
          # no room for a decode group next to the geometry scratch: the rest of the run reads
          # its depth through the host decoder (same samples, zero-copy pinned path)
          self.depth_source.close()
          self.depth_source = None
          os.environ['BOP_B200_DEPTH_DECODE'] = 'host'
          self.torch.cuda.empty_cache()
      if depth is None:
        depth, sc = load_depth_stack(paths, [scene_camera[i]['depth_scale'] for i in im_ids], h, w)
      rows = np.nonzero(np.isin(im_pos, chunk))[0]
      local = np.searchsorted(chunk, im_pos[rows]).astype(np.int32)
      # every estimate of an object meets every GT pose of it (:281-289): the pairs that share a GT
      # pose go to the kernel as one group (GT geometry and raster work once per group)
      order, goff = batch.group_pairs_by_gt(oi[rows], pg[rows], local)
      rows, local = rows[order], local[order]
      err = batch.vsd_batch(self.ms, oi[rows], pe[rows], pg[rows], depth, local,
                            self._K_rows(scene_camera, im_ids), p['vsd_deltas'][self.dataset],
                            p['vsd_taus'], p['vsd_normalized_by_diameter'], 'step', 'bop19',
                            gate=True, depth_scale=sc,
                            zero_copy=(not depth.is_cuda) and depth.is_pinned(),
                            group_off=(goff if len(goff) - 1 < len(rows) else None))
      out[rows] = err.cpu().numpy()
    return out


# ----------------------------------------------------------------------------- driver
def _dist_info():
  try:
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
      return dist.get_rank(), dist.get_world_size()
  except ImportError:
    pass
  return 0, 1


def _shard_pairs(sp, rank, world):
  """Pairs owned by `rank`: images are the unit (depth_test stays local); greedy LPT on the
  number of pairs per image (sharding.assign_images)."""
  n = len(sp.rec)
  if world == 1:
    return np.arange(n)
  from . import sharding
  cost = np.bincount(sp.im_pos, minlength=len(sp.im_ids))
  owner = sharding.assign_images(cost, world)
  return np.nonzero(owner[sp.im_pos] == rank)[0]


def _gather_errors(errs, sel, n_pairs, device, sp=None, world=None):
  """All-gathers (pair index, error row) from every rank -> float64 [n_pairs, T] on every rank.
  With `sp` the per-rank row counts follow from the (deterministic) image assignment, so the
  gather is a single collective."""
  import torch
  from . import sharding
  counts = None
  if sp is not None and world:
    cost = np.bincount(sp.im_pos, minlength=len(sp.im_ids))
    owner = sharding.assign_images(cost, world)
    counts = np.bincount(owner[sp.im_pos], minlength=world).tolist()
  ids = torch.zeros((len(sel), 5), dtype=torch.int32, device=device)
  ids[:, 0] = torch.from_numpy(sel.astype(np.int32)).to(device)
  e = torch.from_numpy(np.ascontiguousarray(errs)).to(device)
  ids_all, e_all = sharding.gather_records(ids, e, counts=counts)
  out = np.full((n_pairs, errs.shape[1]), np.nan)
  out[ids_all[:, 0].cpu().numpy().astype(np.int64)] = e_all.cpu().numpy()
  return out


def fill_records(sp, errs):
  """errs [n_pairs, T] -> records[...]['errors'][gt_id] = [e_0 .. e_T-1] (Python floats)."""
  rows = errs.tolist()
  for k in range(len(rows)):
    sp.records[sp.rec[k]]['errors'][int(sp.gt_id[k])] = rows[k]
  return sp.records


def records_for_tau(records, tau_id):
  """Per-tau view of VSD records (eval_calc_errors.py:408-414), without the deep copy."""
  return [{'im_id': r['im_id'], 'obj_id': r['obj_id'], 'est_id': r['est_id'], 'score': r['score'],
           'errors': {g: [e[tau_id]] for g, e in r['errors'].items()}} for r in records]


def _record_arrays(sp, errs):
  """Flat arrays for bop_write_error_file_host, or None when the records hold anything the native
  writer does not format exactly like json.dumps (non-int ids, non-float scores, unsorted pairs)."""
  recs = sp.records
  cols = [[r[k] for r in recs] for k in ('im_id', 'obj_id', 'est_id', 'score')]
  if not all(type(v) is int for c in cols[:3] for v in c) or not all(type(v) is float for v in cols[3]):
    return None
  errs = np.ascontiguousarray(errs, dtype=np.float64)
  rec, gt = np.ascontiguousarray(sp.rec, np.int64), np.ascontiguousarray(sp.gt_id, np.int64)
  if errs.ndim != 2 or errs.shape[0] != rec.size:
    return None
  if rec.size > 1:
    d = np.diff(rec)
    if (d < 0).any() or ((d == 0) & (np.diff(gt) <= 0)).any():
      return None
  try:
    ids = [np.array(c, dtype=np.int64) for c in cols[:3]]
  except OverflowError:
    return None
  return ids + [np.array(cols[3], dtype=np.float64), rec, gt, errs]


def save_scene_errors(p, dataset, result_name, scene_id, records, log=misc.log, sp=None, errs=None):
  """The error files of one scene in the reference's layout (eval_calc_errors.py:380-418,
  inout.save_json): VSD -> one file per tau holding 1-element error lists.  With the scene's
  flat arrays (`sp`, `errs`) the text is produced by the native writer (same bytes, no GIL);
  otherwise by inout.dumps_error_records."""
  arrays = _record_arrays(sp, errs) if (sp is not None and errs is not None) else None
  tails = inout.error_record_tails(records) if arrays is None else None

  def save(sign, tau_id):
    path = p['out_errors_tpath'].format(eval_path=p['eval_path'], result_name=result_name,
                                        error_sign=sign, scene_id=scene_id)
    misc.ensure_dir(os.path.dirname(path))
    log('Saving errors to: {}'.format(path))
    if arrays is not None:
      from . import _lib
      im, ob, es, sc, rec, gt, e = arrays
      _lib.check(_lib.lib().bop_write_error_file_host(
        os.fsencode(path), len(im), im.ctypes.data, ob.ctypes.data, es.ctypes.data, sc.ctypes.data,
        rec.size, rec.ctypes.data, gt.ctypes.data, e.ctypes.data, e.shape[1],
        -1 if tau_id is None else tau_id))
      return
    with open(path, 'w') as f:
      f.write(inout.dumps_error_records(records, tau_id, tails))
  if p['error_type'] == 'vsd':
    for tau_id, tau in enumerate(p['vsd_taus']):
      sign = misc.get_error_signature(p['error_type'], p['n_top'],
                                      vsd_delta=p['vsd_deltas'][dataset], vsd_tau=tau)
      save(sign, tau_id)
  else:
    save(misc.get_error_signature(p['error_type'], p['n_top']), None)


def calc_errors(p, device=None, log=misc.log, write=True):
  """Runs the error calculation for every result file in p['result_filenames'].
  Returns {result_name: {scene_id: records}} (on rank 0; other ranks return the same data)."""
  rank, world = _dist_info()
  if device is None:
    import torch
    if torch.cuda.is_available() and torch.distributed.is_available() and torch.distributed.is_initialized():
      device = 'cuda:%d' % torch.cuda.current_device()    # set by sharding.init_from_env (one process per GPU)
    else:
      device = 'cuda:%d' % int(os.environ.get('LOCAL_RANK', '0'))
  all_out = {}
  for result_filename in p['result_filenames']:
    log('Processing: {}'.format(result_filename))
    t_start = time.time()
    result_name, method, dataset, split, split_type = parse_result_filename(result_filename)
    dp_split = dataset_params.get_split_params(p['datasets_path'], dataset, split, split_type)
    dp_model = dataset_params.get_model_params(p['datasets_path'], dataset, 'eval')
    targets = inout.load_json(os.path.join(dp_split['base_path'], p['targets_filename']))
    targets_org = {}
    for t in targets:
      targets_org.setdefault(t['scene_id'], {}).setdefault(t['im_id'], {})[t['obj_id']] = t
    log('Loading pose estimates...')
    ests = _memo(inout.load_bop_results, os.path.join(p['results_path'], result_filename))
    ests_org = {}
    for est in ests:
      ests_org.setdefault(est['scene_id'], {}).setdefault(est['im_id'], {}).setdefault(
        est['obj_id'], []).append(est)
    needed = set(t['obj_id'] for t in targets)
    engine = ErrorEngine(p, dataset, dp_model, dp_split, needed, device)
    n_ests = 0
    scenes = {}
    # Error files are written by one background thread: encoding JSON holds the GIL, but the main
    # thread spends the next scene mostly inside the PNG decoders and the GPU calls, which do not.
    writer = ThreadPoolExecutor(max_workers=1) if (write and rank == 0) else None
    pending = []
    # Pass 1 (host): flatten every scene and announce the depth images its pairs will read, so
    # the GPU PNG decoder gets groups of thousands of images whatever the scene size.
    plan = []
    for scene_id, scene_targets in targets_org.items():
      scene_camera = _memo(inout.load_scene_camera, dp_split['scene_camera_tpath'].format(scene_id=scene_id))
      scene_gt = _memo(inout.load_scene_gt, dp_split['scene_gt_tpath'].format(scene_id=scene_id))
      sp = flatten_scene_cached(
        p, scene_id, scene_targets, ests_org.get(scene_id, {}), scene_gt,
        (os.path.join(p['results_path'], result_filename),
         os.path.join(dp_split['base_path'], p['targets_filename']),
         dp_split['scene_gt_tpath'].format(scene_id=scene_id)))
      sel = _shard_pairs(sp, rank, world)
      engine.register_scene(scene_id, sp, sel)
      plan.append((scene_id, scene_camera, sp, sel))
    # Pass 2: the batched kernels, scene by scene.
    for scene_id, scene_camera, sp, sel in plan:
      n_ests += sp.n_ests
      log('Calculating error {} - method: {}, dataset: {}{}, scene: {}, pairs: {}'.format(
        p['error_type'], method, dataset, ' - ' + split_type if split_type else '', scene_id,
        len(sp.rec)))
      errs = engine.compute(scene_id, sp, scene_camera, sel)
      if world > 1:
        errs = _gather_errors(errs, sel, len(sp.rec), engine.device, sp=sp, world=world)
      records = fill_records(sp, errs)
      scenes[scene_id] = records
      if writer is not None:
        pending.append(writer.submit(save_scene_errors, p, dataset, result_name, scene_id, records, log,
                                     sp, errs))
    engine.close()
    if writer is not None:
      for f in pending:
        f.result()          # re-raises I/O errors
      writer.shutdown()
    all_out[result_name] = scenes
    log('Calculation of errors for {} estimates took {}s.'.format(n_ests, time.time() - t_start))
  log('Done.')
  return all_out


def main(argv=None):
  p = parse_args(argv)
  misc.log('-----------')
  misc.log('Parameters:')
  for k, v in p.items():
    misc.log('- {}: {}'.format(k, v))
  misc.log('-----------')
  world = int(os.environ.get('WORLD_SIZE', '1'))
  if world > 1:
    import torch.distributed as dist
    from . import sharding
    sharding.init_from_env()
  calc_errors(p)
  if world > 1:
    dist.destroy_process_group()


if __name__ == '__main__':
  main()

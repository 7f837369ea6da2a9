#!/usr/bin/env python
"""Benchmark of the pose-error hot path: pose-error pairs/sec (VSD + MSSD + MSPD).

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference ...                      (CPU oracle port on all host cores)

Workload at every N (weak scaling, per GPU): BASELINE.json configs[1] -- VSD (delta = 15 mm,
taus 0.05..0.5, step cost) on 640x480 synthetic depth, 1 object / image, 10,000 pose pairs (one
distinct depth image per pair) -- with MSSD and MSPD evaluated on the same pairs using the 4
symmetry transforms of configs[0].  A "step" is one pass of the hot path over all pairs; a pair
counts once when all three errors are computed.  Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)

H, W = 480, 640
N_OBJ = 8
DEPTH_SCALE = 0.1          # uint16 samples x 0.1 = mm (BOP depth_scale convention)
DELTA = 15.0
TAUS = [float(t) for t in np.arange(0.05, 0.51, 0.05)]
N_SYMS = 4
WORKLOAD = ('cfg2: VSD(delta=15,taus=.05-.5,step)+MSSD+MSPD(S=4), 640x480, 1 obj/image, '
            '%d pose pairs/GPU, 8 models of 4902 verts/9800 faces')


def build_models():
  from bop_toolkit_b200 import synth, misc
  models = []
  for k in range(N_OBJ):
    diameter = 80.0 + (300.0 - 80.0) * k / (N_OBJ - 1)
    pts, faces = synth.make_mesh(50, 100, diameter, seed=100 + k)
    info = {'diameter': diameter, 'symmetries_discrete': synth.discrete_z_symmetries(N_SYMS)}
    syms = misc.get_symmetry_transformations(info, 0.01)
    models.append(dict(obj_id=k + 1, pts=pts, faces=faces, syms=syms, diameter=diameter))
  return models


def build_pairs(n_pairs, seed, n_images=0):
  """n_images = 0: every pair has its own image (cfg2 as quoted).  n_images > 0: the "shared
  depth_test" variant of SURVEY 8d -- n_images images with one GT instance each and
  n_pairs / n_images estimates per instance (the perturbations of the per-pair workload applied
  to the image's GT pose)."""
  from bop_toolkit_b200 import synth
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n_pairs, seed=seed)
  obj = (np.arange(n_pairs) % N_OBJ).astype(np.int32)
  if n_images:
    per = n_pairs // n_images
    first = (np.arange(n_pairs) // per) * per          # the pair whose GT the image shows
    dR = np.einsum('nji,njk->nik', R_g, R_e)           # R_g^T R_e
    dt = t_e - t_g
    R_g, t_g, obj = R_g[first], t_g[first], obj[first]
    R_e, t_e = np.einsum('nij,njk->nik', R_g, dR), t_g + dt
  return R_e, t_e, R_g, t_g, obj


# ----------------------------------------------------------------------------- clocks
class ClockSampler(object):
  """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
  Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
       'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
       'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

  def __init__(self, gpu_index):
    self.gpu = gpu_index
    self.rows = []
    self.proc = None

  def start(self):
    try:
      self.proc = subprocess.Popen(
        ['nvidia-smi', '-i', str(self.gpu), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits',
         '-lms', '100'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
      self.t = threading.Thread(target=self._read, daemon=True)
      self.t.start()
    except Exception:
      self.proc = None

  def _read(self):
    for line in self.proc.stdout:
      self.rows.append((time.time(), line.strip()))

  def stop(self):
    if self.proc is not None:
      self.proc.terminate()
      try:
        self.proc.wait(timeout=2)
      except Exception:
        pass

  def summary(self, t0, t1):
    sm, mx, reasons = [], [], set()
    for ts, line in self.rows:
      if ts < t0 or ts > t1 + 0.2:
        continue
      f = [x.strip() for x in line.split(',')]
      if len(f) < 9:
        continue
      try:
        sm.append(float(f[1])); mx.append(float(f[2]))
      except ValueError:
        continue
      for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), f[5:9]):
        if v.lower().startswith('active'):
          reasons.add(name)
    if not sm:
      return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'samples': 0}
    return {'sm_mhz': float(np.median(sm)), 'sm_max_mhz': float(max(mx)), 'reasons': sorted(reasons),
            'samples': len(sm)}


# ----------------------------------------------------------------------- CPU (oracle) arm
# The reference's pose_error.py + eval_calc_errors.py path, restated (oracle/), timed on the host
# cores.  Everything a worker needs -- models, poses, the renderer with its models and the depth
# images of the sample -- is built in the parent BEFORE the timed pool forks, so the workers
# inherit it (copy-on-write) and a task is an index range: nothing but a few integers goes out
# and a few floats per pair come back.  (r01 shipped a 1.2 MB depth image with every task through
# the pool's single feeder thread, which capped the arm at ~8x one core whatever the core count.)
def _cpu_init():
  os.environ['OMP_NUM_THREADS'] = '1'
  os.environ['OPENBLAS_NUM_THREADS'] = '1'
  os.environ['MKL_NUM_THREADS'] = '1'
  try:
    from threadpoolctl import threadpool_limits
    _CPU['tp_limit'] = threadpool_limits(1)   # numpy's BLAS is already loaded in a forked worker
  except Exception:
    pass


_CPU = {}


def _cpu_prepare(n_pairs_cfg, sample, seed):
  """Untimed, parent process: models, pose pairs, the oracle renderer."""
  from oracle import renderer_oracle
  renderer_oracle.build()
  _CPU['models'] = build_models()
  _CPU['pairs'] = build_pairs(n_pairs_cfg, seed)
  ren = _CPU['ren'] = renderer_oracle.RendererOracle(W, H)
  for mm in _CPU['models']:
    ren.add_object(mm['obj_id'], pts=mm['pts'], faces=mm['faces'])
  _CPU['seed'] = seed
  _CPU['depth'] = np.zeros((sample, H, W), dtype=np.float32)


def _cpu_setup_range(args):
  """Untimed: synthesises the depth images of sample pairs lo..hi-1 with the oracle renderer."""
  lo, hi = args
  from bop_toolkit_b200 import synth
  from oracle import renderer_oracle
  R_e, t_e, R_g, t_g, obj = _CPU['pairs']
  K = synth.LM_K
  out = []
  for i in range(lo, hi):
    m = _CPU['models'][obj[i]]
    d = renderer_oracle.render_depth(m['pts'], m['faces'], R_g[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2], H, W)
    d = synth.degrade_depth(d, seed=_CPU['seed'] * 100003 + i)
    d16 = np.round(d / DEPTH_SCALE).astype(np.uint16)
    out.append(d16)
  return lo, np.stack(out) if out else np.zeros((0, H, W), np.uint16)


def _cpu_eval_pair(i):
  """Timed: what scripts/eval_calc_errors.py does for one (est, gt) pair for vsd + mssd + mspd
  (:281-389), with the oracle port of pose_error.py and the C oracle rasteriser as the renderer."""
  from bop_toolkit_b200 import synth
  from oracle import pose_error_oracle as po
  R_e, t_e, R_g, t_g, obj = _CPU['pairs']
  m = _CPU['models'][obj[i]]
  ren = _CPU['ren']
  depth = _CPU['depth'][i]
  K = synth.LM_K
  te, tg = t_e[i].reshape(3, 1), t_g[i].reshape(3, 1)
  if po.overlapping_sphere_projections(0.5 * m['diameter'], t_e[i], t_g[i]):
    e_vsd = po.vsd(R_e[i], te, R_g[i], tg, depth, K, DELTA, TAUS, True, m['diameter'], ren, m['obj_id'], 'step')
  else:
    e_vsd = [1.0] * len(TAUS)
  if np.linalg.norm(te - tg) < m['diameter']:
    e_mssd = po.mssd(R_e[i], te, R_g[i], tg, m['pts'], m['syms'])
  else:
    e_mssd = float('inf')
  e_mspd = po.mspd(R_e[i], te, R_g[i], tg, K, m['pts'], m['syms'])
  return i, e_vsd, e_mssd, e_mspd


def _cpu_eval_range(args):
  lo, hi = args
  return [_cpu_eval_pair(i) for i in range(lo, hi)]


def _ranges(n, parts):
  edges = [n * k // parts for k in range(parts + 1)]
  return [(edges[k], edges[k + 1]) for k in range(parts)]


def cpu_arm(n_pairs_cfg, per_core, cores, steps, warmup, seed=0, one_core_pairs=64):
  """Times the oracle port on `cores` worker processes, each holding a fixed shard of
  `per_core` sample pairs, and on one core alone.  Returns a dict (pairs/s for both, the
  scaling efficiency between them, ms per step, the results of the last step)."""
  import multiprocessing as mp
  cores = max(1, int(cores))
  sample = per_core * cores
  _cpu_prepare(n_pairs_cfg, sample, seed)
  ctx = mp.get_context('fork')
  # depth images of the sample: rendered by a set-up pool (untimed), kept in the parent
  with ctx.Pool(cores, initializer=_cpu_init) as pool:
    for lo, block in pool.map(_cpu_setup_range, _ranges(sample, 4 * cores), chunksize=1):
      _CPU['depth'][lo:lo + block.shape[0]] = block.astype(np.float32) * np.float32(DEPTH_SCALE)
  # one core, in this process (also loads oracle/liboracle.so here, where the driver's hook sees it)
  _cpu_init()
  n1 = min(sample, one_core_pairs)
  _cpu_eval_range((0, min(4, n1)))
  t0 = time.perf_counter()
  _cpu_eval_range((0, n1))
  one_core = n1 / (time.perf_counter() - t0)
  def time_pool(workers, n_steps, n_warm):
    """`workers` processes, each with its own shard of per_core pairs -> (pairs/s, seconds, results)."""
    shards = _ranges(per_core * workers, workers)
    with ctx.Pool(workers, initializer=_cpu_init) as pool:   # forked now: inherits _CPU incl. the depth images
      for _ in range(max(1, n_warm)):
        pool.map(_cpu_eval_range, shards, chunksize=1)
      t0 = time.perf_counter()
      for _ in range(n_steps):
        out = [r for part in pool.map(_cpu_eval_range, shards, chunksize=1) for r in part]
      dt = time.perf_counter() - t0
    return per_core * workers * n_steps / dt, dt, out

  half = None
  if cores == 1:
    for _ in range(warmup):
      _cpu_eval_range((0, sample))
    t0 = time.perf_counter()
    for _ in range(steps):
      res = _cpu_eval_range((0, sample))
    dt = time.perf_counter() - t0
    v = sample * steps / dt
  else:
    v, dt, res = time_pool(cores, steps, warmup)
    if cores >= 4 and v < 0.8 * cores * one_core:
      # Where does the scaling go?  The same shards on half the workers: if these scale, the box's
      # "cores" are hardware threads sharing execution units and caches (the arm is numpy on
      # 2.4 MB float64 images: memory-bound), not a serial section in this harness.
      hv, _, _ = time_pool(cores // 2, 1, 1)
      half = {'cores': cores // 2, 'value': hv, 'scaling_efficiency': hv / ((cores // 2) * one_core)}
  out = {'value': v, 'ms_per_step': 1e3 * dt / steps, 'one_core': one_core, 'cores': cores, 'sample': sample,
         'scaling_efficiency': v / (cores * one_core), 'results': res}
  if half:
    out['half_cores'] = half
  return out


def host_cores():
  try:
    return max(1, len(os.sched_getaffinity(0)))
  except Exception:
    return os.cpu_count() or 1


# ------------------------------------------------------------------------------ GPU arm
def make_depth_gpu(batch, torch, ms, obj, pose_gt, t_gt, diam, K, device, seed, chunk=400):
  """Synthetic sensor depth (uint16 samples, scale 0.1 mm) for every pair, generated on the GPU:
  GT render over a tilted background plane + N(0, 2 mm) noise + 3 % missing pixels + an occluder
  box in front of ~30 % of the objects.  Not timed."""
  P = pose_gt.shape[0]
  gen = torch.Generator(device=device)
  gen.manual_seed(seed)
  d16 = torch.empty((P, H, W), dtype=torch.int16, device=device)
  ys = torch.arange(H, device=device, dtype=torch.float32).view(1, H, 1)
  xs = torch.arange(W, device=device, dtype=torch.float32).view(1, 1, W)
  cams = torch.tensor([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], dtype=torch.float64, device=device)
  for c0 in range(0, P, chunk):
    c1 = min(P, c0 + chunk)
    n = c1 - c0
    d = batch.render_batch(ms, obj[c0:c1], pose_gt[c0:c1], cams, H, W)
    u = lambda lo, hi: torch.rand((n, 1, 1), device=device, generator=gen) * (hi - lo) + lo
    bg = u(1200.0, 2000.0) + u(-0.3, 0.3) * (xs - W / 2) + u(-0.3, 0.3) * (ys - H / 2)
    out = torch.where(d > 0, d, bg)
    tz = t_gt[c0:c1, 2].view(n, 1, 1).float()
    pu = (t_gt[c0:c1, 0].view(n, 1, 1).float() / tz) * float(K[0, 0]) + float(K[0, 2])
    pv = (t_gt[c0:c1, 1].view(n, 1, 1).float() / tz) * float(K[1, 1]) + float(K[1, 2])
    half = 0.25 * diam[c0:c1].view(n, 1, 1).float() * float(K[0, 0]) / tz
    ou, ov = pu + u(-1.0, 1.0) * half, pv + u(-1.0, 1.0) * half
    occ = (u(0.0, 1.0) < 0.3) & ((xs - ou).abs() < half) & ((ys - ov).abs() < half)
    out = torch.where(occ, 0.6 * tz.expand_as(out), out)
    out = out + 2.0 * torch.randn(out.shape, device=device, generator=gen)
    out = torch.where(torch.rand(out.shape, device=device, generator=gen) < 0.03, torch.zeros_like(out), out)
    v = torch.round(out / DEPTH_SCALE).clamp_(0, 65535).to(torch.int32)
    d16[c0:c1] = torch.where(v >= 32768, v - 65536, v).to(torch.int16)
  return d16


def kernel_counters():
  """Per-pair instruction and DRAM counts of the dominant kernel from ONE committed ncu --set full
  capture (profiles/kernel_counters.json names the capture); constants, not measured in this run."""
  try:
    return json.load(open(os.path.join(ROOT, 'profiles', 'kernel_counters.json')))
  except Exception:
    return {}


def gpu_arm(args):
  import torch
  import torch.distributed as dist
  from bop_toolkit_b200 import _lib, batch, misc, sharding, synth

  world = int(os.environ.get('WORLD_SIZE', '1'))
  rank = int(os.environ.get('RANK', '0'))
  local_rank = int(os.environ.get('LOCAL_RANK', '0'))
  if world > 1:
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
  assert torch.cuda.is_available(), 'bench.py needs a GPU (there is no CPU fallback)'
  torch.cuda.set_device(local_rank)
  from bop_toolkit_b200 import affinity
  bound = None if os.environ.get('BOP_NO_BIND') else affinity.bind_to_gpu(local_rank)
  device = torch.device('cuda', local_rank)
  _lib.require_cuda()
  lib = _lib.lib()
  if os.environ.get('BOP_TUNE'):   # "tile_px,threads,ctas_per_sm,small_tri_px" (0 = keep)
    _lib.check(lib.bop_vsd_set_tuning(*[int(x) for x in os.environ['BOP_TUNE'].split(',')]))
  P = args.pairs
  K = synth.LM_K

  # ---- per-GPU model residency + this rank's shard of pose pairs (weak scaling) -------------
  # `ms`: the drop-in default, no back-face culling = the reference's GL state (renderer_vispy.py:470).
  # `ms_cull`: the opt-in cull='auto' (closed, consistently wound meshes only; verified by probe renders).
  models = build_models()
  ms, ms_cull = batch.ModelSet(device), batch.ModelSet(device)
  for m in models:
    ms.add(m['obj_id'], m['pts'], m['faces'], m['syms'], m['diameter'], cull=False)
    ms_cull.add(m['obj_id'], m['pts'], m['faces'], m['syms'], m['diameter'], cull='auto')
  n_verts, n_faces = models[0]['pts'].shape[0], models[0]['faces'].shape[0]
  n_images = args.images if args.images else P
  assert P % n_images == 0, '--images must divide --pairs'
  per_im = P // n_images
  R_e, t_e, R_g, t_g, obj_np = build_pairs(P, seed=rank, n_images=args.images)
  pe_np, pg_np = batch.pack_poses(R_e, t_e), batch.pack_poses(R_g, t_g)
  diam_np = np.array([models[o]['diameter'] for o in obj_np])
  gate_ok = np.array([misc.overlapping_sphere_projections(0.5 * diam_np[i], t_e[i], t_g[i]) for i in range(P)])
  n_active = int(gate_ok.sum())

  pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
  h_pe, h_pg, h_obj = pin(pe_np), pin(pg_np), pin(obj_np)
  h_im = pin((np.arange(P) // per_im).astype(np.int32))
  h_K = pin(np.tile(K.reshape(1, 9), (n_images, 1)))
  h_Kp = h_K if n_images == P else pin(np.tile(K.reshape(1, 9), (P, 1)))   # mspd takes K per pair
  h_scale = pin(np.full(n_images, DEPTH_SCALE, dtype=np.float32))
  d_pe, d_pg, d_obj = h_pe.to(device), h_pg.to(device), h_obj.to(device)
  d_im, d_K, d_Kp = h_im.to(device), h_K.to(device), h_Kp.to(device)
  d16 = make_depth_gpu(batch, torch, ms, d_obj[::per_im].contiguous(), d_pg[::per_im].contiguous(),
                       torch.from_numpy(t_g[::per_im].copy()).to(device),
                       torch.from_numpy(diam_np[::per_im].copy()).to(device), K, device, seed=1234 + rank)
  h_d16 = torch.empty((n_images, H, W), dtype=torch.int16).pin_memory()
  h_d16.copy_(d16)
  # Resident float32 millimetres: bit-identical to what the u16 path computes (u16 * 0.1f).
  d_f32 = torch.empty((n_images, H, W), dtype=torch.float32, device=device)
  for c0 in range(0, n_images, 1000):
    v = d16[c0:c0 + 1000].to(torch.int32)
    d_f32[c0:c0 + 1000] = torch.where(v < 0, v + 65536, v).to(torch.float32) * DEPTH_SCALE
  del d16
  ids = torch.stack([torch.full((P,), rank, dtype=torch.int32, device=device), d_im, d_obj + 1,
                     torch.zeros(P, dtype=torch.int32, device=device),
                     torch.zeros(P, dtype=torch.int32, device=device)], dim=1)
  h_out = torch.empty((P, len(TAUS) + 2), dtype=torch.float64).pin_memory()
  torch.cuda.synchronize()

  def step_resident(mset=ms):
    vsd = batch.vsd_batch(mset, d_obj, d_pe, d_pg, d_f32, d_im, d_K, DELTA, TAUS, True, 'step', 'bop19', gate=True)
    mssd = batch.mssd_batch(mset, d_obj, d_pe, d_pg, gate=True)
    mspd = batch.mspd_batch(mset, d_obj, d_pe, d_pg, d_Kp)
    errs = torch.cat([vsd, mssd.unsqueeze(1), mspd.unsqueeze(1)], dim=1)
    if world > 1:   # the path's one exchange step: all-gather of fixed-width error records
      return sharding.gather_records(ids, errs, counts=[P] * world)[1]
    return errs

  def step_e2e(zero_copy=True):
    # HOST buffers in, host results out: H2D of every input of the step + D2H of the errors.
    # zero_copy: the pinned uint16 depth images are handed to the kernel as they are; it pulls
    # only the samples under the rendered silhouettes across PCIe (no whole-image copies).
    vsd = batch.vsd_batch(ms, h_obj, h_pe, h_pg, h_d16, h_im, h_K, DELTA, TAUS, True, 'step', 'bop19',
                          gate=True, depth_scale=h_scale, zero_copy=zero_copy)
    o, a, b = (x.to(device, non_blocking=True) for x in (h_obj, h_pe, h_pg))
    mssd = batch.mssd_batch(ms, o, a, b, gate=True)
    mspd = batch.mspd_batch(ms, o, a, b, h_Kp)
    errs = torch.cat([vsd, mssd.unsqueeze(1), mspd.unsqueeze(1)], dim=1)
    if world > 1:
      sharding.gather_records(ids, errs, counts=[P] * world)
    h_out.copy_(errs, non_blocking=True)
    return errs

  # The same end-to-end step on an SM partition (sm_partition.py): the pairs are split in two
  # sub-batches in proportion to the SMs of the two green contexts and issued on their streams;
  # nothing orders the two streams against each other from step to step, so one partition's
  # PCIe-bound pixel pass runs while the other one's geometry / MSSD / MSPD kernels compute.
  part_state = {}

  def part_setup(sms_a):
    from bop_toolkit_b200 import sm_partition
    part = sm_partition.SmPartition(local_rank, sms_a)
    n_a = int(round(P * part.sm_counts[0] / float(sum(part.sm_counts))))
    # a partition may take its share in several pieces per step: pieces of different length on the
    # two streams keep them from falling into lockstep (both in their PCIe-bound phase at once)
    pieces = [max(1, int(x)) for x in args.e2e_pieces.split(',')]
    cuts = []
    for (lo, hi), n_p in zip(((0, n_a), (n_a, P)), pieces):
      edges = [lo + (hi - lo) * j // n_p for j in range(n_p + 1)]
      cuts.append([slice(edges[j], edges[j + 1]) for j in range(n_p) if edges[j + 1] > edges[j]])
    part_state.update(part=part, cuts=cuts, ws=[batch.Workspace(device) for _ in range(max(2, args.e2e_depth))])
    return part

  def step_e2e_part_default(only=(0, 1)):
    part, outs = part_state['part'], []
    for k in only:
      with torch.cuda.stream(part.streams[k]), batch.private_workspace(part_state['ws'][k]):
        for sl in part_state['cuts'][k]:
          vsd = batch.vsd_batch(ms, h_obj[sl], h_pe[sl], h_pg[sl], h_d16, h_im[sl], h_K, DELTA, TAUS, True, 'step',
                                'bop19', gate=True, depth_scale=h_scale, zero_copy=True)
          o, a, b = (x[sl].to(device, non_blocking=True) for x in (h_obj, h_pe, h_pg))
          mssd = batch.mssd_batch(ms, o, a, b, gate=True)
          mspd = batch.mspd_batch(ms, o, a, b, h_Kp[sl])
          errs = torch.cat([vsd, mssd.unsqueeze(1), mspd.unsqueeze(1)], dim=1)
          h_out[sl].copy_(errs, non_blocking=True)
          outs.append(errs)
    if world > 1:   # the exchange runs on the main stream behind both partitions; they do not wait for it
      cur = torch.cuda.current_stream()
      for k in (0, 1):
        cur.wait_stream(part.streams[k])
      for o_ in outs:
        o_.record_stream(cur)
      sharding.gather_records(ids, torch.cat(outs, dim=0), counts=[P] * world)
    return outs

  # The same step as a two-stage pipeline over the two partitions: geometry + MSSD + MSPD of every
  # step on the compute partition, the PCIe-bound raster + pixel pass (bop_vsd_split_dev), the
  # assembly of the records and the D2H copy on the raster partition; two workspaces alternate.
  def step_e2e_pipe():
    part = part_state['part']
    s_r, s_c = part.streams
    if part_state.pop('pipeline_empty', False):
      # nothing is in flight (first step after a synchronisation): this step's geometry + MSSD + MSPD
      # may use the whole GPU instead of filling the pipeline from the small partition
      s_c = torch.cuda.current_stream()
    k = part_state['k'] = (part_state.get('k', 0) + 1) % len(part_state['ws'])
    with torch.cuda.stream(s_c), batch.private_workspace(part_state['ws'][k]):
      if part_state.get('done%d' % k) is not None:   # this workspace's previous raster pass
        s_c.wait_event(part_state['done%d' % k])
      vsd = batch.vsd_batch(ms, h_obj, h_pe, h_pg, h_d16, h_im, h_K, DELTA, TAUS, True, 'step', 'bop19',
                            gate=True, depth_scale=h_scale, zero_copy=True, raster_stream=s_r)
      o, a, b = (x.to(device, non_blocking=True) for x in (h_obj, h_pe, h_pg))
      mssd = batch.mssd_batch(ms, o, a, b, gate=True)
      mspd = batch.mspd_batch(ms, o, a, b, h_Kp)
      mssd.record_stream(s_r)
      mspd.record_stream(s_r)
      ev_c = torch.cuda.Event()
      ev_c.record(s_c)
    with torch.cuda.stream(s_r):
      s_r.wait_event(ev_c)
      errs = torch.cat([vsd, mssd.unsqueeze(1), mspd.unsqueeze(1)], dim=1)
      h_out.copy_(errs, non_blocking=True)
      ev_r = torch.cuda.Event()
      ev_r.record(s_r)
      part_state['done%d' % k] = ev_r
    if world > 1:
      cur = torch.cuda.current_stream()
      cur.wait_stream(s_r)
      errs.record_stream(cur)
      sharding.gather_records(ids, errs, counts=[P] * world)
    return [errs]

  def timed_part(steps, warmup, step_e2e_part=None):
    step_e2e_part = step_e2e_part or step_e2e_part_default
    part = part_state['part']
    cur = torch.cuda.current_stream()
    for _ in range(warmup):
      step_e2e_part()
    torch.cuda.synchronize()
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()
    l0 = batch.kernel_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for st in part.streams:
      st.wait_stream(cur)
    part_state['pipeline_empty'] = True
    for _ in range(steps):
      outs = step_e2e_part()
    for st in part.streams:
      cur.wait_stream(st)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
      dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if world > 1:
      dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), batch.kernel_launch_count() - l0, torch.cat(outs, dim=0)

  def timed(fn, steps, warmup):
    for _ in range(warmup):
      fn()
    torch.cuda.synchronize()
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()
    lib.bop_profile_reset()
    l0 = batch.kernel_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.time()
    e0.record()
    for _ in range(steps):
      out = fn()
    e1.record()
    torch.cuda.synchronize()
    w1 = time.time()
    if world > 1:
      dist.barrier()
    ms_total = e0.elapsed_time(e1)
    t = torch.tensor([ms_total], dtype=torch.float64, device=device)
    if world > 1:
      dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), batch.kernel_launch_count() - l0, (w0, w1), out

  def kernel_parts(steps):
    s_ms, s_n = ctypes_double(), ctypes_int()
    parts = {}
    for name, slot in (('fused', 0), ('geometry', 1), ('mssd', 2), ('mspd', 3)):
      lib.bop_profile_sum_ms(slot, steps, byref(s_ms), byref(s_n))
      parts[name] = s_ms.value / max(1, s_n.value)
    return parts

  sampler = ClockSampler(local_rank)
  if rank == 0:
    sampler.start()
    time.sleep(0.3)
  # (1) eager pass with the library's per-kernel CUDA events: where a step's time goes
  lib.bop_profile_enable(1)
  ms_eager, launches, _, out_res = timed(step_resident, args.steps, args.warmup)
  parts = kernel_parts(args.steps)
  lib.bop_profile_enable(0)
  # (2) the timed resident pass: the same step captured ONCE into a CUDA graph and replayed, so that
  # host-side launch jitter (python, driver) cannot starve the GPU between the 8 kernels of a step.
  # Falls back to the eager loop if capture is not possible (e.g. NCCL inside the step).
  graph_note = 'eager launches'
  ms_res, win_res = ms_eager, None
  use_graph = world == 1 and not os.environ.get('BOP_NO_GRAPH')
  if use_graph:
    try:
      side = torch.cuda.Stream(device)
      side.wait_stream(torch.cuda.current_stream())
      with torch.cuda.stream(side):
        step_resident()
      torch.cuda.current_stream().wait_stream(side)
      torch.cuda.synchronize()
      graph = torch.cuda.CUDAGraph()
      with torch.cuda.graph(graph):
        g_out = step_resident()
      l_graph = launches // max(1, args.steps)
      ms_res, _, win_res, _ = timed(graph.replay, args.steps, args.warmup)
      assert torch.equal(g_out, out_res), 'graph replay differs from the eager step'
      launches = l_graph * args.steps
      graph_note = 'CUDA graph replay of the step (%d kernels of this library per step)' % l_graph
    except Exception as e:   # keep the eager number
      graph_note = 'eager launches (graph capture failed: %s)' % (str(e).splitlines()[0][:120],)
      ms_res = ms_eager
  if win_res is None:
    ms_res, launches, win_res, out_res = timed(step_resident, args.steps, args.warmup)
  # (3) end to end through host buffers
  lib.bop_profile_enable(1)
  ms_e2e, _, win_e2e, out_e2e = timed(step_e2e, args.steps, args.warmup)
  e2e_parts = kernel_parts(args.steps)
  depth_pulled = batch.depth_bytes_fetched(device, _lib.DEPTH_U16)
  ms_e2e_copy, _, _, out_e2e_copy = timed(lambda: step_e2e(False), max(1, args.steps // 2), 1)
  # (3b) the same step on two SM partitions (skipped with --e2e-sms 0 or when green contexts are unavailable)
  e2e_part = None
  if args.e2e_sms > 0:
    lib.bop_profile_enable(0)
    try:
      part = part_setup(args.e2e_sms)
      if args.e2e_mode == 'pipe':   # every kernel has a partition to itself: the per-kernel events are meaningful
        lib.bop_profile_enable(1)
        lib.bop_profile_reset()
      ms_part, _, out_part = timed_part(args.steps, args.warmup, step_e2e_pipe if args.e2e_mode == 'pipe' else None)
      diag = None
      if args.e2e_mode == 'pipe':
        diag = {'kernels_ms': kernel_parts(args.steps)}
        lib.bop_profile_enable(0)
      if args.e2e_diag and args.e2e_mode != 'pipe':   # each sub-batch alone on its partition, with the library's per-kernel events
        diag = {}
        lib.bop_profile_enable(1)
        for k in (0, 1):
          for _ in range(2):
            step_e2e_part_default((k,))
          torch.cuda.synchronize()
          lib.bop_profile_reset()
          t0 = time.time()
          for _ in range(4):
            step_e2e_part_default((k,))
          torch.cuda.synchronize()
          diag['alone_%d' % k] = {'ms_per_step': (time.time() - t0) * 250.0, 'kernels_ms': kernel_parts(4)}
        lib.bop_profile_enable(0)
      e2e_part = {'ms': ms_part, 'sms': list(part.sm_counts), 'diag': diag, 'mode': args.e2e_mode,
                  'pairs': [[c.stop - c.start for c in cs] for cs in part_state['cuts']],
                  'same': bool(world > 1 or torch.equal(out_part, out_e2e))}
      # the partition lives until the process ends: torch's allocator keeps blocks tied to its streams
    except Exception as e:
      e2e_part = {'error': str(e).splitlines()[0][:200]}
    lib.bop_profile_enable(1)
  # (4) the opt-in back-face culling on the same step (closed synthetic meshes)
  ms_cull_t, _, _, out_cull = timed(lambda: step_resident(ms_cull), max(2, args.steps // 2), 2)
  cull_parts = kernel_parts(max(2, args.steps // 2))
  lib.bop_profile_enable(0)
  same = bool(torch.equal(out_res[:, :len(TAUS) + 2] if world == 1 else out_res[:0], out_e2e if world == 1 else out_e2e[:0]))
  cull_max_dev = float((out_cull[:, :len(TAUS)] - out_res[:, :len(TAUS)]).abs().max().item()) if world == 1 else None
  clocks = None
  if rank == 0:
    time.sleep(0.2)
    sampler.stop()
    clocks = sampler.summary(win_res[0], win_res[1])

  # (5) BASELINE configs[4]: strong-scaling leg (its own data; every rank takes part)
  cfg5 = None
  if args.cfg5_images > 0:
    del d_f32
    torch.cuda.empty_cache()
    sys.path.insert(0, os.path.join(ROOT, 'tools'))
    import cfg5_leg
    cfg5 = cfg5_leg.run(torch, dist, device, rank, world, args.cfg5_images, steps=1)

  if rank != 0:
    if world > 1:
      dist.destroy_process_group()
    return None

  value = P * world * args.steps / (ms_res * 1e-3)
  e2e_v = P * world * args.steps / (ms_e2e * 1e-3)
  peaks = {}
  try:
    peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
  except Exception:
    pass
  hbm_peak = float(peaks.get('hbm_gbs', 6650.0))
  fused_ms = parts['fused']
  kc = kernel_counters()
  sm_mhz = (clocks or {}).get('sm_mhz') or float(peaks.get('sm_max_mhz', 1965.0))
  n_sm = torch.cuda.get_device_properties(device).multi_processor_count
  # The fused kernel is bound by instruction issue, not by HBM (ncu: DRAM ~10-15 % busy): its
  # roofline is the SM's thread-instruction rate, SMs x 128 lanes x clock.  achieved = thread
  # instructions per launch (per-pair count of the committed ncu capture x pairs past the gate)
  # / the kernel's CUDA-event duration in THIS run.
  tinst = kc.get('fused_thread_inst_per_pair')
  winst = kc.get('fused_warp_inst_per_pair')
  peak_ti = n_sm * 128 * sm_mhz * 1e6 / 1e9                      # G thread-inst/s
  ach_ti = (tinst * n_active / (fused_ms * 1e-3) / 1e9) if tinst else None
  # Algorithmic bytes of a NON-fused pipeline: per evaluated pair 3 reads + 2
  # z-buffer writes of 4 B x H x W plus the mesh twice; gated pairs move nothing.
  alg_pair = 5 * 4 * H * W + 2 * (12 * n_verts + 12 * n_faces)
  eff_gbs = n_active * alg_pair / (fused_ms * 1e-3) / 1e9
  dram_pair = kc.get('fused_dram_bytes_per_pair')
  small = sum(x.numel() * x.element_size() for x in (h_obj, h_pe, h_pg, h_im, h_K, h_scale)) + \
    sum(x.numel() * x.element_size() for x in (h_obj, h_pe, h_pg, h_Kp))
  h2d = small + depth_pulled
  h2d_copy = small + h_d16.numel() * h_d16.element_size()
  e2e_copy_v = P * world * max(1, args.steps // 2) / (ms_e2e_copy * 1e-3)
  e2e_serial = {'value': e2e_v, 'ms_per_step': ms_e2e / args.steps, 'kernels_ms': e2e_parts,
                'mode': 'one stream: geometry, fused (PCIe-bound), MSSD, MSPD one after the other'}
  e2e_mode = 'pinned host buffers; depth pulled on demand over PCIe by the fused kernel'
  e2e_ms_step = ms_e2e / args.steps
  if e2e_part and 'ms' in e2e_part and e2e_part['same'] and e2e_part['ms'] < ms_e2e:
    e2e_v = P * world * args.steps / (e2e_part['ms'] * 1e-3)
    e2e_ms_step = e2e_part['ms'] / args.steps
    if e2e_part['mode'] == 'pipe':
      e2e_mode += ('; two-stage pipeline over two SM partitions (CUDA green contexts): the raster + pixel pass, the '
                   'record assembly and the D2H copy on %d SMs, geometry + MSSD + MSPD of the next step on %d SMs '
                   '(bop_vsd_split_dev; two workspaces alternate)' % (e2e_part['sms'][0], e2e_part['sms'][1]))
    else:
      e2e_mode += ('; the step is issued as sub-batches (%s + %s pairs) on the streams of two SM partitions '
                   '(CUDA green contexts, %d + %d SMs) that run independently, so one partition\'s PCIe-bound pixel '
                   'pass overlaps the other one\'s geometry / MSSD / MSPD kernels' %
                   (e2e_part['pairs'][0], e2e_part['pairs'][1], e2e_part['sms'][0], e2e_part['sms'][1]))
  line = {
    'metric': 'pose-error pairs/sec (VSD+MSSD+MSPD)', 'value': value, 'unit': 'pairs/s',
    'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
    'ms_per_step': ms_res / args.steps, 'higher_is_better': True, 'scaling': 'weak',
    'vs_baseline': None, 'dtype': 'f32 raster + f64 pixel/vertex arithmetic', 'data': 'synthetic',
    'config': {'workload': workload_config(P, args.images), 'pairs_per_gpu': P, 'pairs_past_gate': n_active,
               'culling': 'off (the reference\'s GL state; `cull_auto` below is the opt-in)',
               'launch': graph_note, 'eager_ms_per_step': ms_eager / args.steps,
               'l2': 'inputs larger than L2: resident pass reads its model pixels out of %.1f GB of depth; '
                     'e2e pass fetches ~%.2f GB of host depth; both write and re-read ~%.1f GB of assembled-'
                     'triangle lists per step (L2 = 126 MB)' % (n_images * H * W * 4 / 1e9, depth_pulled / 1e9,
                                                                P * 2 * n_faces * 40 / 1e9),
               'resident_depth': 'float32 mm in HBM', 'e2e_depth': 'uint16 samples + depth_scale (pinned host)',
               'exchange': 'all-gather of error records per step' if world > 1 else 'none',
               'cpu_affinity': bound},
    'e2e': {'value': e2e_v, 'unit': 'pairs/s', 'ms_per_step': e2e_ms_step,
            'h2d_bytes_per_step': int(h2d), 'd2h_bytes_per_step': int(h_out.numel() * 8),
            'matches_resident': same, 'kernels_ms': e2e_parts,
            'mode': e2e_mode, 'serial': e2e_serial, 'sm_partition': e2e_part,
            'full_copy': {'value': e2e_copy_v, 'h2d_bytes_per_step': int(h2d_copy),
                          'note': 'same call with whole-image H2D copies of the uint16 depth'}},
    'cull_auto': {'value': P * world * max(2, args.steps // 2) / (ms_cull_t * 1e-3), 'unit': 'pairs/s',
                  'ms_per_step': ms_cull_t / max(2, args.steps // 2), 'kernels_ms': cull_parts,
                  'max_abs_vsd_difference_vs_no_cull': cull_max_dev,
                  'note': 'same resident step (eager launches) with ModelSet.add(cull="auto"): back faces of closed, '
                          'consistently wound meshes skipped after probe renders agree; not the drop-in default'},
    'gpu_launches': int(launches),
    'clocks': clocks,
    'roofline': {'bound': 'issue', 'kernel': 'fused_kernel<STEP> (raster x2 + pixel pass)',
                 'achieved': ach_ti, 'peak': peak_ti, 'unit': 'Gthread-inst/s',
                 'frac': (ach_ti / peak_ti) if ach_ti else None,
                 'peak_source': '%d SMs x 128 lanes x %.0f MHz (SM clock sampled during the timed region)' % (n_sm, sm_mhz),
                 'kernel_ms': fused_ms, 'thread_inst_per_pair': tinst, 'warp_inst_per_pair': winst,
                 'issue_slot_frac': (winst * n_active / (fused_ms * 1e-3) / (n_sm * 4 * sm_mhz * 1e6)) if winst else None,
                 'lanes_per_inst': (tinst / winst) if tinst and winst else None,
                 'counters_source': kc.get('source'),
                 'traffic': (dram_pair * n_active) if dram_pair else None,
                 'traffic_note': 'DRAM bytes per launch = per-pair bytes of the same committed capture x pairs past the '
                                 'gate (a profile constant, not measured in this run)',
                 'hbm': {'physical_gbs': (dram_pair * n_active / (fused_ms * 1e-3) / 1e9) if dram_pair else None,
                         'peak_gbs': hbm_peak, 'peak_source': 'MEASURED_PEAKS.json hbm_gbs' if peaks else 'fallback 6650',
                         'physical_frac': (dram_pair * n_active / (fused_ms * 1e-3) / 1e9 / hbm_peak) if dram_pair else None,
                         'effective_gbs_vs_unfused': eff_gbs, 'effective_frac_vs_unfused': eff_gbs / hbm_peak,
                         'algorithmic_bytes_per_pair_unfused': alg_pair,
                         'note': 'SURVEY 8d counts the bytes of a NON-fused pipeline (3 reads + 2 z-buffer writes of '
                                 'H*W*4 B + mesh); the fused kernel keeps the rendered images in shared memory, so '
                                 'its effective rate exceeds the HBM peak and HBM is not its roofline'},
                 'other_kernels_ms': {k: v for k, v in parts.items() if k != 'fused'}},
  }
  if cfg5 is not None:
    line['cfg5'] = cfg5
  if world == 1 and not args.no_cpu_baseline and n_images == P:   # the CPU sample reproduces the per-pair-image workload
    cores = host_cores()
    r = cpu_arm(P, args.cpu_per_core, cores, 1, 1)
    line['cpu_baseline'] = {
      'value': r['value'], 'unit': 'pairs/s', 'cores': cores, 'kind': 'port',
      'one_core': r['one_core'], 'scaling_efficiency': r['scaling_efficiency'], 'half_cores': r.get('half_cores'),
      'sample': '%d of the %d pairs (%d per worker, every worker holds its shard before the clock starts), '
                'oracle port (numpy pose_error + C rasteriser) on %d processes; one_core = the first %d of them '
                'on one core' % (r['sample'], P, args.cpu_per_core, cores, min(r['sample'], 64))}
  if world > 1:
    dist.destroy_process_group()
  return line


def ctypes_double():
  import ctypes
  return ctypes.c_double(0.0)


def ctypes_int():
  import ctypes
  return ctypes.c_int(0)


def byref(x):
  import ctypes
  return ctypes.byref(x)


def workload_config(P, n_images=0):
  per_im = P // n_images if n_images else 1
  w = WORKLOAD % P
  if n_images and n_images != P:
    w = w.replace('1 obj/image', '%d shared images x %d estimates' % (n_images, per_im))
  return w


def reference_arm(args):
  """The reference's CPU implementation of the path = the oracle port (the reference is Python
  + OpenGL and cannot run on the box), on all host cores, bounded sample per step."""
  rank = int(os.environ.get('RANK', '0'))
  if rank != 0:
    return None
  cores = min(host_cores(), 64)
  r = cpu_arm(args.pairs, args.cpu_per_core, cores, max(1, args.steps), min(args.warmup, 1))
  v = r['value']
  cb = {'value': v, 'unit': 'pairs/s', 'cores': cores, 'kind': 'port', 'one_core': r['one_core'],
        'scaling_efficiency': r['scaling_efficiency'], 'half_cores': r.get('half_cores'),
        'sample': '%d of the %d pairs per step (%d per worker, held by the worker before the clock starts), '
                  'oracle port on %d processes' % (r['sample'], args.pairs, args.cpu_per_core, cores)}
  return {
    'impl': 'reference', 'metric': 'pose-error pairs/sec (VSD+MSSD+MSPD)', 'value': v, 'unit': 'pairs/s',
    'n_gpus': int(os.environ.get('WORLD_SIZE', '1')), 'steps': args.steps, 'warmup': args.warmup,
    'ms_per_step': r['ms_per_step'], 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
    'dtype': 'f64', 'data': 'synthetic', 'config': {'workload': workload_config(args.pairs)},
    'cpu_baseline': cb,
    'e2e': {'value': v, 'unit': 'pairs/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    'gpu_launches': 0,
  }


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--gpus', type=int, default=1)
  ap.add_argument('--steps', type=int, default=10)
  ap.add_argument('--warmup', type=int, default=3)
  ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
  ap.add_argument('--pairs', type=int, default=10000)
  ap.add_argument('--images', type=int, default=0,
                  help='0: one image per pair (cfg2 as quoted); N: N shared images x pairs/N estimates each')
  ap.add_argument('--cpu-per-core', type=int, default=32, help='sample pairs per host worker of the CPU arm')
  ap.add_argument('--no-cpu-baseline', action='store_true')
  ap.add_argument('--e2e-mode', default='pipe', choices=['pipe', 'part'],
                  help='pipe: raster pass on one SM partition, geometry + point kernels on the other; part: two independent sub-batch streams')
  ap.add_argument('--e2e-depth', type=int, default=2, help='workspaces the pipelined end-to-end leg alternates')
  ap.add_argument('--e2e-pieces', default='1,1', help='sub-batches per step on the two partitions')
  ap.add_argument('--e2e-diag', action='store_true', help='time each sub-batch alone on its SM partition')
  ap.add_argument('--e2e-sms', type=int, default=96,
                  help='SMs of the first of the two partitions of the end-to-end leg (0 = one stream only)')
  ap.add_argument('--cfg5-images', type=int, default=20000,
                  help='images of the cfg5 strong-scaling leg (BASELINE configs[4]); 0 skips it')
  args = ap.parse_args()
  if args.impl == 'reference':
    line = reference_arm(args)
  else:
    args.warmup = max(args.warmup, 3)
    line = gpu_arm(args)
  if line is not None:
    print(json.dumps(line))


if __name__ == '__main__':
  main()

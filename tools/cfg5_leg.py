#!/usr/bin/env python
"""BASELINE.json configs[4] as a STRONG-scaling bench leg, all in memory: a synthetic YCB-V-scale
test set (default 20,000 images x ~5 instances of 21 models, 2 estimates per target) is split over
the ranks by image (sharding.assign_images, greedy LPT on pairs per image), every rank evaluates
vsd + mssd + mspd for its (estimate, GT) pairs against the depth images it holds, ONE all-gather
moves the fixed-width error records (unequal shard sizes), and rank 0 computes the BOP19 recalls
at 120 settings with the in-process score stage (score_sweep.ThresholdSweep).

What is shared by all ranks -- poses, object ids, scores -- comes from one seeded numpy generator,
and the depth image of global image i depends on i alone (renders of its GT instances + a
background plane + counter-based noise), so the gathered records and the recalls are the same
bits at every world size: the line carries their SHA-256.

Used by bench.py (`"cfg5": {...}` in the JSON line) and tests/test_cfg5_leg.py; the loop it
shards is scripts/eval_calc_errors.py:216-389 feeding scripts/eval_bop19_pose.py:147-230.
"""
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)

H, W = 480, 640
N_OBJ = 21
N_SCENES = 12
DEPTH_SCALE = 0.1
DELTA = 15.0
TAUS = [float(t) for t in np.arange(0.05, 0.51, 0.05)]
SYMS_OF = [1, 1, 1, 1, 2, 1, 4, 1, 1, 1, 1, 1, 2, 1, 1, 1, 1, 1, 4, 1, 1]
ESTS_PER_TARGET = 2


def build_models(mesh_res=(50, 100)):
  from bop_toolkit_b200 import synth, misc
  models = []
  for k in range(N_OBJ):
    diameter = 60.0 + (300.0 - 60.0) * ((k * 8) % N_OBJ) / (N_OBJ - 1)
    pts, faces = synth.make_mesh(mesh_res[0], mesh_res[1], diameter, seed=300 + k)
    info = {'diameter': diameter}
    if SYMS_OF[k] > 1:
      info['symmetries_discrete'] = synth.discrete_z_symmetries(SYMS_OF[k])
    syms = misc.get_symmetry_transformations(info, 0.01)
    models.append(dict(obj_id=k + 1, pts=pts, faces=faces, syms=syms, diameter=diameter))
  return models


def build_table(n_images, seed=7):
  """Everything but the depth images, identical on every rank.  Returns a dict of flat arrays:
  instances (GT poses) and pairs (estimate x GT of the same object in the same image)."""
  from bop_toolkit_b200 import synth
  rng = np.random.default_rng(seed)
  K = synth.LM_K
  n_inst = rng.integers(3, 8, size=n_images)                     # 3..7 instances, 5 on average
  inst_im = np.repeat(np.arange(n_images), n_inst)
  inst_gt = np.concatenate([np.arange(c) for c in n_inst])       # gt_id inside the image
  order = np.argsort(rng.random((n_images, N_OBJ)), axis=1)      # distinct objects per image
  inst_obj = order[inst_im, inst_gt].astype(np.int32)            # model row 0..20
  n = len(inst_im)
  R_g = synth.random_rotations(rng, n)
  z = rng.uniform(600.0, 1400.0, size=n)
  u = rng.uniform(0.1 * W, 0.9 * W, size=n)
  v = rng.uniform(0.1 * H, 0.9 * H, size=n)
  t_g = np.stack([(u - K[0, 2]) / K[0, 0] * z, (v - K[1, 2]) / K[1, 1] * z, z], axis=1)
  # estimates: ESTS_PER_TARGET per instance; est 0 close (score 0.9), est 1 looser (score 0.5), 8 % gross outliers
  pe_R, pe_t, p_inst, p_est, p_score = [], [], [], [], []
  for e in range(ESTS_PER_TARGET):
    ang = np.radians(rng.uniform(0.0, 6.0 + 18.0 * e, size=n))
    ax = rng.normal(size=(n, 3))
    ax /= np.linalg.norm(ax, axis=1, keepdims=True)
    Kx = np.zeros((n, 3, 3))
    Kx[:, 0, 1], Kx[:, 0, 2], Kx[:, 1, 0] = -ax[:, 2], ax[:, 1], ax[:, 2]
    Kx[:, 1, 2], Kx[:, 2, 0], Kx[:, 2, 1] = -ax[:, 0], -ax[:, 1], ax[:, 0]
    dR = np.eye(3)[None] + np.sin(ang)[:, None, None] * Kx + (1 - np.cos(ang))[:, None, None] * (Kx @ Kx)
    R_e = R_g @ dR
    t_e = t_g + rng.normal(size=(n, 3)) * (np.array([4.0, 4.0, 12.0]) * (1 + 2 * e))[None]
    out = rng.random(n) < 0.08
    t_e[out] += rng.normal(size=(int(out.sum()), 3)) * 150.0
    pe_R.append(R_e); pe_t.append(t_e); p_inst.append(np.arange(n)); p_est.append(np.full(n, e))
    p_score.append(np.full(n, 0.9 - 0.4 * e) + 0.05 * rng.random(n))
  p_inst = np.concatenate(p_inst)
  srt = np.lexsort((np.concatenate(p_est), p_inst))               # pairs grouped by image, then instance
  p_inst = p_inst[srt]
  R_e, t_e = np.concatenate(pe_R)[srt], np.concatenate(pe_t)[srt]
  p_score = np.concatenate(p_score)[srt]
  p_im = inst_im[p_inst]
  first = np.searchsorted(p_im, np.arange(n_images))
  p_est_id = np.arange(len(p_im)) - first[p_im]                   # est_id inside the image
  return dict(n_images=n_images, n_inst=n_inst, inst_im=inst_im, inst_gt=inst_gt, inst_obj=inst_obj,
              R_g=R_g, t_g=t_g, p_inst=p_inst, p_im=p_im, p_est_id=p_est_id, p_score=p_score, R_e=R_e, t_e=t_e)


def scene_of(im):
  return 48 + (im % N_SCENES), im // N_SCENES      # (scene_id, im_id) of global image `im`


def _hash_u64(torch, x):
  """splitmix64-style counter hash on int64 tensors (wrapping arithmetic)."""
  x = x * -7046029254386353131            # 0x9E3779B97F4A7C15
  x = (x ^ (x >> 30)) * -4658895280553007687   # 0xBF58476D1CE4E5B9
  x = (x ^ (x >> 27)) * -7723592293110705685   # 0x94D049BB133111EB
  return x ^ (x >> 31)


def make_depth(torch, batch, ms, tab, images, device, chunk=250):
  """uint16 depth (scale 0.1 mm) of the given global images -> int16 tensor [len(images),H,W] on
  `device`.  Image i: z-buffer of the renders of its GT instances over a tilted background plane,
  + ~N(0, 2 mm) noise and 3 % missing pixels from a counter-based hash of (i, pixel)."""
  from bop_toolkit_b200 import synth
  K = synth.LM_K
  cams = torch.tensor([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], dtype=torch.float64, device=device)
  first = np.concatenate([[0], np.cumsum(tab['n_inst'])])
  out = torch.empty((len(images), H, W), dtype=torch.int16, device=device)
  pix = torch.arange(H * W, device=device, dtype=torch.int64).view(1, H, W)
  ys = torch.arange(H, device=device, dtype=torch.float32).view(1, H, 1)
  xs = torch.arange(W, device=device, dtype=torch.float32).view(1, 1, W)
  inf = float('inf')
  for c0 in range(0, len(images), chunk):
    ims = np.asarray(images[c0:c0 + chunk], dtype=np.int64)
    n = len(ims)
    inst = np.concatenate([np.arange(first[i], first[i + 1]) for i in ims])
    local = np.repeat(np.arange(n), tab['n_inst'][ims])
    poses = torch.from_numpy(batch.pack_poses(tab['R_g'][inst], tab['t_g'][inst])).to(device)
    obj = torch.from_numpy(tab['inst_obj'][inst]).to(device)
    d = batch.render_batch(ms, obj, poses, cams, H, W)
    d = torch.where(d > 0, d, torch.full_like(d, inf))
    img = torch.full((n, H, W), inf, dtype=torch.float32, device=device)
    img.index_reduce_(0, torch.from_numpy(local).to(device), d, 'amin', include_self=True)
    del d
    gid = torch.from_numpy(ims).to(device).view(n, 1, 1)
    hp = _hash_u64(torch, gid * 1000003 + 17)                       # per-image parameters
    f = lambda sh: ((hp >> sh) & 0xFFFF).to(torch.float32) / 65535.0
    bg = (1500.0 + 500.0 * f(0)) + (0.6 * f(16) - 0.3) * (xs - W / 2) + (0.6 * f(32) - 0.3) * (ys - H / 2)
    img = torch.where(torch.isinf(img), bg.expand_as(img), img)
    h = _hash_u64(torch, gid * (H * W) + pix)
    s = ((h & 0xFF) + ((h >> 8) & 0xFF) + ((h >> 16) & 0xFF) + ((h >> 24) & 0xFF)).to(torch.float32)
    img = img + (s - 510.0) * (2.0 / 147.8)                         # Irwin-Hall(4) ~ N(0, 2 mm)
    img = torch.where(((h >> 32) & 0x3FF) < 31, torch.zeros_like(img), img)   # 3 % missing
    v = torch.round(img / DEPTH_SCALE).clamp_(0, 65535).to(torch.int32)
    out[c0:c0 + n] = torch.where(v >= 32768, v - 65536, v).to(torch.int16)
  return out


def records_sha256(ids, errs):
  """SHA-256 of the records ordered by (scene, image, object, estimate, GT)."""
  ids = np.ascontiguousarray(ids, dtype=np.int32)
  errs = np.ascontiguousarray(errs, dtype=np.float64)
  order = np.lexsort((ids[:, 4], ids[:, 3], ids[:, 2], ids[:, 1], ids[:, 0]))
  h = hashlib.sha256()
  h.update(np.ascontiguousarray(ids[order]).tobytes())
  h.update(np.ascontiguousarray(errs[order]).tobytes())
  return h.hexdigest()


def recalls(tab, models, ids, errs):
  """The BOP19 recalls (eval_bop19_pose.py:47-66,147-230): vsd 10 taus x 10 thresholds, mssd and
  mspd 10 thresholds each, through ThresholdSweep on an in-memory context.  -> (dict, list of 120)."""
  from bop_toolkit_b200 import eval_calc_scores as ecs, score_sweep
  T = len(TAUS)
  n_images = tab['n_images']
  scene_ids = [48 + k for k in range(N_SCENES)]
  p = ecs.parse_args([])
  # in-memory context instead of the dataset files (eval_calc_scores.load_context)
  scene_gt = {s: {} for s in scene_ids}
  scene_gt_info = {s: {} for s in scene_ids}
  targets = {s: {} for s in scene_ids}
  first = np.concatenate([[0], np.cumsum(tab['n_inst'])])
  for im in range(n_images):
    s, im_id = scene_of(im)
    objs = (tab['inst_obj'][first[im]:first[im + 1]] + 1).tolist()
    scene_gt[s][im_id] = [{'obj_id': o} for o in objs]
    scene_gt_info[s][im_id] = [{'visib_fract': 1.0} for _ in objs]
    targets[s][im_id] = {o: {'scene_id': s, 'im_id': im_id, 'obj_id': o, 'inst_count': 1} for o in objs}
  ctx = {'key': ('synth', 'test', None, p['targets_filename'], p['datasets_path']),
         'dp_split': {'scene_ids': scene_ids, 'im_size': (W, H)},
         'dp_model': {'obj_ids': list(range(1, N_OBJ + 1))},
         'models_info': {m['obj_id']: {'diameter': m['diameter']} for m in models},
         'targets_org': targets, 'scene_gt': scene_gt, 'scene_gt_info': scene_gt_info}
  order = np.lexsort((ids[:, 4], ids[:, 3], ids[:, 1], ids[:, 0]))
  ids, errs = ids[order], errs[order]
  score = tab['score_of'](ids)
  ths = {'vsd': [round(x, 2) for x in np.arange(0.05, 0.51, 0.05)],
         'mssd': [round(x, 2) for x in np.arange(0.05, 0.51, 0.05)],
         'mspd': [float(x) for x in np.arange(5, 51, 5)]}
  ar, flat = {}, []
  for et, cols in (('vsd', slice(0, T)), ('mssd', slice(T, T + 1)), ('mspd', slice(T + 1, T + 2))):
    recs = {s: [] for s in scene_ids}
    vals = errs[:, cols].tolist()
    for k in range(len(ids)):
      recs[int(ids[k, 0])].append({'im_id': int(ids[k, 1]), 'obj_id': int(ids[k, 2]), 'est_id': int(ids[k, 3]),
                                   'score': float(score[k]), 'errors': {int(ids[k, 4]): vals[k]}})
    sweep = score_sweep.ThresholdSweep(p, et, -1, 'synth', 'test', None, recs, ctx)
    r = []
    for col in range(T if et == 'vsd' else 1):
      r += [sc['recall'] for sc, _ in sweep.run(ths[et], col=col, want_matches=False)]
    ar[et] = float(np.mean(r))
    flat += r
  return ar, flat


def run(torch, dist, device, rank, world, n_images=20000, steps=1, cull=False, seed=7, models=None):
  """-> dict for the bench line (rank 0) / None (other ranks)."""
  from bop_toolkit_b200 import batch, sharding, synth
  models = models if models is not None else build_models()
  ms = batch.ModelSet(device)
  for m in models:
    ms.add(m['obj_id'], m['pts'], m['faces'], m['syms'], m['diameter'], cull=('auto' if cull else False))
  tab = build_table(n_images, seed)
  score_lut = {}
  pairs_per_image = np.bincount(tab['p_im'], minlength=n_images)
  owner = sharding.assign_images(pairs_per_image, world)
  counts = [int(pairs_per_image[owner == r].sum()) for r in range(world)]
  my_images = np.nonzero(owner == rank)[0]
  local_of = np.full(n_images, -1, dtype=np.int64)
  local_of[my_images] = np.arange(len(my_images))
  sel = np.nonzero(owner[tab['p_im']] == rank)[0]
  # the estimates of one GT pose are consecutive in the table (pairs are sorted by instance): groups
  # for bop_vsd_grouped_dev -- the GT pose is projected, sorted and rasterised once per group
  inst_sel = tab['p_inst'][sel]
  cut = np.flatnonzero(np.diff(inst_sel)) + 1
  group_off = np.concatenate([[0], cut, [len(sel)]]).astype(np.int32) if len(sel) else np.zeros(1, np.int32)
  if os.environ.get('BOP_CFG5_NO_GROUPS'):
    group_off = None
  t0 = time.time()
  d16 = make_depth(torch, batch, ms, tab, my_images, device)
  torch.cuda.synchronize()
  gen_s = time.time() - t0
  K = synth.LM_K
  inst = tab['p_inst'][sel]
  pe = torch.from_numpy(batch.pack_poses(tab['R_e'][sel], tab['t_e'][sel])).to(device)
  pg = torch.from_numpy(batch.pack_poses(tab['R_g'][inst], tab['t_g'][inst])).to(device)
  obj = torch.from_numpy(tab['inst_obj'][inst]).to(device)
  im_local = torch.from_numpy(local_of[tab['p_im'][sel]].astype(np.int32)).to(device)
  Kd = torch.from_numpy(np.tile(K.reshape(1, 9), (max(1, len(my_images)), 1))).to(device)
  Kp = torch.from_numpy(np.tile(K.reshape(1, 9), (max(1, len(sel)), 1))).to(device)
  scale = torch.full((max(1, len(my_images)),), DEPTH_SCALE, dtype=torch.float32, device=device)
  sc_id = np.stack(scene_of(tab['p_im'][sel]), axis=1).astype(np.int32).reshape(-1, 2)
  ids = torch.from_numpy(np.stack([sc_id[:, 0], sc_id[:, 1], tab['inst_obj'][inst] + 1,
                                   tab['p_est_id'][sel].astype(np.int32),
                                   tab['inst_gt'][inst].astype(np.int32)], axis=1).astype(np.int32)).to(device)

  def step():
    vsd = batch.vsd_batch(ms, obj, pe, pg, d16, im_local, Kd, DELTA, TAUS, True, 'step', 'bop19', gate=True,
                          depth_scale=scale, group_off=group_off)
    mssd = batch.mssd_batch(ms, obj, pe, pg, gate=True)
    mspd = batch.mspd_batch(ms, obj, pe, pg, Kp)
    errs = torch.cat([vsd, mssd.unsqueeze(1), mspd.unsqueeze(1)], dim=1)
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    out = sharding.gather_records(ids, errs, counts=counts)
    g1.record()
    return out, (g0, g1)

  step()                                   # warm-up (workspace, NCCL channels)
  torch.cuda.synchronize()
  if world > 1:
    dist.barrier()
  torch.cuda.synchronize()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record()
  for _ in range(steps):
    (ids_all, errs_all), gev = step()
  e1.record()
  torch.cuda.synchronize()
  t = torch.tensor([e0.elapsed_time(e1) / steps, gev[0].elapsed_time(gev[1])], dtype=torch.float64, device=device)
  if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
  if rank != 0:
    return None
  ms_step, gather_ms = float(t[0].item()), float(t[1].item())
  ids_np, errs_np = ids_all.cpu().numpy(), errs_all.cpu().numpy()
  # scores travel with the table (every rank has it): look them up by (image, est_id)
  key_all = (tab['p_im'].astype(np.int64) << 8) | tab['p_est_id'].astype(np.int64)
  srt = np.argsort(key_all)

  def score_of(idr):
    im = (idr[:, 0].astype(np.int64) - 48) + idr[:, 1].astype(np.int64) * N_SCENES
    k = (im << 8) | idr[:, 3].astype(np.int64)
    return tab['p_score'][srt[np.searchsorted(key_all[srt], k)]]
  tab['score_of'] = score_of
  t0 = time.time()
  ar, flat = recalls(tab, models, ids_np, errs_np)
  recall_s = time.time() - t0
  n_pairs = int(sum(counts))
  return {
    'workload': 'cfg5: %d images x %.2f instances (21 models of 4902 verts, %d estimates per target, n_top=-1), '
                'vsd+mssd+mspd, images sharded over %d rank(s), one all-gather of the records, 120 BOP19 recall '
                'settings on rank 0' % (n_images, len(tab['inst_im']) / float(n_images), ESTS_PER_TARGET, world),
    'scaling': 'strong', 'n_gpus': world, 'images': int(n_images), 'pairs_total': n_pairs,
    'ms': ms_step, 'pairs_per_s': n_pairs / (ms_step * 1e-3), 'gather_ms': gather_ms,
    'per_rank_pairs': counts, 'lpt_imbalance': max(counts) / (n_pairs / float(world)),
    'record_bytes': int(errs_np.shape[1] * 8 + 24), 'recall_s': recall_s, 'depth_synthesis_s': gen_s,
    'AR': ar, 'AR_mean': float(np.mean(list(ar.values()))),
    'records_sha256': records_sha256(ids_np, errs_np),
    'recalls_sha256': hashlib.sha256(np.asarray(flat, dtype=np.float64).tobytes()).hexdigest(),
    'cull': bool(cull), 'gt_groups': (int(len(group_off)) - 1) if group_off is not None else None}


def main():
  import argparse
  import json
  import torch
  import torch.distributed as dist
  ap = argparse.ArgumentParser()
  ap.add_argument('--images', type=int, default=20000)
  ap.add_argument('--steps', type=int, default=1)
  a = ap.parse_args()
  from bop_toolkit_b200 import sharding
  rank, world, device = sharding.init_from_env()
  out = run(torch, dist, device, rank, world, a.images, a.steps)
  if out is not None:
    print(json.dumps(out))
  if world > 1:
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
  main()

#!/usr/bin/env python
"""Timings of the BASELINE.json configs that are NOT the bench line (bench.py measures configs[1]):

  cfg1  MSSD / MSPD / ADD / ADI, 5k-vertex model, 1,000 perturbed pairs, 4 symmetries
  cfg3b the per-vertex part of cfg3: 1,600 pairs, 8 models, 3 of them with 36 symmetries
  cfg4  ADD / ADI on a 100k-vertex model (64 pairs), MSSD / MSPD with 64 symmetries

Device time per call with CUDA events (inputs resident, best of `--reps`), pairs/s, and for ADI
the FFMA-pipe view: distance evaluations/s and executed fp32 instructions/s against the B200
fp32 issue peak (148 SMs x 128 lanes x SM clock).  A bounded sample of every config is also
timed with the CPU oracle port on one core.  Prints one JSON line per config.

  python tools/bench_configs.py [--only cfg4] [--no-cpu]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(torch, fn, reps):
  fn()
  torch.cuda.synchronize()
  best = 1e30
  for _ in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
  return best, out


def cpu_time(fn, idx):
  t0 = time.perf_counter()
  vals = [fn(i) for i in idx]
  return (time.perf_counter() - t0) / len(idx), vals


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--only', default='')
  ap.add_argument('--reps', type=int, default=5)
  ap.add_argument('--no-cpu', action='store_true')
  args = ap.parse_args()
  import torch
  from bop_toolkit_b200 import batch, misc, synth
  from oracle import pose_error_oracle as po
  dev = torch.device('cuda:0')
  K = synth.LM_K
  col = lambda t: t.reshape(3, 1)
  sm_mhz = 1965.0
  fp32_issue_peak = 148 * 128 * sm_mhz * 1e6      # fp32 lane-instructions / s

  def run(name, models, obj_np, n, metrics, cpu_sample, seed):
    ms = batch.ModelSet(dev)
    for m in models:
      ms.add(m['obj_id'], m['pts'], None, m['syms'], m['diameter'])
    R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=seed, outlier_frac=0.0)
    pe = torch.from_numpy(batch.pack_poses(R_e, t_e)).to(dev)
    pg = torch.from_numpy(batch.pack_poses(R_g, t_g)).to(dev)
    oi = torch.from_numpy(obj_np.astype(np.int32)).to(dev)
    Kd = torch.from_numpy(np.tile(K.reshape(1, 9), (n, 1))).to(dev)
    calls = {'mssd': lambda: batch.mssd_batch(ms, oi, pe, pg), 'mspd': lambda: batch.mspd_batch(ms, oi, pe, pg, Kd),
             'add': lambda: batch.add_batch(ms, oi, pe, pg), 'adi': lambda: batch.adi_batch(ms, oi, pe, pg)}
    ref = {'mssd': lambda i: po.mssd(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), models[obj_np[i]]['pts'], models[obj_np[i]]['syms']),
           'mspd': lambda i: po.mspd(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), K, models[obj_np[i]]['pts'], models[obj_np[i]]['syms']),
           'add': lambda i: po.add(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), models[obj_np[i]]['pts']),
           'adi': lambda i: po.adi(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), models[obj_np[i]]['pts'])}
    line = {'config': name, 'pairs': n, 'models': [dict(n_pts=int(m['pts'].shape[0]), n_syms=len(m['syms'])) for m in models]}
    for met in metrics:
      ms_call, out = timed(torch, calls[met], args.reps)
      r = {'ms': round(ms_call, 4), 'pairs_per_s': round(n / (ms_call * 1e-3), 1)}
      if met == 'adi':
        evals = float(sum(models[o]['pts'].shape[0] ** 2 for o in obj_np))
        r['dist_evals_per_s'] = evals / (ms_call * 1e-3)
        r['fp32_instr_per_s'] = 7.0 * evals / (ms_call * 1e-3)
        r['frac_of_fp32_issue_peak'] = round(r['fp32_instr_per_s'] / fp32_issue_peak, 3)
      if not args.no_cpu:
        idx = list(range(0, n, max(1, n // cpu_sample)))[:cpu_sample]
        s_per, vals = cpu_time(ref[met], idx)
        got = out.cpu().numpy()[idx]
        vals = np.array(vals)
        r['cpu_1core_pairs_per_s'] = round(1.0 / s_per, 2)
        r['cpu_sample'] = len(idx)
        r['max_rel_err_vs_oracle'] = float(np.max(np.abs(got - vals) / np.maximum(np.abs(vals), 1e-9)))
      line[met] = r
    print(json.dumps(line), flush=True)

  def model(obj_id, n_lat, n_lon, diameter, seed, info):
    pts, _ = synth.make_mesh(n_lat, n_lon, diameter, seed=seed)
    info = dict(info, diameter=diameter)
    return dict(obj_id=obj_id, pts=pts, syms=misc.get_symmetry_transformations(info, 0.085), diameter=diameter)

  if args.only in ('', 'cfg1'):
    m = model(1, 50, 100, 200.0, 1, {'symmetries_discrete': synth.discrete_z_symmetries(4)})
    run('cfg1: 4902-vertex model, 1000 pairs, 4 symmetries', [m], np.zeros(1000, np.int64), 1000,
        ['mssd', 'mspd', 'add', 'adi'], 20, 2)
  if args.only in ('', 'cfg3b'):
    models = []
    for k in range(8):
      info = {'symmetries_continuous': [{'axis': [0, 0, 1], 'offset': [0, 0, 0]}]} if k < 3 else {}
      models.append(model(k + 1, 50, 100, 100.0 + 25 * k, 10 + k, info))
    run('cfg3b: 1600 pairs over 8 models (3 with a 36-fold discretised continuous symmetry)', models,
        np.arange(1600) % 8, 1600, ['mssd', 'mspd'], 16, 3)
  if args.only in ('', 'cfg4'):
    pts, _ = synth.make_dense_mesh(100000, 250.0, seed=5)
    info = {'diameter': 250.0, 'symmetries_discrete': synth.discrete_z_symmetries(64)}
    m = dict(obj_id=1, pts=pts, syms=misc.get_symmetry_transformations(info, 0.01), diameter=250.0)
    run('cfg4: %d-vertex model, 64 pairs, 64 symmetries' % pts.shape[0], [m], np.zeros(64, np.int64), 64,
        ['add', 'adi', 'mssd', 'mspd'], 2, 4)


if __name__ == '__main__':
  main()

#!/usr/bin/env python
"""Where a resident bench step spends its time: device time (CUDA events) and host time (wall clock
until the call returns, without synchronising) per batched call.  usage: step_breakdown.py [pairs]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from bop_toolkit_b200 import _lib, batch, synth

P = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
dev = torch.device('cuda', 0)
_lib.require_cuda()
models = bench.build_models()
ms = batch.ModelSet(dev)
for m in models:
  ms.add(m['obj_id'], m['pts'], m['faces'], m['syms'], m['diameter'])
R_e, t_e, R_g, t_g, obj = bench.build_pairs(P, seed=0)
pe, pg = torch.from_numpy(batch.pack_poses(R_e, t_e)).to(dev), torch.from_numpy(batch.pack_poses(R_g, t_g)).to(dev)
d_obj = torch.from_numpy(obj).to(dev)
K = synth.LM_K
d_K = torch.from_numpy(np.tile(K.reshape(1, 9), (P, 1))).to(dev)
d_im = torch.arange(P, dtype=torch.int32, device=dev)
diam = torch.tensor([models[o]['diameter'] for o in obj], device=dev)
d16 = bench.make_depth_gpu(batch, torch, ms, d_obj, pg, torch.from_numpy(t_g.copy()).to(dev), diam, K, dev, seed=1)
d_f32 = torch.empty((P, bench.H, bench.W), dtype=torch.float32, device=dev)
for c0 in range(0, P, 1000):
  v = d16[c0:c0 + 1000].to(torch.int32)
  d_f32[c0:c0 + 1000] = torch.where(v < 0, v + 65536, v).to(torch.float32) * bench.DEPTH_SCALE
del d16
calls = {
  'vsd': lambda: batch.vsd_batch(ms, d_obj, pe, pg, d_f32, d_im, d_K, bench.DELTA, bench.TAUS, True, 'step', 'bop19', gate=True),
  'mssd': lambda: batch.mssd_batch(ms, d_obj, pe, pg, gate=True),
  'mspd': lambda: batch.mspd_batch(ms, d_obj, pe, pg, d_K),
}
for name, fn in calls.items():
  for _ in range(3):
    fn()
  torch.cuda.synchronize()
  dev_ms, host_ms = [], []
  for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(); fn(); e1.record(); t1 = time.perf_counter()
    torch.cuda.synchronize()
    dev_ms.append(e0.elapsed_time(e1)); host_ms.append(1e3 * (t1 - t0))
  print('%-5s device %.3f ms  host %.3f ms' % (name, np.median(dev_ms), np.median(host_ms)))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
  for fn in calls.values():
    fn()
e1.record(); torch.cuda.synchronize()
print('all three, 10 steps back to back: %.3f ms per step' % (e0.elapsed_time(e1) / 10))

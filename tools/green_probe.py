#!/usr/bin/env python
"""Probe: do CUDA green contexts (an SM partition of one GPU) apply to kernels this library launches
through the runtime API on a stream created inside a green context?  Times the resident VSD call on
the whole GPU, on each partition alone and on both partitions at once.

  python tools/green_probe.py [--sms 88] [--pairs 2000]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--sms', type=int, default=88)
  ap.add_argument('--pairs', type=int, default=2000)
  args = ap.parse_args()
  import torch
  from bop_toolkit_b200 import batch, sm_partition, synth
  dev = torch.device('cuda', 0)
  torch.zeros(1, device=dev)
  part = sm_partition.SmPartition(0, args.sms)
  print(json.dumps({'sms': part.sm_counts}))
  K = synth.LM_K
  pts, faces = synth.make_mesh(50, 100, 200.0, seed=1)
  ms = batch.ModelSet(dev)
  ms.add(1, pts, faces, None, 200.0)
  n = args.pairs
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=2, z_range=(500.0, 1200.0), outlier_frac=0.05)
  pe = torch.from_numpy(batch.pack_poses(R_e, t_e)).to(dev)
  pg = torch.from_numpy(batch.pack_poses(R_g, t_g)).to(dev)
  depth = torch.full((1, 480, 640), 900.0, dtype=torch.float32, device=dev)
  im = torch.zeros(n, dtype=torch.int32, device=dev)
  Kd = torch.from_numpy(K.reshape(1, 9).copy()).to(dev)
  taus = [0.05 * (k + 1) for k in range(10)]
  run = lambda: batch.vsd_batch(ms, 0, pe, pg, depth, im, Kd, 15, taus, True, 'step')
  want = run().cpu()
  wss = [batch.Workspace(dev), batch.Workspace(dev)]

  def timed(streams, reps=5):
    for rep in range(reps + 2):
      if rep == 2:
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True)
        e0.record()
      outs = []
      for k, s in enumerate(streams):
        if s is None:
          outs.append(run())
          continue
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s), batch.private_workspace(wss[k]):
          outs.append(run())
      for s in streams:
        if s is not None:
          torch.cuda.current_stream().wait_stream(s)
    e1 = torch.cuda.Event(enable_timing=True)
    e1.record()
    torch.cuda.synchronize()
    for o in outs:
      assert torch.equal(o.cpu(), want)
    return e0.elapsed_time(e1) / reps

  out = {'whole_gpu_ms': timed([None])}
  out['partition_a_ms'] = timed([part.streams[0]])
  out['partition_b_ms'] = timed([part.streams[1]])
  out['both_ms'] = timed([part.streams[0], part.streams[1]])
  print(json.dumps(out))


if __name__ == '__main__':
  main()

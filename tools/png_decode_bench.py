#!/usr/bin/env python
"""Depth-PNG decode: OpenCV (libpng + zlib) on the host threads vs the GPU inflate + unfilter
kernels, on synthetic 640x480 16-bit depth images written the way OpenCV writes them (Sub filter,
RLE strategy: what this repo's synthetic datasets hold) and the way pypng wrote the BOP datasets
(filter None, zlib level 6: inout.save_depth, bop_toolkit_lib/inout.py:49-63).

  python tools/png_decode_bench.py --images 512
Prints one JSON line per encoder.
"""
import argparse
import ctypes
import json
import os
import shutil
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--images', type=int, default=512)
  ap.add_argument('--reps', type=int, default=3)
  args = ap.parse_args()
  import cv2
  import torch
  from bop_toolkit_b200 import _lib, batch, eval_calc_errors as ece
  from test_png_depth import write_png16
  H, W = 480, 640
  rng = np.random.default_rng(0)
  yy, xx = np.mgrid[:H, :W]
  root = tempfile.mkdtemp(prefix='png_bench_')
  ims = []
  for i in range(16):   # distinct images, cycled
    plane = 1200 + rng.uniform(-0.4, 0.4) * xx + rng.uniform(-0.4, 0.4) * yy
    cx, cy, r = rng.uniform(150, 490), rng.uniform(100, 380), rng.uniform(40, 120)
    blob = np.where((xx - cx) ** 2 + (yy - cy) ** 2 < r * r, -300 + 0.002 * ((xx - cx) ** 2 + (yy - cy) ** 2), 0)
    d = plane + blob + rng.normal(0, 2, (H, W))
    d[rng.random((H, W)) < 0.03] = 0
    ims.append(np.clip(d, 0, 65535).astype(np.uint16))
  try:
    for enc in ('opencv', 'pypng_style'):
      paths = []
      for i in range(args.images):
        p = os.path.join(root, '%s_%06d.png' % (enc, i))
        if i < 16:
          if enc == 'opencv':
            cv2.imwrite(p, ims[i])
          else:
            write_png16(p, ims[i], [0], 6)
        else:
          shutil.copyfile(paths[i % 16], p)
        paths.append(p)
      comp = sum(os.path.getsize(p) for p in paths)
      threads = ece._io_threads()
      scales = [1.0] * len(paths)
      best = {}
      for rep in range(args.reps):
        t0 = time.time()
        depth, _ = ece.load_depth_stack(paths, scales, H, W)
        best['host_opencv_s'] = min(best.get('host_opencv_s', 1e9), time.time() - t0)
        ref = depth.numpy().view(np.uint16).copy()
        torch.cuda.synchronize()
        t0 = time.time()
        staged = batch.png_stage(paths, threads)
        t1 = time.time()
        _lib.lib().bop_profile_enable(1)
        out = batch.decode_depth_staged(staged, H, W, 'cuda:0')
        torch.cuda.synchronize()
        t2 = time.time()
        ms, cnt = ctypes.c_double(), ctypes.c_int()
        _lib.lib().bop_profile_sum_ms(8, 1, ctypes.byref(ms), ctypes.byref(cnt))
        _lib.lib().bop_profile_enable(0)
        best['gpu_stage_s'] = min(best.get('gpu_stage_s', 1e9), t1 - t0)
        best['gpu_h2d_decode_s'] = min(best.get('gpu_h2d_decode_s', 1e9), t2 - t1)
        best['gpu_total_s'] = min(best.get('gpu_total_s', 1e9), t2 - t0)
        best['gpu_kernels_ms'] = min(best.get('gpu_kernels_ms', 1e9), ms.value)
        assert np.array_equal(out.cpu().numpy().view(np.uint16), ref)
      n = len(paths)
      print(json.dumps({
        'encoder': enc, 'images': n, 'host_threads': threads, 'compressed_mb': round(comp / 1e6, 1),
        'raw_mb': round(n * H * W * 2 / 1e6, 1), 'seconds': {k: round(v, 4) for k, v in best.items()},
        'images_per_s': {'host_opencv': round(n / best['host_opencv_s']),
                         'gpu_total': round(n / best['gpu_total_s']),
                         'gpu_kernels': round(n / (best['gpu_kernels_ms'] * 1e-3))},
        'identical_to_opencv': True}))
  finally:
    shutil.rmtree(root, ignore_errors=True)


if __name__ == '__main__':
  main()

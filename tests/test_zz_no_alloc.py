"""bop_vsd_from_depth_static_dev -- the allocation-free variant of bop_vsd_from_depth_dev whose
scratch is a static __device__ array inside the library (include/bop_b200.h; no counterpart in the
reference) -- called through the C ABI (ctypes) and
compared with the existing entry point on the same inputs, on two calls in a row with different
batches (the array keeps the previous call's contents, so a missing clear would show).

Tolerance: the one tests/test_gpu_parity.py uses for the K2 pass -- step cost bit-identical
(integer counts), tlinear 1e-12 (float64 sums added in a different order).

This file is named to be collected LAST (pytest -x must not lose other tests to
a failure here); it passes alone in a fresh process and after the rest of the suite.
"""
import ctypes

import numpy as np
import pytest


def _case(n, n_im, H, W, seed):
  """Pre-rendered est / GT depth (blobs over a zero background) + a sensor image per image."""
  rng = np.random.default_rng(seed)
  de = np.zeros((n, H, W), np.float32)
  dg = np.zeros((n, H, W), np.float32)
  ys, xs = np.mgrid[0:H, 0:W]
  for i in range(n):
    cy, cx, r = rng.uniform(0.3, 0.7) * H, rng.uniform(0.3, 0.7) * W, rng.uniform(0.1, 0.25) * H
    z = rng.uniform(500.0, 1200.0)
    m_g = (ys - cy) ** 2 + (xs - cx) ** 2 < r * r
    m_e = (ys - cy - rng.uniform(-6, 6)) ** 2 + (xs - cx - rng.uniform(-6, 6)) ** 2 < (r * rng.uniform(0.9, 1.1)) ** 2
    dg[i][m_g] = (z + 0.2 * (xs - cx))[m_g]
    de[i][m_e] = (z + rng.uniform(-25, 25) + 0.2 * (xs - cx))[m_e]
  im = rng.integers(0, n_im, size=n).astype(np.int32)
  dt = rng.uniform(400.0, 1400.0, size=(n_im, 1, 1)).astype(np.float32) + \
    rng.normal(0, 20.0, size=(n_im, H, W)).astype(np.float32)
  dt[:, ::5, ::3] = 0.0                                     # missing depth
  K = np.tile(np.array([572.4114, 0, 325.2611, 0, 573.57043, 242.04899, 0, 0, 1.0]), (n_im, 1))
  diam = rng.uniform(80.0, 300.0, size=n)
  return de, dg, dt, im, K, diam


def _call(lib, _lib, torch, static, case, taus, cost, H, W):
  """One call through the C ABI; device buffers are torch tensors (the ENTRY POINT allocates nothing;
  the caller owns inputs and outputs).  -> (errors [n, T], counts [n, 2 + T])."""
  from bop_toolkit_b200 import batch
  de, dg, dt, im, K, diam = case
  n, T = de.shape[0], len(taus)
  dev = torch.device('cuda', 0)
  t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
  d_de, d_dg, d_dt, d_im, d_K, d_diam = t(de), t(dg), t(dt), t(im), t(K), t(diam)
  err = torch.full((n, T), -7.0, dtype=torch.float64, device=dev)
  cnt = torch.full((n, 2 + T), -7, dtype=torch.int64, device=dev)
  prm = batch.make_vsd_params(H, W, 15.0, taus, True, cost, 'bop19', False, _lib.DEPTH_F32)
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
  if static:
    rc = lib.bop_vsd_from_depth_static_dev(d_dt.data_ptr(), None, d_im.data_ptr(), d_K.data_ptr(), d_de.data_ptr(),
                                           d_dg.data_ptr(), d_diam.data_ptr(), n, ctypes.byref(prm),
                                           err.data_ptr(), cnt.data_ptr(), stream)
  else:
    wsb = lib.bop_vsd_workspace_bytes(n, H, W, 1)
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    rc = lib.bop_vsd_from_depth_dev(d_dt.data_ptr(), None, d_im.data_ptr(), d_K.data_ptr(), d_de.data_ptr(),
                                    d_dg.data_ptr(), d_diam.data_ptr(), n, ctypes.byref(prm), err.data_ptr(),
                                    cnt.data_ptr(), ws.data_ptr(), wsb, stream)
  _lib.check(rc)
  torch.cuda.synchronize()
  return err.cpu().numpy(), cnt.cpu().numpy()


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_workspace_entry_point():
  import torch
  from bop_toolkit_b200 import _lib
  _lib.require_cuda()
  lib = _lib.lib()
  H, W = 120, 160
  taus = [0.05 * (k + 1) for k in range(10)]
  big, small = _case(48, 5, H, W, seed=1), _case(17, 3, H, W, seed=2)
  # two calls in a row on the static scratch: a large batch, then a smaller, different one
  got = [_call(lib, _lib, torch, True, big, taus, 'step', H, W),
         _call(lib, _lib, torch, True, small, taus, 'step', H, W)]
  want = [_call(lib, _lib, torch, False, big, taus, 'step', H, W),
          _call(lib, _lib, torch, False, small, taus, 'step', H, W)]
  for (e1, c1), (e0, c0) in zip(got, want):
    assert np.array_equal(c1, c0)
    assert np.array_equal(e1, e0)
    assert (c0[:, 1] > 0).sum() >= len(c0) // 2 and np.isfinite(e0).all()      # the cases do have intersections
  assert not np.array_equal(got[0][0][:17], got[1][0])                         # the two batches differ
  # tlinear right after step on the same array (its sums must start at zero), twice, and step again
  for case in (big, small):
    e1, c1 = _call(lib, _lib, torch, True, case, taus[:4], 'tlinear', H, W)
    e0, c0 = _call(lib, _lib, torch, False, case, taus[:4], 'tlinear', H, W)
    assert np.array_equal(c1[:, :2], c0[:, :2])
    assert np.abs(e1 - e0).max() <= 1e-12
  e1, c1 = _call(lib, _lib, torch, True, big, taus, 'step', H, W)
  assert np.array_equal(e1, want[0][0]) and np.array_equal(c1, want[0][1])


@pytest.mark.gpu
def test_static_scratch_entry_point_limit_and_no_allocation():
  import torch
  from bop_toolkit_b200 import _lib
  _lib.require_cuda()
  lib = _lib.lib()
  # beyond the stated maximum: rejected before anything is touched, with a clear message
  rc = lib.bop_vsd_from_depth_static_dev(None, None, None, None, None, None, None, _lib.STATIC_MAX_PAIRS + 1,
                                         None, None, None, None)
  assert rc != 0
  msg = lib.bop_last_error().decode()
  assert 'BOP_STATIC_MAX_PAIRS' in msg and '4096' in msg
  assert lib.bop_vsd_from_depth_static_dev(None, None, None, None, None, None, None, 0, None, None, None, None) == 0
  # steady state: device memory in use does not move across calls (everything is warm: kernels
  # loaded, symbol address looked up, the test's own tensors allocated by torch's caching allocator)
  H, W = 120, 160
  taus = [0.1, 0.3]
  case = _case(8, 2, H, W, seed=3)
  for cost in ('step', 'tlinear'):
    _call(lib, _lib, torch, True, case, taus, cost, H, W)
  torch.cuda.synchronize()
  free0, _ = torch.cuda.mem_get_info()
  outs = [_call(lib, _lib, torch, True, case, taus, cost, H, W) for cost in ('step', 'tlinear', 'step')]
  torch.cuda.synchronize()
  free1, _ = torch.cuda.mem_get_info()
  assert free1 == free0, (free0, free1)
  assert np.array_equal(outs[0][0], outs[2][0])
  # the full maximum fits
  n = _lib.STATIC_MAX_PAIRS
  dev = torch.device('cuda', 0)
  de = torch.zeros((n, 8, 16), dtype=torch.float32, device=dev)
  de[:, 2:6, 4:12] = 700.0
  dt = torch.full((1, 8, 16), 705.0, dtype=torch.float32, device=dev)
  im = torch.zeros(n, dtype=torch.int32, device=dev)
  K = torch.tensor([[100.0, 0, 8, 0, 100.0, 4, 0, 0, 1]], dtype=torch.float64, device=dev)
  diam = torch.full((n,), 100.0, dtype=torch.float64, device=dev)
  err = torch.empty((n, 2), dtype=torch.float64, device=dev)
  from bop_toolkit_b200 import batch
  prm = batch.make_vsd_params(8, 16, 15.0, taus, True, 'step', 'bop19', False, _lib.DEPTH_F32)
  _lib.check(lib.bop_vsd_from_depth_static_dev(dt.data_ptr(), None, im.data_ptr(), K.data_ptr(), de.data_ptr(),
                                               de.data_ptr(), diam.data_ptr(), n, ctypes.byref(prm), err.data_ptr(),
                                               None, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
  torch.cuda.synchronize()
  assert (err == 0.0).all()          # est == gt, everything visible: cost 0 at every tau

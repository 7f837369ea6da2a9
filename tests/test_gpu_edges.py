"""GPU edge cases and size-independent properties of the CUDA path (through the C ABI).

  * empty batches for every batched entry point;
  * close-up poses (image-filling object, large triangles) against the oracle rasteriser and
    under different kernel tunings -- tile size, the small/large triangle split and the
    deferred-triangle queue (overflow included) must not change a single bit of any depth image
    or VSD count (raster.cuh: the depth of a pixel depends on the triangle and the pixel only);
  * objects behind / straddling the camera plane (outside the parity domain: triangles touching
    Z <= 1e-6 are dropped, same rule as the oracle), fully outside the image, all-zero depth;
  * the tau limit of one launch (BOP_MAX_TAUS) and chunking above it;
  * at BASELINE.json's full size (10,000 pairs): results do not depend on batch composition
    (shuffled batch == shuffled results, bit for bit), VSD(gt, gt) is 0 wherever something is
    visible and 1 otherwise, MSSD / MSPD of a GT pose composed with a model symmetry is ~0.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

DEFAULT_TUNING = (20480, 1024, 1, 256)


@pytest.fixture(scope='module')
def B():
  import torch
  from bop_toolkit_b200 import _lib, batch
  assert torch.cuda.is_available(), 'GPU tests need a CUDA device'
  _lib.require_cuda()
  return batch


def _tune(*t):
  from bop_toolkit_b200 import _lib
  _lib.check(_lib.lib().bop_vsd_set_tuning(*t))


def _cams(K, n):
  return np.tile([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], (n, 1))


def test_empty_batches(B):
  from bop_toolkit_b200 import synth
  pts, faces = synth.make_mesh(10, 20, 100.0, seed=1)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 100.0)
  z = np.zeros((0, 12))
  K = synth.LM_K
  assert B.mssd_batch(ms, 0, z, z).shape == (0,)
  assert B.mspd_batch(ms, 0, z, z, K).shape == (0,)
  assert B.add_batch(ms, 0, z, z).shape == (0,)
  assert B.adi_batch(ms, 0, z, z).shape == (0,)
  re, te = B.re_te_batch(z, z)
  assert re.shape == (0,) and te.shape == (0,)
  depth = np.ones((1, 480, 640), np.float32)
  assert B.vsd_batch(ms, 0, z, z, depth, None, K.reshape(1, 9), 15, [0.1, 0.2]).shape == (0, 2)
  assert B.render_batch(ms, 0, z, np.zeros((0, 4)), 480, 640).shape == (0, 480, 640)


def _closeup_case(n=10, seed=3):
  from bop_toolkit_b200 import synth
  pts, faces = synth.make_mesh(24, 48, 220.0, seed=seed)     # coarse mesh: big triangles up close
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=seed + 1, z_range=(140.0, 320.0), outlier_frac=0.0)
  t_g[0] = [0.0, 0.0, 125.0]          # surface ~15 mm from the camera plane: fills the whole image
  t_e[0] = [3.0, -2.0, 128.0]
  t_g[1] = [250.0, 150.0, 160.0]      # mostly outside
  t_e[1] = [245.0, 155.0, 165.0]
  return pts, faces, R_e, t_e, R_g, t_g


def test_closeup_render_vs_oracle(B, oracle_built):
  from bop_toolkit_b200 import synth
  from oracle import renderer_oracle
  from test_gpu_parity import _render_compare
  pts, faces, _, _, R_g, t_g = _closeup_case()
  K = synth.LM_K
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 220.0)
  d = B.render_batch(ms, 0, B.pack_poses(R_g, t_g), _cams(K, len(R_g)), 480, 640).cpu().numpy()
  tot_sil = tot_mis = tot_far = n_bad = n_px = 0
  for i in range(len(R_g)):
    ref = renderer_oracle.render_depth(pts, faces, R_g[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2], 480, 640, 8)
    sil, mis, far, rel, bad = _render_compare(d[i], ref)
    tot_sil += sil; tot_mis += mis; tot_far += far
    n_bad += int(bad.sum()); n_px += rel.size
  assert (d[0] > 0).mean() > 0.95                      # the first pose really fills the image
  assert tot_far == 0
  assert tot_mis <= 0.0005 * tot_sil + 2, (tot_mis, tot_sil)
  assert n_bad <= 1e-4 * n_px + 2, (n_bad, n_px)
  print('close-up render: %d covered px, %d silhouette px, %d coverage mismatches, %d depth outliers'
        % (n_px, tot_sil, tot_mis, n_bad))


def test_results_do_not_depend_on_tuning(B):
  """Tile height, CTA shape, the small/large triangle threshold and the deferred queue
  (small_tri_px = 1 sends every triangle through it and overflows its 2048 slots) are
  scheduling choices: depth images and VSD counts stay bit-identical."""
  from bop_toolkit_b200 import synth
  K = synth.LM_K
  pts, faces, R_e, t_e, R_g, t_g = _closeup_case()
  fine_pts, fine_faces = synth.make_mesh(50, 100, 200.0, seed=9)
  R2e, t2e, R2g, t2g = synth.make_pose_pairs(12, seed=10, z_range=(300.0, 900.0), outlier_frac=0.1)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 220.0)
  ms.add(2, fine_pts, fine_faces, None, 200.0)
  obj = np.array([0] * len(R_e) + [1] * len(R2e), np.int32)
  pe = B.pack_poses(np.concatenate([R_e, R2e]), np.concatenate([t_e, t2e]))
  pg = B.pack_poses(np.concatenate([R_g, R2g]), np.concatenate([t_g, t2g]))
  n = len(obj)
  rng = np.random.default_rng(0)
  taus = [0.05, 0.2, 0.5]

  def run():
    dg = B.render_batch(ms, obj, pg, _cams(K, n), 480, 640)
    depth = dg.clone()
    depth[depth == 0] = 1500.0
    out = [dg.cpu().numpy()]
    for cost in ('step', 'tlinear'):
      e, c = B.vsd_batch(ms, obj, pe, pg, depth, np.arange(n, dtype=np.int32), np.tile(K.reshape(1, 9), (n, 1)),
                         15, taus, True, cost, return_counts=True)
      out += [e.cpu().numpy(), c.cpu().numpy()]
    st = B.mask_overlap_batch(ms, obj, pe, pg, K, 480, 640)
    out += [st[k].cpu().numpy() for k in sorted(st)]
    return out

  try:
    base = run()
    assert (base[0][0] > 0).mean() > 0.95
    for tuning in ((4096, 256, 2, 4), (18432, 1024, 1, 1), (4096, 128, 4, 1 << 20), (9216, 512, 2, 64)):
      _tune(*tuning)
      for a, b in zip(base, run()):
        if a.dtype.kind == 'f' and a.ndim == 2 and a.shape[1] == len(taus):
          # tlinear sums float64 costs in a tiling-dependent order: 1e-11 (step costs are exact)
          assert np.allclose(a, b, rtol=0, atol=1e-11, equal_nan=True), tuning
        else:
          assert np.array_equal(a, b, equal_nan=True), tuning
  finally:
    _tune(*DEFAULT_TUNING)


def test_behind_camera_outside_image_and_zero_depth(B, oracle_built):
  from bop_toolkit_b200 import synth
  from oracle import pose_error_oracle as po, renderer_oracle
  K = synth.LM_K
  pts, faces = synth.make_mesh(24, 48, 200.0, seed=5)
  ms = B.ModelSet()
  ms.add(1, pts, faces, [{'R': np.eye(3), 't': np.zeros((3, 1))}], 200.0, cull='auto')   # opt-in culling
  assert ms.info[1]['cull'] != 0                     # culled model: the inside view must survive
  R = np.tile(np.eye(3), (5, 1, 1))
  t_g = np.array([[0, 0, 60.0],        # camera INSIDE the object: triangles straddle Z = 0
                  [0, 0, -500.0],      # entirely behind the camera
                  [9000.0, 0, 600.0],  # far outside the image
                  [0, 0, 600.0],       # normal pose
                  [0, 0, 600.0]])      # normal pose, all-zero depth image (pair 4)
  t_e = t_g + np.array([2.0, 1.0, 5.0])
  n = len(t_g)
  d = B.render_batch(ms, 0, B.pack_poses(R, t_g), _cams(K, n), 480, 640).cpu().numpy()
  assert np.isfinite(d).all() and (d >= 0).all()
  assert not d[1].any() and not d[2].any() and d[3].any()
  for i in (0, 3):
    ref = renderer_oracle.render_depth(pts, faces, R[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2], 480, 640, 8)
    cov = (d[i] > 0) != (ref > 0)
    assert cov.sum() <= 0.002 * max(1, (ref > 0).sum()) + 2, (i, int(cov.sum()))
  depth = np.full((n, 480, 640), 1500.0, np.float32)
  depth[4] = 0.0
  taus = [0.05, 0.5]
  err = B.vsd_batch(ms, 0, B.pack_poses(R, t_e), B.pack_poses(R, t_g), depth, None,
                    np.tile(K.reshape(1, 9), (n, 1)), 15, taus).cpu().numpy()
  assert np.isfinite(err).all()
  assert np.array_equal(err[1], [1.0, 1.0]) and np.array_equal(err[2], [1.0, 1.0])   # empty union
  # all-zero depth: bop19 visibility keeps every model pixel (visibility.py:44-47)
  ren = renderer_oracle.RendererOracle(640, 480)
  ren.add_object(1, pts=pts, faces=faces)
  for i in (3, 4):
    ref = po.vsd(R[i], t_e[i].reshape(3, 1), R[i], t_g[i].reshape(3, 1), depth[i], K, 15, taus, True, 200.0,
                 ren, 1, 'step')
    assert np.all(np.abs(err[i] - np.array(ref)) <= 1e-3), (i, err[i], ref)
  # point errors are plain arithmetic behind the camera too
  mssd = B.mssd_batch(ms, 0, B.pack_poses(R, t_e), B.pack_poses(R, t_g)).cpu().numpy()
  assert np.allclose(mssd, np.linalg.norm([2.0, 1.0, 5.0]), rtol=1e-6)


def test_tau_limit_and_chunking(B):
  from bop_toolkit_b200 import _lib, synth
  K = synth.LM_K
  pts, faces = synth.make_mesh(24, 48, 200.0, seed=6)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 200.0)
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(6, seed=7, z_range=(400.0, 900.0), outlier_frac=0.0)
  pe, pg = B.pack_poses(R_e, t_e), B.pack_poses(R_g, t_g)
  dg = B.render_batch(ms, 0, pg, _cams(K, 6), 480, 640)
  depth = dg.clone()
  depth[depth == 0] = 1400.0
  taus = [0.01 * (k + 1) for k in range(2 * _lib.BOP_MAX_TAUS + 3)]     # three launches
  Ks = np.tile(K.reshape(1, 9), (6, 1))
  full = B.vsd_batch(ms, 0, pe, pg, depth, None, Ks, 15, taus).cpu().numpy()
  assert full.shape == (6, len(taus))
  for k in (0, _lib.BOP_MAX_TAUS - 1, _lib.BOP_MAX_TAUS, len(taus) - 1):
    one = B.vsd_batch(ms, 0, pe, pg, depth, None, Ks, 15, [taus[k]]).cpu().numpy()
    assert np.array_equal(one[:, 0], full[:, k])
  assert np.all(np.diff(full, axis=1) <= 0)          # step cost is monotone in tau
  prm = B.make_vsd_params(480, 640, 15, taus[:_lib.BOP_MAX_TAUS])
  prm.n_taus = _lib.BOP_MAX_TAUS + 1                  # one launch refuses more than its limit
  import ctypes
  import torch
  out = torch.empty((6, _lib.BOP_MAX_TAUS + 1), dtype=torch.float64, device='cuda')
  wsb = _lib.lib().bop_vsd_workspace_bytes(6, 480, 640, ms.max_geom)
  ws = torch.empty(wsb, dtype=torch.uint8, device='cuda')
  oi = torch.zeros(6, dtype=torch.int32, device='cuda')
  ii = torch.arange(6, dtype=torch.int32, device='cuda')
  rc = _lib.lib().bop_vsd_dev(ms.table.data_ptr(), oi.data_ptr(), torch.from_numpy(pe).cuda().data_ptr(),
                              torch.from_numpy(pg).cuda().data_ptr(), depth.data_ptr(), None, ii.data_ptr(),
                              torch.from_numpy(Ks).cuda().data_ptr(), None, 6, ms.max_geom, ctypes.byref(prm),
                              out.data_ptr(), None, ws.data_ptr(), wsb, None)
  assert rc != 0 and _lib.lib().bop_last_error()


def test_full_size_properties(B):
  """BASELINE.json configs[1] size: 10,000 pairs, 8 models of 4,902 vertices."""
  import torch
  from bop_toolkit_b200 import misc, synth
  K = synth.LM_K
  P, n_im = 10000, 48
  ms = B.ModelSet()
  syms = []
  for k in range(8):
    pts, faces = synth.make_mesh(50, 100, 120.0 + 20 * k, seed=20 + k)
    info = {'diameter': 120.0 + 20 * k}
    if k < 3:
      info['symmetries_discrete'] = synth.discrete_z_symmetries(4)
    s = misc.get_symmetry_transformations(info, 0.01)
    syms.append(s)
    ms.add(k + 1, pts, faces, s, info['diameter'])
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(P, seed=77, z_range=(450.0, 1400.0), outlier_frac=0.05)
  obj = (np.arange(P) % 8).astype(np.int32)
  im = (np.arange(P) % n_im).astype(np.int32)
  pe, pg = B.pack_poses(R_e, t_e), B.pack_poses(R_g, t_g)
  Ks = np.tile(K.reshape(1, 9), (n_im, 1))
  # test images: the GT render of the first pair that uses the image over a far background
  first = np.array([np.flatnonzero(im == i)[0] for i in range(n_im)])
  depth = B.render_batch(ms, obj[first], pg[first], _cams(K, n_im), 480, 640)
  depth[depth == 0] = 1800.0
  taus = [0.05 * (k + 1) for k in range(10)]

  def run(order):
    v, c = B.vsd_batch(ms, obj[order], pe[order], pg[order], depth, im[order], Ks, 15, taus, True, 'step',
                       gate=True, return_counts=True)
    return (v, c, B.mssd_batch(ms, obj[order], pe[order], pg[order], gate=True),
            B.mspd_batch(ms, obj[order], pe[order], pg[order], K),
            B.add_batch(ms, obj[order], pe[order], pg[order]))

  ident = np.arange(P)
  perm = np.random.default_rng(5).permutation(P)
  a, b = run(ident), run(perm)
  from bop_toolkit_b200 import _lib
  try:                                               # float32 decision path == all-float64 pixel pass
    _lib.check(_lib.lib().bop_vsd_set_exact(1))
    ex = run(ident)
  finally:
    _lib.check(_lib.lib().bop_vsd_set_exact(0))
  assert torch.equal(a[0], ex[0]) and torch.equal(a[1], ex[1])
  pt = torch.from_numpy(perm).cuda()
  for x, y in zip(a, b):
    assert torch.equal(x[pt], y)                       # batch composition does not matter
  # VSD of the GT pose against itself: 0 where anything is visible, 1 where the union is empty
  v, c = B.vsd_batch(ms, obj, pg, pg, depth, im, Ks, 15, taus, True, 'step', return_counts=True)
  v, c = v.cpu().numpy(), c.cpu().numpy()
  vis = c[:, 0] > 0
  assert vis.sum() > 0.5 * P
  assert np.all(v[vis] == 0.0) and np.all(v[~vis] == 1.0) and np.all(c[:, 2:] == 0)
  assert np.array_equal(c[:, 0], c[:, 1])              # union == intersection
  # GT composed with a model symmetry: MSSD / MSPD see the same pose (float rounding only)
  R_s, t_s = R_g.copy(), t_g.copy()
  for i in range(P):
    s = syms[obj[i]][i % len(syms[obj[i]])]
    R_s[i] = R_g[i] @ s['R']
    t_s[i] = R_g[i] @ s['t'][:, 0] + t_g[i]
  ps = B.pack_poses(R_s, t_s)
  assert float(B.mssd_batch(ms, obj, ps, pg).max()) < 1e-3       # mm, at |t| ~ 1e3 mm
  assert float(B.mspd_batch(ms, obj, ps, pg, K).max()) < 1e-3    # px
  # ADD >= MSSD-without-symmetry is not a theorem, but max >= mean is: max over points (MSSD of
  # the identity symmetry alone) bounds the mean (ADD) from above for the 5 asymmetric models
  add, mssd = a[4].cpu().numpy(), a[2].cpu().numpy()
  asym = (obj >= 3) & np.isfinite(mssd)
  assert np.all(mssd[asym] >= add[asym] * (1 - 1e-6))


def test_float32_decision_path_equals_exact(B):
  """The step-cost pixel pass decides comparisons in float32 when they are far from their
  thresholds and falls back to the reference's float64 arithmetic otherwise.  Counts must be
  bit-identical to the all-float64 run (bop_vsd_set_exact) -- on a normal workload and on
  adversarial inputs built to sit ON the thresholds: depth_test at exactly delta behind the GT
  surface (+- a few float32 ulps) and taus equal to exact per-pixel costs."""
  import torch
  from bop_toolkit_b200 import _lib, synth
  lib = _lib.lib()
  K = synth.LM_K
  pts, faces = synth.make_mesh(50, 100, 180.0, seed=31)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 180.0)
  n = 48
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=32, z_range=(400.0, 1200.0), outlier_frac=0.05)
  pe, pg = B.pack_poses(R_e, t_e), B.pack_poses(R_g, t_g)
  cams = _cams(K, n)
  de = B.render_batch(ms, 0, pe, cams, 480, 640).cpu().numpy()
  dg = B.render_batch(ms, 0, pg, cams, 480, 640).cpu().numpy()
  ys, xs = np.mgrid[0:480, 0:640]
  s = np.sqrt(((xs - K[0, 2]) / K[0, 0]) ** 2 + ((ys - K[1, 2]) / K[1, 1]) ** 2 + 1.0)
  delta = 15.0
  rng = np.random.default_rng(33)
  # depth_test: delta behind the GT surface along the ray, nudged by -3..3 float32 ulps; elsewhere far
  dt = np.where(dg > 0, dg + (delta / s).astype(np.float32), np.float32(1500.0)).astype(np.float32)
  steps = rng.integers(-3, 4, size=dt.shape)
  for k in range(1, 4):
    dt = np.where(steps >= k, np.nextafter(dt, np.float32(np.inf)), dt)
    dt = np.where(steps <= -k, np.nextafter(dt, np.float32(-np.inf)), dt)
  dt[:, ::7, ::5] = 0.0                                     # missing depth
  # taus that ARE per-pixel costs (float64, as the reference computes them)
  both = (dg > 0) & (de > 0)
  Dg, De = dg.astype(np.float64) * s, de.astype(np.float64) * s
  costs = np.abs(Dg - De)[both] / 180.0
  costs = costs[(costs > 1e-4) & (costs < 0.6)]
  taus = sorted(float(c) for c in rng.choice(costs, 10, replace=False))
  Ks = np.tile(K.reshape(1, 9), (n, 1))
  im = np.arange(n, dtype=np.int32)
  results = {}
  noisy = np.where(dg > 0, dg + rng.normal(0, 3.0, dg.shape).astype(np.float32), 1500.0).astype(np.float32)
  try:
    for exact in (1, 0, 2):      # 1: all float64; 0: float32 pre-test + tau histogram; 2: float32 pre-test + per-tau counters
      _lib.check(lib.bop_vsd_set_exact(exact))
      out = []
      unsorted = [taus[k] for k in (3, 0, 7, 1)]              # not ascending: per-tau counters whatever the mode
      for depth, tt in ((dt, taus), (dt, [0.05 * (k + 1) for k in range(10)]), (noisy, taus[:3]), (dt, unsorted),
                        (dt, [taus[2], taus[2], taus[5]])):    # duplicate taus
        for mode in ('bop19', 'bop18'):
          for norm in (True, False):
            t2 = tt if norm else [t * 180.0 for t in tt]
            e, c = B.vsd_batch(ms, 0, pe, pg, depth, im, Ks, delta, t2, norm, 'step', mode, return_counts=True)
            out += [e.cpu().numpy(), c.cpu().numpy()]
      results[exact] = out
  finally:
    _lib.check(lib.bop_vsd_set_exact(0))
  for a, b, c in zip(results[1], results[0], results[2]):
    assert np.array_equal(a, b) and np.array_equal(a, c)
  assert results[1][1][:, 1].sum() > 1000                    # the cases do have intersections
  # and the exact run equals the stand-alone float64 pixel pass on the same renders
  e, c = B.vsd_from_depth_batch(torch.from_numpy(de).cuda(), torch.from_numpy(dg).cuda(), dt, im, Ks, 180.0,
                                delta, taus, True, 'step', 'bop19', return_counts=True)
  assert np.array_equal(c.cpu().numpy(), results[0][1])


def test_dense_model_takes_the_global_vertex_scratch(B, oracle_built):
  """Models above 12,288 vertices do not fit the assemble kernel's shared memory: their projected
  vertices go through the global scratch.  Same renders (vs the oracle) and fused VSD == render +
  stand-alone pixel pass, also when a small and a dense model share one batch."""
  from bop_toolkit_b200 import synth
  from oracle import renderer_oracle
  from test_gpu_parity import _render_compare
  K = synth.LM_K
  dense_pts, dense_faces = synth.make_mesh(100, 200, 210.0, seed=41)
  small_pts, small_faces = synth.make_mesh(24, 48, 150.0, seed=42)
  assert dense_pts.shape[0] > 12288
  ms = B.ModelSet()
  ms.add(1, dense_pts, dense_faces, None, 210.0)
  ms.add(2, small_pts, small_faces, None, 150.0)
  n = 6
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=43, z_range=(350.0, 900.0), outlier_frac=0.0)
  obj = np.array([0, 1, 0, 1, 0, 0], np.int32)
  pe, pg = B.pack_poses(R_e, t_e), B.pack_poses(R_g, t_g)
  cams = _cams(K, n)
  dg = B.render_batch(ms, obj, pg, cams, 480, 640)
  de = B.render_batch(ms, obj, pe, cams, 480, 640)
  tot_sil = tot_mis = tot_far = 0
  for i in (0, 1, 2):
    pts, faces = (dense_pts, dense_faces) if obj[i] == 0 else (small_pts, small_faces)
    ref = renderer_oracle.render_depth(pts, faces, R_g[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2], 480, 640, 8)
    sil, mis, far, rel, bad = _render_compare(dg[i].cpu().numpy(), ref)
    tot_sil += sil; tot_mis += mis; tot_far += far
    assert int(bad.sum()) <= 1e-4 * rel.size + 2
  assert tot_far == 0 and tot_mis <= 0.0005 * tot_sil + 2, (tot_mis, tot_sil)
  depth = dg.clone()
  depth[depth == 0] = 1600.0
  Ks = np.tile(K.reshape(1, 9), (n, 1))
  im = np.arange(n, dtype=np.int32)
  taus = [0.05, 0.2, 0.5]
  diam = np.where(obj == 0, 210.0, 150.0)
  a, ca = B.vsd_batch(ms, obj, pe, pg, depth, im, Ks, 15, taus, True, 'step', return_counts=True)
  b, cb = B.vsd_from_depth_batch(de, dg, depth, im, Ks, diam, 15, taus, True, 'step', return_counts=True)
  assert np.array_equal(ca.cpu().numpy(), cb.cpu().numpy())
  assert np.array_equal(a.cpu().numpy(), b.cpu().numpy())


def test_wide_image(B, oracle_built):
  """4000-pixel-wide images: the per-column ray tables grow with the width and the z-tile shrinks
  to fit shared memory; renders and VSD must not care."""
  from bop_toolkit_b200 import synth
  from oracle import renderer_oracle
  from test_gpu_parity import _render_compare
  H, W = 200, 4000
  K = np.array([[2500.0, 0, 2000.3], [0, 2500.0, 99.6], [0, 0, 1]])
  pts, faces = synth.make_mesh(24, 48, 200.0, seed=51)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 200.0)
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(3, seed=52, K=K, width=W, height=H, z_range=(1500.0, 3000.0),
                                             outlier_frac=0.0)
  t_g[0] = [-900.0, 0.0, 1500.0]; t_e[0] = [-897.0, 2.0, 1510.0]     # far left
  t_g[1] = [950.0, 10.0, 1600.0]; t_e[1] = [955.0, 8.0, 1590.0]      # far right
  pe, pg = B.pack_poses(R_e, t_e), B.pack_poses(R_g, t_g)
  cams = _cams(K, 3)
  dg = B.render_batch(ms, 0, pg, cams, H, W)
  de = B.render_batch(ms, 0, pe, cams, H, W)
  for i in range(3):
    ref = renderer_oracle.render_depth(pts, faces, R_g[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2], H, W, 8)
    sil, mis, far, rel, bad = _render_compare(dg[i].cpu().numpy(), ref)
    assert sil > 0 and far == 0 and mis <= 0.0005 * sil + 2 and int(bad.sum()) <= 1e-4 * rel.size + 2
  depth = dg.clone()
  depth[depth == 0] = 4000.0
  Ks = np.tile(K.reshape(1, 9), (3, 1))
  im = np.arange(3, dtype=np.int32)
  taus = [0.05, 0.2, 0.5]
  for cost in ('step', 'tlinear'):
    a, ca = B.vsd_batch(ms, 0, pe, pg, depth, im, Ks, 15, taus, True, cost, return_counts=True)
    b, cb = B.vsd_from_depth_batch(de, dg, depth, im, Ks, 200.0, 15, taus, True, cost, return_counts=True)
    assert np.array_equal(ca.cpu().numpy()[:, :2], cb.cpu().numpy()[:, :2])
    assert np.allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=0, atol=1e-11)
  st = B.mask_overlap_batch(ms, 0, pe, pg, K, H, W)
  assert int(st['counts'][:, 0].min()) > 0


def test_point_errors_float32_fold_vs_float64(B):
  """MSSD / MSPD / ADD / PROJ evaluate float32-exact model points in float32 AFTER folding the rigid
  parts in float64 (pose_points.cu); bop_points_set_exact(1) forces the reference's float64 per vertex.
  The two must agree to 1e-6 relative (bar: 1e-5) on close estimates, gross outliers, a skewed K,
  estimates that ARE a symmetry of the GT (errors ~ 0) and poses partly behind the camera."""
  from bop_toolkit_b200 import _lib, synth, misc
  lib = _lib.lib()
  K = synth.LM_K.copy()
  K[0, 1] = 3.5                                   # skew
  pts, faces = synth.make_mesh(40, 80, 220.0, seed=41)
  syms = misc.get_symmetry_transformations({'diameter': 220.0, 'symmetries_discrete': synth.discrete_z_symmetries(6)}, 0.01)
  ms = B.ModelSet()
  ms.add(1, pts, None, syms, 220.0)
  n = 600
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=42, z_range=(350.0, 1500.0), outlier_frac=0.1)
  # estimates that are exactly a symmetry of the GT
  for i in range(0, 40):
    s_ = syms[i % len(syms)]
    R_e[i] = R_g[i] @ s_['R']
    t_e[i] = (R_g[i] @ s_['t']).reshape(3) + t_g[i]
  # tiny perturbations (sub-micron errors) and poses straddling the camera plane
  t_e[40:80] = t_g[40:80] + 1e-4
  R_e[40:80] = R_g[40:80]
  t_g[80:90, 2] = 60.0
  t_e[80:90, 2] = 58.0
  pe, pg = B.pack_poses(R_e, t_e), B.pack_poses(R_g, t_g)
  Ks = np.tile(K.reshape(1, 9), (n, 1))
  out = {}
  try:
    for exact in (1, 0):
      _lib.check(lib.bop_points_set_exact(exact))
      out[exact] = (B.mssd_batch(ms, 0, pe, pg).cpu().numpy(), B.mspd_batch(ms, 0, pe, pg, Ks).cpu().numpy(),
                    B.add_batch(ms, 0, pe, pg).cpu().numpy(), B.proj_batch(ms, 0, pe, pg, Ks).cpu().numpy())
  finally:
    _lib.check(lib.bop_points_set_exact(0))
  for k, name in enumerate(('mssd', 'mspd', 'add', 'proj')):
    a, b = out[1][k], out[0][k]
    fin = np.isfinite(a)
    assert np.array_equal(fin, np.isfinite(b)), name
    rel = np.abs(a[fin] - b[fin]) / np.maximum(np.abs(a[fin]), 1e-6)   # 1e-6 mm / px floor for the ~0 cases
    assert rel.max() <= 1e-6, (name, float(rel.max()), int(np.argmax(rel)))
  assert out[1][0][:40].max() < 1e-9 and out[0][0][:40].max() < 1e-3      # symmetric poses: ~0 either way


def test_triangle_lists_are_sorted_by_band_and_class(B):
  """The geometry pass writes every (pair, pose) triangle list sorted by (row band, candidate-
  rectangle class, largest first) -- what keeps the lanes of a raster warp converged.  Checked
  through the debug hook on a batch with small and multi-band ROIs."""
  from bop_toolkit_b200 import synth
  K = synth.LM_K
  pts, faces = synth.make_mesh(50, 100, 260.0, seed=51)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 260.0)
  n = 12
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=52, z_range=(380.0, 1400.0), outlier_frac=0.0)
  pe, pg = B.pack_poses(R_e, t_e), B.pack_poses(R_g, t_g)
  depth = np.full((n, 480, 640), 1000.0, np.float32)
  B.vsd_batch(ms, 0, pe, pg, depth, np.arange(n, dtype=np.int32), np.tile(K.reshape(1, 9), (n, 1)), 15.0, [0.2])
  n_multi = 0
  for pair in range(n):
    for pose in (0, 1):
      xy = B.debug_triangle_list(ms, n, 480, 640, pair, pose).astype(np.int64)
      assert len(xy) > 0
      minx, maxx = xy[:, [0, 2, 4]].min(1), xy[:, [0, 2, 4]].max(1)
      miny, maxy = xy[:, [1, 3, 5]].min(1), xy[:, [1, 3, 5]].max(1)
      w = np.minimum(639, maxx >> 8) - np.maximum(0, (minx + 255) >> 8) + 1
      h = np.minimum(479, maxy >> 8) - np.maximum(0, (miny + 255) >> 8) + 1
      assert (w > 0).all() and (h > 0).all()                 # only triangles that can draw
      key = np.minimum(h, 15) * 16 + np.minimum(w, 15)
      # the list is a sequence of segments (bands, then the shared segment); inside a segment the
      # class key never increases
      rises = np.flatnonzero(np.diff(key) > 0)
      assert len(rises) <= 16, len(rises)
      n_multi += int(len(rises) > 0)
      cost = 0
      k32 = len(w) // 32 * 32
      W_, H_ = w[:k32].reshape(-1, 32), h[:k32].reshape(-1, 32)
      for r in range(int(H_.max())):
        cost += int(np.where(H_ > r, W_, 0).max(1).sum())
      assert (W_ * H_).sum() >= 0.8 * 32 * cost           # >= 80 % of the lane-slots of the pixel loops are used
  assert n_multi > 0                                       # some ROI needed more than one band


def test_calls_on_two_streams_share_the_workspace_in_order(B):
  """Every batched call on a device uses ONE scratch buffer (batch.Workspace; the library keeps
  its work lists and counters there).  A call issued on another stream than the previous one
  must therefore wait for it: two different batches launched back to back on two streams give
  the results each gives alone."""
  import torch
  from bop_toolkit_b200 import synth
  K = synth.LM_K
  pts, faces = synth.make_mesh(40, 80, 200.0, seed=51)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 200.0)
  n = 1500
  cases = []
  for seed in (52, 53):
    R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=seed, z_range=(400.0, 1200.0), outlier_frac=0.05)
    cases.append((B.pack_poses(R_e, t_e), B.pack_poses(R_g, t_g)))
  depth = torch.full((1, 480, 640), 900.0, dtype=torch.float32, device=ms.device)
  im = np.zeros(n, np.int32)
  Ks = K.reshape(1, 9)
  taus = [0.05 * (k + 1) for k in range(10)]
  run = lambda pe, pg: B.vsd_batch(ms, 0, pe, pg, depth, im, Ks, 15, taus, True, 'step', return_counts=True)
  want = [tuple(x.cpu().numpy() for x in run(*c)) for c in cases]
  torch.cuda.synchronize()
  streams = [torch.cuda.Stream(ms.device), torch.cuda.Stream(ms.device)]
  for rep in range(3):
    got = []
    for s_, c in zip(streams, cases):
      s_.wait_stream(torch.cuda.current_stream())
      with torch.cuda.stream(s_):
        got.append(run(*c))
    torch.cuda.synchronize()
    for (e1, c1), (e0, c0) in zip(got, want):
      assert np.array_equal(c1.cpu().numpy(), c0) and np.array_equal(e1.cpu().numpy(), e0, equal_nan=True)
  assert not np.array_equal(want[0][1], want[1][1])


def test_split_streams_on_an_sm_partition_are_bit_identical(B):
  """bop_vsd_split_dev (vsd_batch(raster_stream=...)): geometry pass on one stream, raster + pixel
  pass on another, ordered by events inside the call -- on two plain streams and on the two
  streams of an SM partition (sm_partition.SmPartition, CUDA green contexts), pipelined over
  several batches with alternating private workspaces the way bench.py's end-to-end leg does,
  resident float32 depth and host-mapped uint16 depth: results equal bop_vsd_dev's bit for bit."""
  import torch
  from bop_toolkit_b200 import synth, sm_partition
  K = synth.LM_K
  pts, faces = synth.make_mesh(40, 80, 200.0, seed=61)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 200.0)
  dev = ms.device
  n = 1200
  cases = []
  for seed in (62, 63, 64):
    R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=seed, z_range=(400.0, 1200.0), outlier_frac=0.05)
    cases.append((B.pack_poses(R_e, t_e), B.pack_poses(R_g, t_g)))
  rng = np.random.default_rng(65)
  d16 = (rng.uniform(7000, 11000, size=(4, 480, 640))).astype(np.uint16)        # 0.1 mm units
  d16[:, ::7, ::5] = 0
  h_d16 = torch.from_numpy(d16.view(np.int16)).pin_memory()
  scale = np.full(4, 0.1, np.float32)
  d_f32 = torch.from_numpy(d16.astype(np.float32) * np.float32(0.1)).to(dev)
  im = (np.arange(n) % 4).astype(np.int32)
  Ks = np.tile(K.reshape(1, 9), (4, 1))
  taus = [0.05 * (k + 1) for k in range(10)]

  def run(c, host, **kw):
    if host:
      return B.vsd_batch(ms, 0, c[0], c[1], h_d16, im, Ks, 15, taus, True, 'step', return_counts=True,
                         depth_scale=scale, zero_copy=True, **kw)
    return B.vsd_batch(ms, 0, c[0], c[1], d_f32, im, Ks, 15, taus, True, 'step', return_counts=True, **kw)

  want = {(i, host): tuple(x.cpu().numpy() for x in run(c, host)) for i, c in enumerate(cases) for host in (False, True)}
  assert np.array_equal(want[(0, False)][1], want[(0, True)][1])     # u16 * 0.1f == the float32 image
  torch.cuda.synchronize()
  part = sm_partition.SmPartition(dev.index if dev.index is not None else 0, 96)
  assert sum(part.sm_counts) == torch.cuda.get_device_properties(dev).multi_processor_count
  assert min(part.sm_counts) >= 8
  plain = [torch.cuda.Stream(dev), torch.cuda.Stream(dev)]
  for s_r, s_c in (plain, part.streams):
    wss = [B.Workspace(dev), B.Workspace(dev)]
    done = [None, None]
    got = []
    order = [(i % 3, bool(i & 1)) for i in range(7)]
    for s_ in (s_r, s_c):
      s_.wait_stream(torch.cuda.current_stream())
    for j, (i, host) in enumerate(order):
      k = j % 2
      with torch.cuda.stream(s_c), B.private_workspace(wss[k]):
        if done[k] is not None:
          s_c.wait_event(done[k])
        e, c = run(cases[i], host, raster_stream=s_r)
      with torch.cuda.stream(s_r):
        done[k] = torch.cuda.Event()
        done[k].record(s_r)
      got.append((e, c))
    torch.cuda.synchronize()
    for (i, host), (e, c) in zip(order, got):
      assert np.array_equal(c.cpu().numpy(), want[(i, host)][1]), (i, host)
      assert np.array_equal(e.cpu().numpy(), want[(i, host)][0], equal_nan=True), (i, host)
  # more taus than one launch takes: the tau chunks reuse the workspace, ordered inside vsd_batch
  many = [0.01 * (k + 1) for k in range(45)]
  e0 = B.vsd_batch(ms, 0, cases[0][0], cases[0][1], d_f32, im, Ks, 15, many, True, 'step').cpu().numpy()
  with torch.cuda.stream(part.streams[1]), B.private_workspace(B.Workspace(dev)):
    e1 = B.vsd_batch(ms, 0, cases[0][0], cases[0][1], d_f32, im, Ks, 15, many, True, 'step', raster_stream=part.streams[0])
  torch.cuda.synchronize()
  assert np.array_equal(e1.cpu().numpy(), e0, equal_nan=True)

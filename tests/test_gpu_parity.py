"""GPU parity tests (run on the B200 box): the CUDA path, called through the C ABI, against the
committed golden vectors (real reference outputs) and against the CPU oracle on seeded inputs.

Tolerances (BASELINE.json north_star):
  re/te/add/adi/mssd/mspd/proj : 1e-5 relative (+ tiny absolute floor for exact-zero cases)
  rendered depth               : 1e-3 relative per pixel; coverage mismatches only at silhouette
                                 edges, <= 0.5 % of silhouette pixels (+2 px slack)
  VSD                          : 1e-3 absolute per tau; bit-exact given identical depth images
"""
import ctypes

import numpy as np
import pytest

from conftest import syms_from_array

pytestmark = pytest.mark.gpu

RTOL = 1e-5
ATOL = 1e-9


@pytest.fixture(scope='module')
def B():
  import torch
  from bop_toolkit_b200 import _lib, batch
  assert torch.cuda.is_available(), 'GPU tests need a CUDA device'
  _lib.require_cuda()
  return batch


def _poses(B, R, t):
  return B.pack_poses(R, t)


def _close(a, b, rtol=RTOL, atol=ATOL):
  a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
  fin = np.isfinite(b)
  assert np.array_equal(np.isfinite(a), fin)
  assert np.array_equal(a[~fin], b[~fin])
  err = np.abs(a[fin] - b[fin])
  tol = atol + rtol * np.abs(b[fin])
  assert np.all(err <= tol), 'max rel err %.3e' % float((err / np.maximum(np.abs(b[fin]), 1e-300)).max())


# ------------------------------------------------------------------ golden vectors
@pytest.mark.parametrize('mname', ['f32', 'f64'])
def test_point_errors_vs_reference_golden(B, golden, mname):
  g = golden
  pts = g['pts_small'] if mname == 'f32' else g['pts_f64']
  pe, pg = _poses(B, g['R_e'], g['t_e']), _poses(B, g['R_g'], g['t_g'])
  for si in range(4):
    key = 'mssd_%s_%d' % (mname, si)
    if key not in g:
      continue
    ms = B.ModelSet()
    ms.add(1, pts, None, g['sym_%d' % si], 150.0)
    _close(B.mssd_batch(ms, 0, pe, pg).cpu().numpy(), g[key])
    _close(B.mspd_batch(ms, 0, pe, pg, g['K']).cpu().numpy(), g['mspd_%s_%d' % (mname, si)])
    if si == 1:
      _close(B.mspd_batch(ms, 0, pe, pg, g['K_skew']).cpu().numpy(), g['mspd_skew_' + mname])
  ms = B.ModelSet()
  ms.add(1, pts, None, None, 150.0)
  _close(B.add_batch(ms, 0, pe, pg).cpu().numpy(), g['add_' + mname])
  _close(B.proj_batch(ms, 0, pe, pg, g['K']).cpu().numpy(), g['proj_' + mname])
  # ADI: float32 search in the model frame; 1e-5 relative with an absolute floor of 1e-6 mm.
  _close(B.adi_batch(ms, 0, pe, pg).cpu().numpy(), g['adi_' + mname], atol=1e-6)


def test_re_te_vs_reference_golden(B, golden):
  g = golden
  re, te = B.re_te_batch(_poses(B, g['R_e'], g['t_e']), _poses(B, g['R_g'], g['t_g']))
  _close(re.cpu().numpy(), g['re'], atol=1e-6)   # acos is ill-conditioned at 0 deg
  _close(te.cpu().numpy(), g['te'])


def test_facade_signatures_vs_reference_golden(B, golden):
  from bop_toolkit_b200 import pose_error as pe
  g = golden
  syms = syms_from_array(g['sym_1'])
  for i in (0, 1, 5):
    a = (g['R_e'][i], g['t_e'][i].reshape(3, 1), g['R_g'][i], g['t_g'][i].reshape(3, 1))
    _close([pe.mssd(*a, g['pts_small'], syms)], [g['mssd_f32_1'][i]])
    _close([pe.mspd(*a, g['K'], g['pts_small'], syms)], [g['mspd_f32_1'][i]])
    _close([pe.add(*a, g['pts_small'])], [g['add_f32'][i]])
    _close([pe.adi(*a, g['pts_small'])], [g['adi_f32'][i]], atol=1e-6)
    _close([pe.proj(*a, g['K'], g['pts_small'])], [g['proj_f32'][i]])
    _close([pe.re(a[0], a[2])], [g['re'][i]], atol=1e-6)
    _close([pe.te(a[1], a[3])], [g['te'][i]])
    assert isinstance(pe.mssd(*a, g['pts_small'], syms), float)
  with pytest.raises(AssertionError):
    pe.re(np.eye(4), np.eye(3))
  with pytest.raises(AssertionError):
    pe.te(np.zeros(4), np.zeros(3))


def test_vsd_vs_reference_golden(B, golden):
  """Reference pose_error.vsd (run on the oracle renders) vs the fused CUDA kernel."""
  from bop_toolkit_b200 import pose_error as pe
  from bop_toolkit_b200.renderer import create_renderer
  g = golden
  H, W = g['vsd_depth_test'].shape[1:]
  ren = create_renderer(W, H, 'cuda', mode='depth')
  ren.add_object(7, pts=g['vsd_pts'], faces=g['vsd_faces'], diameter=180.0)
  taus = list(g['vsd_taus'])
  n = g['vsd_R_e'].shape[0]
  worst = 0.0
  for i in range(n):
    a = (g['vsd_R_e'][i], g['vsd_t_e'][i].reshape(3, 1), g['vsd_R_g'][i], g['vsd_t_g'][i].reshape(3, 1),
         g['vsd_depth_test'][i], g['vsd_K'], 15, taus)
    e = pe.vsd(*a, True, 180.0, ren, 7, 'step')
    assert isinstance(e, list) and len(e) == len(taus) and isinstance(e[0], float)
    worst = max(worst, np.abs(np.array(e) - g['vsd_step'][i]).max())
    e = pe.vsd(*a, True, 180.0, ren, 7, 'tlinear')
    worst = max(worst, np.abs(np.array(e) - g['vsd_tlinear'][i]).max())
    e = pe.vsd(*a[:6], 15, [5.0, 20.0], False, 180.0, ren, 7, 'step')
    worst = max(worst, np.abs(np.array(e) - g['vsd_unnorm'][i]).max())
  # Small 160x120 images: one silhouette pixel is ~1e-3 of the union, so allow 3e-3 here; the
  # full-size test below holds the 1e-3 bar.
  assert worst <= 3e-3, worst
  with pytest.raises(ValueError):
    pe.vsd(*a, True, 180.0, ren, 7, 'nope')


def test_k2_bit_exact_vs_reference_golden(B, golden):
  """Given the SAME rendered depth images, counts and errors are bit-identical to numpy's."""
  g = golden
  taus = list(g['vsd_taus'])
  n = g['vsd_R_e'].shape[0]
  for cost, key in (('step', 'vsd_step'), ('tlinear', 'vsd_tlinear')):
    err = B.vsd_from_depth_batch(g['vsd_render_est'], g['vsd_render_gt'], g['vsd_depth_test'],
                                 np.arange(n, dtype=np.int32), np.tile(g['vsd_K'].reshape(1, 9), (n, 1)),
                                 180.0, 15, taus, True, cost).cpu().numpy()
    if cost == 'step':
      assert np.array_equal(err, g[key])
    else:
      assert np.allclose(err, g[key], rtol=0, atol=1e-13)
  err = B.vsd_from_depth_batch(g['vsd_render_est'], g['vsd_render_gt'], g['vsd_depth_test'],
                               None, np.tile(g['vsd_K'].reshape(1, 9), (n, 1)), 180.0, 15,
                               [5.0, 20.0], False, 'step').cpu().numpy()
  assert np.array_equal(err, g['vsd_unnorm'])
  # 'bop18' visibility (visibility.py:29-32; pose_error.vsd itself always asks for 'bop19'): against the
  # oracle, whose masks are pinned to the reference's visibility module in tests/test_oracle_golden.py
  from oracle import pose_error_oracle as po
  err, cnt = B.vsd_from_depth_batch(g['vsd_render_est'], g['vsd_render_gt'], g['vsd_depth_test'],
                                    np.arange(n, dtype=np.int32), np.tile(g['vsd_K'].reshape(1, 9), (n, 1)),
                                    180.0, 15, taus, True, 'step', visib_mode='bop18', return_counts=True)
  err, cnt = err.cpu().numpy(), cnt.cpu().numpy()
  differs = False
  for i in range(n):
    ref, rc = po.vsd_from_depth(g['vsd_render_est'][i], g['vsd_render_gt'][i], g['vsd_depth_test'][i], g['vsd_K'],
                                15, taus, True, 180.0, 'step', visib_mode='bop18', return_counts=True)
    assert int(cnt[i, 0]) == rc[0] and int(cnt[i, 1]) == rc[1]
    assert np.array_equal(err[i], np.array(ref))
    differs = differs or not np.array_equal(err[i], g['vsd_step'][i])
  assert differs            # the mode matters on this data (missing depth pixels)


# ------------------------------------------------------------------ oracle, seeded
def test_point_errors_vs_oracle_cfg1(B):
  """BASELINE configs[0] shape: ~5k-vertex model, 1,000 perturbed pairs, 4 symmetries."""
  from bop_toolkit_b200 import synth
  from oracle import pose_error_oracle as po
  pts, faces = synth.make_mesh(seed=1)
  info = {'diameter': 200.0, 'symmetries_discrete': synth.discrete_z_symmetries(4)}
  syms = po.get_symmetry_transformations(info, 0.01)
  ms = B.ModelSet()
  ms.add(5, pts, faces, syms, 200.0)
  n = 1000
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=2)
  pe, pg = _poses(B, R_e, t_e), _poses(B, R_g, t_g)
  mssd = B.mssd_batch(ms, 0, pe, pg).cpu().numpy()
  mspd = B.mspd_batch(ms, 0, pe, pg, synth.LM_K).cpu().numpy()
  add = B.add_batch(ms, 0, pe, pg).cpu().numpy()
  adi = B.adi_batch(ms, 0, pe, pg).cpu().numpy()
  mssd_g = B.mssd_batch(ms, 0, pe, pg, gate=True).cpu().numpy()
  sel = np.arange(0, n, 25)   # the oracle is slow: check every 25th pair
  col = lambda t: t.reshape(3, 1)
  _close(mssd[sel], [po.mssd(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), pts, syms) for i in sel])
  _close(mspd[sel], [po.mspd(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), synth.LM_K, pts, syms) for i in sel])
  _close(add[sel], [po.add(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), pts) for i in sel])
  _close(adi[sel], [po.adi(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), pts) for i in sel], atol=1e-6)
  # driver gate: inf exactly where ||t_e - t_g|| >= diameter (eval_calc_errors.py:305-322)
  far = np.linalg.norm(t_e - t_g, axis=1) >= 200.0
  assert far.any() and not far.all()
  assert np.all(np.isinf(mssd_g[far])) and np.array_equal(mssd_g[~far], mssd[~far])
  assert np.all(adi <= add * (1 + 1e-6) + 1e-9)


def test_many_symmetries_and_mixed_models(B):
  """36-fold discretised continuous symmetry (configs[2]) + heterogeneous objects per batch."""
  from bop_toolkit_b200 import synth
  from oracle import pose_error_oracle as po
  infos = [{'diameter': 120.0},
           {'diameter': 150.0, 'symmetries_continuous': [{'axis': [0, 0, 1], 'offset': [0, 0, 0]}]},
           {'diameter': 90.0, 'symmetries_discrete': synth.discrete_z_symmetries(2)}]
  ms = B.ModelSet()
  models = []
  for k, info in enumerate(infos):
    pts, faces = synth.make_mesh(14 + 3 * k, 30, info['diameter'], seed=10 + k)
    syms = po.get_symmetry_transformations(info, 0.085)
    ms.add(10 + k, pts, faces, syms, info['diameter'])
    models.append((pts, syms))
  assert ms.max_syms == 36
  n = 90
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=3)
  obj = (np.arange(n) % 3).astype(np.int32)
  pe, pg = _poses(B, R_e, t_e), _poses(B, R_g, t_g)
  mssd = B.mssd_batch(ms, obj, pe, pg).cpu().numpy()
  mspd = B.mspd_batch(ms, obj, pe, pg, np.tile(synth.LM_K.reshape(1, 9), (n, 1))).cpu().numpy()
  col = lambda t: t.reshape(3, 1)
  _close(mssd, [po.mssd(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), *models[obj[i]]) for i in range(n)])
  _close(mspd, [po.mspd(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), synth.LM_K, *models[obj[i]]) for i in range(n)])


def _render_compare(d_gpu, d_ref):
  """Returns (n_silhouette, n_coverage_mismatch, n_interior_bad, max_rel_err)."""
  cg, cr = d_gpu > 0, d_ref > 0
  both = cg & cr
  rel = np.abs(d_gpu[both] - d_ref[both]) / d_ref[both]
  # silhouette = covered pixels with an uncovered 4-neighbour
  pad = np.pad(cr, 1)
  sil = cr & ~(pad[:-2, 1:-1] & pad[2:, 1:-1] & pad[1:-1, :-2] & pad[1:-1, 2:])
  mism = cg != cr
  near = np.pad(sil, 1)
  near = near[:-2, 1:-1] | near[2:, 1:-1] | near[1:-1, :-2] | near[1:-1, 2:] | near[1:-1, 1:-1]
  # depth outliers (> 1e-3) are tolerated only where two surfaces meet within a pixel of an
  # occlusion edge; count those that are NOT next to any depth discontinuity of the reference
  bad = np.zeros_like(cr)
  bad[both] = rel > 1e-3
  return int(sil.sum()), int(mism.sum()), int((mism & ~near).sum()), rel, bad


def test_render_vs_oracle_full_size(B, oracle_built):
  from bop_toolkit_b200 import synth
  from bop_toolkit_b200.renderer import create_renderer
  from oracle import renderer_oracle
  pts, faces = synth.make_mesh(seed=4, diameter=220.0)
  ren = create_renderer(640, 480, 'cuda', mode='depth')
  ren.add_object(3, pts=pts, faces=faces)
  K = synth.LM_K
  n = 12
  _, _, R_g, t_g = synth.make_pose_pairs(n, seed=5, z_range=(350.0, 1400.0))
  t_g[0] = [0.0, 0.0, 330.0]            # very close: ROI needs several z-tile bands
  t_g[1] = [400.0, 250.0, 700.0]        # partly outside the image
  t_g[2] = [5000.0, 0.0, 700.0]         # fully outside: empty image
  tot_sil = tot_mis = tot_far = n_bad = n_px = 0
  snap_mis = snap_far = 0
  for i in range(n):
    d = ren.render_object(3, R_g[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2])['depth']
    assert d.dtype == np.float32 and d.shape == (480, 640)
    ref = renderer_oracle.render_depth(pts, faces, R_g[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2], 480, 640)
    sil, mis, far, rel, bad = _render_compare(d, ref)
    tot_sil += sil; tot_mis += mis; tot_far += far
    n_bad += int(bad.sum()); n_px += rel.size
    # oracle with the same GL-style 1/256 px vertex snapping as the CUDA rasteriser
    ref8 = renderer_oracle.render_depth(pts, faces, R_g[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2], 480, 640, 8)
    _, mis8, far8, _, _ = _render_compare(d, ref8)
    snap_mis += mis8; snap_far += far8
    if i == 2:
      assert not (d > 0).any() and not (ref > 0).any()
  assert tot_far == 0 and snap_far == 0             # zero interior coverage mismatches
  # vs exact positions: snapping moves an edge by <= 1/512 px -> <= 0.5 % of silhouette pixels
  assert tot_mis <= 0.005 * tot_sil + 2, (tot_mis, tot_sil)
  # vs the snapped oracle only float32-vs-float64 projection rounding is left
  assert snap_mis <= 0.0005 * tot_sil + 2, (snap_mis, tot_sil)
  # per-pixel depth: 1e-3 relative, except pixels straddling a self-occlusion edge
  assert n_bad <= 1e-4 * n_px + 2, (n_bad, n_px)
  print('render parity: silhouette px %d, coverage mismatches exact-oracle %d, snapped-oracle %d, '
        'depth outliers %d of %d' % (tot_sil, tot_mis, snap_mis, n_bad, n_px))


def test_render_known_answer_top_left(B):
  """Pixel-aligned square: exact coverage under the top-left rule, exact depth, both windings,
  and the warp-cooperative large-triangle path (two triangles of 200 px)."""
  from bop_toolkit_b200.renderer import create_renderer
  pts = np.array([[-10, -10, 0], [10, -10, 0], [10, 10, 0], [-10, 10, 0]], dtype=np.float64)
  for faces in ([[0, 1, 2], [0, 2, 3]], [[2, 1, 0], [3, 2, 0]]):
    ren = create_renderer(40, 40, 'cuda', mode='depth')
    ren.add_object(1, pts=pts, faces=np.array(faces))
    d = ren.render_object(1, np.eye(3), np.array([0, 0, 100.0]), 100.0, 100.0, 20.5, 20.5)['depth']
    cov = d > 0
    rows, cols = np.nonzero(cov)
    assert (rows.min(), rows.max(), cols.min(), cols.max()) == (10, 29, 10, 29)
    assert cov.sum() == 400 and np.all(np.abs(d[cov] - 100.0) <= 1e-4)


def test_render_batch_and_odd_width(B, oracle_built):
  """Width not a multiple of 4 (scalar path), several objects per batch."""
  from bop_toolkit_b200 import synth
  from oracle import renderer_oracle
  W, H = 150, 97
  K = synth.LM_K.copy(); K[:2] *= 0.22
  ms = B.ModelSet()
  meshes = [synth.make_mesh(10, 20, 160.0, seed=s) for s in (1, 2)]
  for k, (p, f) in enumerate(meshes):
    ms.add(k + 1, p, f, None, 160.0)
  n = 6
  _, _, R, t = synth.make_pose_pairs(n, seed=8, K=K, width=W, height=H, z_range=(500, 900))
  obj = (np.arange(n) % 2).astype(np.int32)
  cams = np.tile([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], (n, 1))
  d = B.render_batch(ms, obj, B.pack_poses(R, t), cams, H, W).cpu().numpy()
  for i in range(n):
    p, f = meshes[obj[i]]
    ref = renderer_oracle.render_depth(p, f, R[i], t[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2], H, W)
    sil, mis, far, rel, bad = _render_compare(d[i], ref)
    assert far == 0 and mis <= 2 and bad.sum() <= 2


def _vsd_case(B, n, seed, H=480, W=640, scale=1.0, n_lat=50, n_lon=100):
  from bop_toolkit_b200 import synth
  from oracle import renderer_oracle
  K = synth.LM_K.copy(); K[:2] *= scale
  pts, faces = synth.make_mesh(n_lat, n_lon, 200.0, seed=seed)
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=seed + 1, K=K, width=W, height=H,
                                             z_range=(450.0, 1400.0), outlier_frac=0.2)
  R_e[0], t_e[0] = R_g[0], t_g[0]
  ren = renderer_oracle.RendererOracle(W, H)
  ren.add_object(1, pts=pts, faces=faces)
  depth = np.stack([synth.degrade_depth(
    ren.render_object(1, R_g[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2])['depth'], seed=seed + 50 + i, K=K)
    for i in range(n)])
  return K, pts, faces, R_e, t_e, R_g, t_g, depth, ren


def test_vsd_fused_vs_oracle_full_size(B, oracle_built):
  """configs[1] shape (640x480, delta 15, taus .05-.5, step), 64 pairs against the oracle."""
  from oracle import pose_error_oracle as po
  n = 64
  K, pts, faces, R_e, t_e, R_g, t_g, depth, ren = _vsd_case(B, n, seed=30)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 200.0)
  taus = list(np.arange(0.05, 0.51, 0.05))
  Ks = np.tile(K.reshape(1, 9), (n, 1))
  pe, pg = _poses(B, R_e, t_e), _poses(B, R_g, t_g)
  err, cnt = B.vsd_batch(ms, 0, pe, pg, depth, np.arange(n, dtype=np.int32), Ks, 15, taus,
                         True, 'step', gate=True, return_counts=True)
  err, cnt = err.cpu().numpy(), cnt.cpu().numpy()
  n_gated = 0
  for i in range(n):
    if not po.overlapping_sphere_projections(100.0, t_e[i], t_g[i]):
      ref = [1.0] * len(taus)       # eval_calc_errors.py:312-313
      n_gated += 1
      assert cnt[i, 0] == 0
    else:
      ref, rc = po.vsd_from_depth(
        ren.render_object(1, R_e[i], t_e[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2])['depth'],
        ren.render_object(1, R_g[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2])['depth'],
        depth[i], K, 15, taus, True, 200.0, 'step', return_counts=True)
      # union / inter pixel counts agree to a handful of silhouette pixels
      assert abs(int(cnt[i, 0]) - rc[0]) <= 0.002 * rc[0] + 3
    assert np.abs(err[i] - np.array(ref)).max() <= 1e-3, (i, err[i], ref)
  assert err[0].max() == 0.0 or err[0].max() < 0.2    # est == gt
  assert np.all(np.diff(err, axis=1) <= 1e-15)        # step VSD non-increasing in tau
  # u16 depth input with depth_scale gives the same result as the float32 image it encodes
  d16 = np.round(depth / 0.1).astype(np.uint16)
  err16 = B.vsd_batch(ms, 0, pe, pg, d16, np.arange(n, dtype=np.int32), Ks, 15, taus, True, 'step',
                      gate=True, depth_scale=np.full(n, 0.1, np.float32)).cpu().numpy()
  errf = B.vsd_batch(ms, 0, pe, pg, d16.astype(np.float32) * np.float32(0.1), np.arange(n, dtype=np.int32),
                     Ks, 15, taus, True, 'step', gate=True).cpu().numpy()
  assert np.array_equal(err16, errf)


def test_vsd_fused_equals_render_plus_k2(B):
  """Internal consistency at full size: fused kernel == (our renderer -> K2 pixel pass), bit for
  bit, for step / tlinear / bop18, shared depth image, small z-tiles (many bands)."""
  from bop_toolkit_b200 import _lib, synth
  n = 16
  K, pts, faces, R_e, t_e, R_g, t_g, depth, _ = _vsd_case(B, 2, seed=40)
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=41, z_range=(350.0, 900.0), outlier_frac=0.1)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 200.0)
  pe, pg = _poses(B, R_e, t_e), _poses(B, R_g, t_g)
  im = (np.arange(n) % 2).astype(np.int32)
  Ks = np.tile(K.reshape(1, 9), (2, 1))
  cams = np.tile([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], (n, 1))
  de = B.render_batch(ms, 0, pe, cams, 480, 640)
  dg = B.render_batch(ms, 0, pg, cams, 480, 640)
  taus = [0.05, 0.2, 0.5]
  try:
    for tile in (0, 4096):
      if tile:
        _lib.check(_lib.lib().bop_vsd_set_tuning(tile, 256, 2, 4))
      for cost, mode in (('step', 'bop19'), ('tlinear', 'bop19'), ('step', 'bop18')):
        a, ca = B.vsd_batch(ms, 0, pe, pg, depth, im, Ks, 15, taus, True, cost, mode, return_counts=True)
        b, cb = B.vsd_from_depth_batch(de, dg, depth, im, Ks, 200.0, 15, taus, True, cost, mode,
                                       return_counts=True)
        assert np.array_equal(ca.cpu().numpy()[:, :2], cb.cpu().numpy()[:, :2])
        if cost == 'step':
          assert np.array_equal(a.cpu().numpy(), b.cpu().numpy())
        else:
          assert np.allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=0, atol=1e-11)
  finally:
    _lib.check(_lib.lib().bop_vsd_set_tuning(20480, 1024, 1, 256))


def test_c_abi_host_entry_points(B, golden):
  """The `_host` C-ABI layer (what a torch-free ctypes/cgo binding would call)."""
  from bop_toolkit_b200 import _lib
  g = golden
  lib = _lib.lib()
  pts = np.ascontiguousarray(g['pts_small'])
  syms = np.ascontiguousarray(g['sym_1'])
  faces = np.ascontiguousarray(g['faces_small'].astype(np.int32))
  desc = _lib.ModelDev()
  desc_dev = ctypes.c_void_p()
  _lib.check(lib.bop_model_upload(pts.ctypes.data, len(pts), faces.ctypes.data, len(faces),
                                  syms.ctypes.data, len(syms), 150.0, ctypes.byref(desc),
                                  ctypes.byref(desc_dev)))
  n = g['R_e'].shape[0]
  pe, pg = _poses(B, g['R_e'], g['t_e']), _poses(B, g['R_g'], g['t_g'])
  out = np.empty(n)
  _lib.check(lib.bop_mssd_host(desc_dev, len(syms), pe.ctypes.data, pg.ctypes.data, n, 0, out.ctypes.data))
  _close(out, g['mssd_f32_1'])
  Ks = np.ascontiguousarray(np.tile(g['K'].reshape(1, 9), (n, 1)))
  _lib.check(lib.bop_mspd_host(desc_dev, len(syms), pe.ctypes.data, pg.ctypes.data, Ks.ctypes.data, n, out.ctypes.data))
  _close(out, g['mspd_f32_1'])
  _lib.check(lib.bop_add_host(desc_dev, len(pts), pe.ctypes.data, pg.ctypes.data, n, 0, out.ctypes.data))
  _close(out, g['add_f32'])
  _lib.check(lib.bop_adi_host(desc_dev, len(pts), pe.ctypes.data, pg.ctypes.data, n, 0, out.ctypes.data))
  _close(out, g['adi_f32'], atol=1e-6)
  re, te = np.empty(n), np.empty(n)
  _lib.check(lib.bop_re_te_host(pe.ctypes.data, pg.ctypes.data, n, re.ctypes.data, te.ctypes.data))
  _close(re, g['re'], atol=1e-6); _close(te, g['te'])
  _lib.check(lib.bop_model_free(ctypes.byref(desc), desc_dev))
  # VSD + render through host buffers
  vp = np.ascontiguousarray(g['vsd_pts']); vf = np.ascontiguousarray(g['vsd_faces'].astype(np.int32))
  _lib.check(lib.bop_model_upload(vp.ctypes.data, len(vp), vf.ctypes.data, len(vf), None, 0, 180.0,
                                  ctypes.byref(desc), ctypes.byref(desc_dev)))
  nv = g['vsd_R_e'].shape[0]
  H, W = g['vsd_depth_test'].shape[1:]
  pe, pg = _poses(B, g['vsd_R_e'], g['vsd_t_e']), _poses(B, g['vsd_R_g'], g['vsd_t_g'])
  taus = list(g['vsd_taus'])
  prm = B.make_vsd_params(H, W, 15, taus)
  depth = np.ascontiguousarray(g['vsd_depth_test'])
  im = np.arange(nv, dtype=np.int32)
  Ks = np.ascontiguousarray(np.tile(g['vsd_K'].reshape(1, 9), (nv, 1)))
  err = np.empty((nv, len(taus)))
  _lib.check(lib.bop_vsd_host(desc_dev, pe.ctypes.data, pg.ctypes.data, depth.ctypes.data, None,
                              im.ctypes.data, Ks.ctypes.data, nv, nv, ctypes.byref(prm), err.ctypes.data))
  assert np.abs(err - g['vsd_step']).max() <= 3e-3
  K = g['vsd_K']
  cams = np.ascontiguousarray(np.tile([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], (nv, 1)))
  d = np.empty((nv, H, W), np.float32)
  _lib.check(lib.bop_render_depth_host(desc_dev, pg.ctypes.data, cams.ctypes.data, nv, H, W, d.ctypes.data))
  ref = g['vsd_render_gt']
  both = (d > 0) & (ref > 0)
  assert ((d > 0) != (ref > 0)).sum() <= 8
  assert (np.abs(d[both] - ref[both]) / ref[both] > 1e-3).sum() <= 8
  _lib.check(lib.bop_model_free(ctypes.byref(desc), desc_dev))
  assert lib.bop_kernel_launch_count() > 0


def test_dense_model_adi_add(B):
  """configs[3] at full size: a 99,906-vertex model (ADI = 10^10 distances per pair), 64
  symmetries, three pairs against the oracle (cKDTree ADI, numpy MSSD / MSPD / ADD)."""
  from bop_toolkit_b200 import synth
  from oracle import pose_error_oracle as po
  pts, faces = synth.make_dense_mesh(100000, 250.0, seed=6)
  assert len(pts) > 99000
  syms = po.get_symmetry_transformations({'symmetries_discrete': synth.discrete_z_symmetries(64)}, 0.01)
  assert len(syms) == 64
  ms = B.ModelSet()
  ms.add(1, pts, faces, syms, 250.0)
  n = 3
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=7, outlier_frac=0.0)
  pe, pg = _poses(B, R_e, t_e), _poses(B, R_g, t_g)
  col = lambda t: t.reshape(3, 1)
  _close(B.adi_batch(ms, 0, pe, pg).cpu().numpy(),
         [po.adi(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), pts) for i in range(n)], atol=1e-6)
  _close(B.add_batch(ms, 0, pe, pg).cpu().numpy(),
         [po.add(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), pts) for i in range(n)])
  _close(B.mssd_batch(ms, 0, pe, pg).cpu().numpy(),
         [po.mssd(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), pts, syms) for i in range(n)])
  _close(B.mspd_batch(ms, 0, pe, pg, synth.LM_K).cpu().numpy(),
         [po.mspd(R_e[i], col(t_e[i]), R_g[i], col(t_g[i]), synth.LM_K, pts, syms) for i in range(n)])


def test_vsd_zero_copy_pinned_depth(B):
  """depth_test read straight from pinned host memory (kernel pulls ROI pixels over PCIe):
  identical results to the device-resident copy, far fewer bytes moved."""
  import torch
  from bop_toolkit_b200 import _lib
  n = 6
  K, pts, faces, R_e, t_e, R_g, t_g, depth, _ = _vsd_case(B, n, seed=60)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 200.0)
  taus = list(np.arange(0.05, 0.51, 0.05))
  Ks = np.tile(K.reshape(1, 9), (n, 1))
  pe, pg = _poses(B, R_e, t_e), _poses(B, R_g, t_g)
  im = np.arange(n, dtype=np.int32)
  d16 = torch.from_numpy(np.round(depth / 0.1).astype(np.uint16).view(np.int16)).pin_memory()
  sc = np.full(n, 0.1, np.float32)
  a = B.vsd_batch(ms, 0, pe, pg, d16, im, Ks, 15, taus, True, 'step', depth_scale=sc, zero_copy=True)
  fetched = B.depth_bytes_fetched(ms.device, _lib.DEPTH_U16)
  b = B.vsd_batch(ms, 0, pe, pg, d16, im, Ks, 15, taus, True, 'step', depth_scale=sc)
  assert torch.equal(a, b)
  assert 0 < fetched < 0.5 * d16.numel() * 2
  f32 = torch.from_numpy(depth).pin_memory()
  c = B.vsd_batch(ms, 0, pe, pg, f32, im, Ks, 15, taus, True, 'step', zero_copy=True)
  d = B.vsd_batch(ms, 0, pe, pg, depth, im, Ks, 15, taus, True, 'step')
  assert torch.equal(c, d)
  with pytest.raises(ValueError):
    B.vsd_batch(ms, 0, pe, pg, torch.from_numpy(depth), im, Ks, 15, taus, zero_copy=True)


# ------------------------------------------------------------------ mask-overlap errors
def test_mask_overlap_bit_exact_given_reference_renders(B, golden):
  """cus / cou_bb_proj / cou_mask from identical depth images: integer pixel statistics, so the
  errors equal the reference's bit for bit."""
  from bop_toolkit_b200 import pose_error as pe
  g = golden
  st = B.mask_overlap_from_depth_batch(g['vsd_render_est'], g['vsd_render_gt'])
  assert np.array_equal(st['cus'].cpu().numpy(), g['cus'])
  assert np.array_equal(st['cus'].cpu().numpy(), g['cou_mask'])
  assert np.array_equal(st['cou_bb'].cpu().numpy(), g['cou_bb_proj'])
  assert np.array_equal(st['bbox'].cpu().numpy()[:, :4], g['bbox2d'])
  n = g['cus'].shape[0]
  for i in range(n):
    assert pe.cou_mask(g['vsd_render_est'][i] > 0, g['vsd_render_gt'][i] > 0) == g['cou_mask'][i]
  for b, e in zip(g['cou_bb_in'], g['cou_bb']):
    assert pe.cou_bb(list(map(int, b[0])), list(map(int, b[1]))) == e
  # empty masks: union == 0 -> 1.0 (pose_error.py:250-252)
  z = np.zeros((1, 8, 12), np.float32)
  st = B.mask_overlap_from_depth_batch(z, z)
  assert float(st['cus'].cpu()[0]) == 1.0 and np.isnan(float(st['cou_bb'].cpu()[0]))

  class Foreign(object):   # duck-typed renderer handing back the reference's renders
    def __init__(self):
      self.k = 0
    def render_object(self, obj_id, R, t, fx, fy, cx, cy):
      i, which = divmod(self.k, 2)
      self.k += 1
      return {'depth': (g['vsd_render_est'] if which == 0 else g['vsd_render_gt'])[i % n]}
  ren = Foreign()
  for i in range(n):
    a = (g['vsd_R_e'][i], g['vsd_t_e'][i].reshape(3, 1), g['vsd_R_g'][i], g['vsd_t_g'][i].reshape(3, 1), g['vsd_K'], ren, 7)
    assert pe.cus(*a) == g['cus'][i]
  ren.k = 0
  for i in range(n):
    a = (g['vsd_R_e'][i], g['vsd_t_e'][i].reshape(3, 1), g['vsd_R_g'][i], g['vsd_t_g'][i].reshape(3, 1), g['vsd_K'], ren, 7)
    assert pe.cou_bb_proj(*a) == g['cou_bb_proj'][i]


def test_cus_fused_vs_reference_golden(B, golden):
  """cus / cou_bb_proj with the library's own rasteriser (fused, nothing leaves the SM but
  counts) against the reference run on the oracle's renders.  Budget: the silhouettes may differ
  in <= 0.5 % of their edge pixels (1/256-px snapping), i.e. a few pixels of some thousands."""
  from bop_toolkit_b200 import pose_error as pe, renderer
  g = golden
  H, W = g['vsd_depth_test'].shape[1:]
  ren = renderer.create_renderer(W, H, 'cuda', mode='depth')
  ren.add_object(7, pts=g['vsd_pts'], faces=g['vsd_faces'], diameter=180.0)
  n = g['cus'].shape[0]
  pe_, pg_ = _poses(B, g['vsd_R_e'], g['vsd_t_e']), _poses(B, g['vsd_R_g'], g['vsd_t_g'])
  st = B.mask_overlap_batch(ren.model_set, 0, pe_, pg_, g['vsd_K'], H, W)
  ref = B.mask_overlap_from_depth_batch(g['vsd_render_est'], g['vsd_render_gt'])
  cnt, cnt_ref = st['counts'].cpu().numpy(), ref['counts'].cpu().numpy()
  assert np.all(np.abs(cnt - cnt_ref) <= 3), (cnt, cnt_ref)
  assert np.allclose(st['cus'].cpu().numpy(), g['cus'], rtol=0, atol=2e-3)
  assert np.all(np.abs(st['bbox'].cpu().numpy() - ref['bbox'].cpu().numpy()) <= 1)
  assert np.allclose(st['cou_bb'].cpu().numpy(), g['cou_bb_proj'], rtol=0, atol=2e-2)
  # fused == (render -> stats), bit for bit
  de = B.render_batch(ren.model_set, 0, pe_, np.tile(g['vsd_K'].reshape(9)[[0, 4, 2, 5]], (n, 1)), H, W)
  dg = B.render_batch(ren.model_set, 0, pg_, np.tile(g['vsd_K'].reshape(9)[[0, 4, 2, 5]], (n, 1)), H, W)
  two = B.mask_overlap_from_depth_batch(de, dg)
  for k in ('cus', 'cou_bb', 'counts', 'bbox'):
    assert np.array_equal(st[k].cpu().numpy(), two[k].cpu().numpy(), equal_nan=True), k
  for i in range(3):
    a = (g['vsd_R_e'][i], g['vsd_t_e'][i].reshape(3, 1), g['vsd_R_g'][i], g['vsd_t_g'][i].reshape(3, 1), g['vsd_K'], ren, 7)
    assert pe.cus(*a) == float(st['cus'].cpu()[i])
    assert pe.cou_bb_proj(*a) == float(st['cou_bb'].cpu()[i])
  # sphere gate of the driver: a far-away estimate -> 1.0
  far = pe_.copy()
  far[:, 9] += 5000.0
  gated = B.cus_batch(ren.model_set, 0, far, pg_, g['vsd_K'], H, W, gate=True).cpu().numpy()
  assert np.all(gated == 1.0)


# ------------------------------------------------------------------ verified back-face culling
def test_backface_culling_is_output_preserving(B):
  """ModelSet.add culls back faces only for closed, consistently oriented meshes and only after
  probe renders agreed; silhouettes must be identical, depths equal up to float rounding."""
  from bop_toolkit_b200 import synth
  pts, faces = synth.make_mesh(30, 60, 180.0, seed=4)
  n = 48
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=9, z_range=(300.0, 1400.0))
  pe, pg = _poses(B, R_e, t_e), _poses(B, R_g, t_g)
  cam = synth.LM_K.reshape(9)[[0, 4, 2, 5]]
  plain, culled, flipped, opened = B.ModelSet(), B.ModelSet(), B.ModelSet(), B.ModelSet()
  plain.add(1, pts, faces, None, 180.0, cull=False)
  culled.add(1, pts, faces, None, 180.0, cull='auto')
  flipped.add(1, pts, faces[:, ::-1].copy(), None, 180.0, cull='auto')       # inside-out winding
  opened.add(1, pts, faces[:-7], None, 180.0, cull='auto')                   # a hole: not closed
  default = B.ModelSet()
  default.add(1, pts, faces, None, 180.0)
  assert default.info[1]['cull'] == 0                           # the default is the reference's GL state: no culling
  assert plain.info[1]['cull'] == 0 and opened.info[1]['cull'] == 0
  assert culled.info[1]['cull'] in (1, -1) and flipped.info[1]['cull'] == -culled.info[1]['cull']
  ref = B.render_batch(plain, 0, pg, cam, 480, 640)
  n_diff = 0
  for ms in (culled, flipped):
    got = B.render_batch(ms, 0, pg, cam, 480, 640)
    assert torch_equal(got > 0, ref > 0)                          # identical silhouettes
    assert bool(((got - ref).abs() <= 1e-5 * ref).all())          # depth: float rounding only
    n_diff += int((got != ref).sum())
  print('culled vs unculled renders: %d of %d covered pixels differ in the last bits'
        % (n_diff // 2, int((ref > 0).sum())))
  depth = ref.clone()
  depth[depth == 0] = 1500.0
  taus = list(np.arange(0.05, 0.51, 0.05))
  Ks = np.tile(synth.LM_K.reshape(1, 9), (n, 1))
  args = (0, pe, pg, depth, np.arange(n, dtype=np.int32), Ks, 15, taus, True, 'step')
  v0, c0 = B.vsd_batch(plain, *args, return_counts=True)
  v1, c1 = B.vsd_batch(culled, *args, return_counts=True)
  assert int((c0 - c1).abs().max()) <= 2 and float((v0 - v1).abs().max()) <= 1e-3
  m0 = B.mask_overlap_batch(plain, 0, pe, pg, synth.LM_K, 480, 640)
  m1 = B.mask_overlap_batch(culled, 0, pe, pg, synth.LM_K, 480, 640)
  assert torch_equal(m0['counts'], m1['counts']) and torch_equal(m0['bbox'], m1['bbox'])


def torch_equal(a, b):
  import torch
  return bool(torch.equal(a, b))


# ------------------------------------------------------------------ analytic known-answer scenes
@pytest.mark.gpu
def test_render_known_answer_scenes(B):
  """The CUDA rasteriser on scenes with a closed-form answer (the same two the oracles are checked
  on in tests/test_oracle_golden.py): a tilted quad (depth = plane depth along the pixel ray) and
  two interpenetrating quads (the nearer plane wins).  Pixels whose hit point is within 3 mm of a
  quad border are left out (edge ownership is implementation-defined)."""
  def plane_depth(n, d, K, H, W):
    ys, xs = np.mgrid[0:H, 0:W]
    rx, ry = (xs + 0.5 - K[0, 2]) / K[0, 0], (ys + 0.5 - K[1, 2]) / K[1, 1]
    return d / (n[0] * rx + n[1] * ry + n[2]), rx, ry
  # tilted plane
  K = np.array([[120.0, 0, 80.0], [0, 118.0, 60.5], [0, 0, 1]])
  H, W = 120, 160
  n, d = np.array([0.25, -0.125, 1.0]), 500.0
  xy = np.array([[-250.0, -200.0], [300.0, -150.0], [280.0, 220.0], [-220.0, 180.0]])
  pts = np.column_stack([xy, (d - n[0] * xy[:, 0] - n[1] * xy[:, 1]) / n[2]])
  ms = B.ModelSet()
  ms.add(1, pts, np.array([[0, 1, 2], [0, 2, 3]]), None, 600.0)
  pose = B.pack_poses(np.eye(3)[None], np.zeros((1, 3)))
  cam = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]])
  got = B.render_batch(ms, 0, pose, cam, H, W)[0].cpu().numpy()
  want, _, _ = plane_depth(n, d, K, H, W)
  cov = got > 0
  assert cov.sum() > 0.5 * H * W
  assert np.abs(got[cov] - want[cov]).max() <= 1e-5 * want[cov].max()      # bar: 1e-3 relative
  # interpenetrating quads
  K = np.array([[100.0, 0, 64.0], [0, 100.0, 48.0], [0, 0, 1]])
  H, W = 96, 128
  quad = np.array([[-200.0, -150.0], [200.0, -150.0], [200.0, 150.0], [-200.0, 150.0]])
  pts = np.vstack([np.column_stack([quad, 400.0 + 0.5 * quad[:, 0]]), np.column_stack([quad, 400.0 - 0.5 * quad[:, 0]])])
  ms2 = B.ModelSet()
  ms2.add(1, pts, np.array([[0, 1, 2], [0, 2, 3], [4, 5, 6], [4, 6, 7]]), None, 600.0)
  cam = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]])
  got = B.render_batch(ms2, 0, pose, cam, H, W)[0].cpu().numpy()
  da, rx, ry = plane_depth(np.array([-0.5, 0.0, 1.0]), 400.0, K, H, W)
  db, _, _ = plane_depth(np.array([0.5, 0.0, 1.0]), 400.0, K, H, W)
  inside = lambda dz, m: (np.abs(rx * dz) < 200.0 - m) & (np.abs(ry * dz) < 150.0 - m)
  want = np.minimum(np.where(inside(da, 0.0), da, np.inf), np.where(inside(db, 0.0), db, np.inf))
  safe = (inside(da, 3.0) | ~inside(da, -3.0)) & (inside(db, 3.0) | ~inside(db, -3.0)) & np.isfinite(want)
  assert safe.sum() > 0.5 * H * W and (got[safe] > 0).all()
  assert np.abs(got[safe] - want[safe]).max() <= 1e-5 * want[safe].max()
  outside = ~inside(da, -3.0) & ~inside(db, -3.0)
  assert not got[outside].any()


# ------------------------------------------------------------------ pairs that share their GT pose
@pytest.mark.gpu
def test_vsd_grouped_equals_ungrouped(B, oracle_built):
  """bop_vsd_grouped_dev (GT pose projected, sorted and rasterised once per group of estimates)
  against bop_vsd_dev on the same pairs: bit-identical errors and counts -- groups of 1..5
  estimates incl. gated ones, estimates outside the image, a group above 32 (cut), step and
  tlinear cost, close-ups that need several row bands."""
  from bop_toolkit_b200 import synth
  K = synth.LM_K
  rng = np.random.default_rng(5)
  pts, faces = synth.make_mesh(40, 80, 240.0, seed=61)
  pts2, faces2 = synth.make_mesh(24, 48, 120.0, seed=62)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 240.0)
  ms.add(2, pts2, faces2, None, 120.0)
  n_gt = 40
  _, _, R_g, t_g = synth.make_pose_pairs(n_gt, seed=63, z_range=(330.0, 1300.0), outlier_frac=0.0)
  k_of = rng.integers(1, 6, size=n_gt)
  k_of[3] = 37                                         # above kMaxGroup
  gt_of = np.repeat(np.arange(n_gt), k_of)
  P = len(gt_of)
  R_e = np.empty((P, 3, 3)); t_e = np.empty((P, 3))
  for i in range(P):
    g = gt_of[i]
    ang = np.radians(rng.uniform(0, 20))
    R_e[i] = R_g[g] @ synth.axis_angle(rng.normal(size=3), ang)
    t_e[i] = t_g[g] + rng.normal(size=3) * np.array([6.0, 6.0, 20.0])
  far = rng.choice(P, 12, replace=False)
  t_e[far[:6], 0] += 700.0                             # fail the sphere gate
  t_e[far[6:], :2] += 4000.0                           # outside the image, gated too
  obj = (gt_of % 2).astype(np.int32)                   # model row: alternates per GT
  im = gt_of.astype(np.int32)
  depth = np.full((n_gt, 480, 640), 900.0, np.float32)
  depth += rng.normal(0, 30.0, depth.shape).astype(np.float32)
  depth[:, ::9, ::7] = 0.0
  Ks = np.tile(K.reshape(1, 9), (n_gt, 1))
  pe, pg = B.pack_poses(R_e, t_e), B.pack_poses(R_g[gt_of], t_g[gt_of])
  perm = rng.permutation(P)                            # the caller's order is arbitrary
  pe, pg, obj, im = pe[perm], pg[perm], obj[perm], im[perm]
  order, goff = B.group_pairs_by_gt(obj, pg, im)
  assert goff[-1] == P and (np.diff(goff) <= 32).all() and len(goff) - 1 >= n_gt + 1
  taus = [0.05 * (k + 1) for k in range(10)]
  for cost in ('step', 'tlinear'):
    e0, c0 = B.vsd_batch(ms, obj, pe, pg, depth, im, Ks, 15.0, taus, True, cost, 'bop19', gate=True,
                         return_counts=True)
    e1, c1 = B.vsd_batch(ms, obj[order], pe[order], pg[order], depth, im[order], Ks, 15.0, taus, True, cost,
                         'bop19', gate=True, return_counts=True, group_off=goff)
    e0, c0, e1, c1 = e0.cpu().numpy(), c0.cpu().numpy(), e1.cpu().numpy(), c1.cpu().numpy()
    assert np.array_equal(c0[order], c1), cost
    if cost == 'step':
      assert np.array_equal(e0[order], e1)
    else:
      assert np.abs(e0[order] - e1).max() <= 1e-12      # float64 sums of the tlinear cost: order of the adds
    assert (c1[:, 0] > 0).sum() > P // 2 and (e1 == 1.0).all(axis=1).sum() >= 12


@pytest.mark.parametrize('H,W,scale,cost', [(97, 150, 0.22, 'step'), (120, 161, 0.25, 'tlinear'), (61, 64, 0.1, 'step'),
                                            (240, 322, 0.5, 'tlinear')])
def test_vsd_fused_vs_oracle_odd_sizes(B, oracle_built, H, W, scale, cost):
  """Image sizes that are not multiples of 4 (scalar depth loads, padded regions of interest at the
  right border), scaled intrinsics and both costs, 8 pairs each against the oracle: 1e-3 per tau."""
  from oracle import pose_error_oracle as po
  n = 8
  K, pts, faces, R_e, t_e, R_g, t_g, depth, ren = _vsd_case(B, n, seed=70 + W, H=H, W=W, scale=scale, n_lat=20, n_lon=40)
  ms = B.ModelSet()
  ms.add(1, pts, faces, None, 200.0)
  taus = list(np.arange(0.05, 0.51, 0.05))
  Ks = np.tile(K.reshape(1, 9), (n, 1))
  err, cnt = B.vsd_batch(ms, 0, _poses(B, R_e, t_e), _poses(B, R_g, t_g), depth, np.arange(n, dtype=np.int32), Ks, 15,
                         taus, True, cost, return_counts=True)
  err, cnt = err.cpu().numpy(), cnt.cpu().numpy()
  seen = 0
  for i in range(n):
    ref, rc = po.vsd_from_depth(
      ren.render_object(1, R_e[i], t_e[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2])['depth'],
      ren.render_object(1, R_g[i], t_g[i], K[0, 0], K[1, 1], K[0, 2], K[1, 2])['depth'],
      depth[i], K, 15, taus, True, 200.0, cost, return_counts=True)
    assert abs(int(cnt[i, 0]) - rc[0]) <= 0.01 * rc[0] + 3, (i, cnt[i, 0], rc[0])
    # a silhouette pixel more or less moves the error of a small mask by 1 / union
    tol = 1e-3 if rc[0] >= 3000 else 1e-3 + 3.0 / max(rc[0], 1)
    assert np.abs(err[i] - np.array(ref)).max() <= tol, (i, err[i], ref, rc)
    seen += rc[0] > 0
  assert seen >= n // 2

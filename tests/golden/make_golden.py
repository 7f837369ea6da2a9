"""Generates the golden vectors under tests/golden/ by running the REAL reference
(gist-ailab/bop_toolkit, mounted read-only at /root/reference) on seeded inputs.

Run here (the reference does not exist on the GPU box):   python tests/golden/make_golden.py
The reference needs a two-line shim on this interpreter: a stub `pytz`
(only misc.log uses it) and `numpy.float = float` (alias removed in numpy >= 1.24).
The reference's renderers cannot run (no graphics stack); pose_error.vsd is driven with the
oracle's CPU rasteriser plugged in through the duck-typed `renderer` argument, so everything
after the two render calls is the reference's own code.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

# --- shim + import of the reference ---------------------------------------------------
pytz = types.ModuleType('pytz')
pytz.utc = types.SimpleNamespace(localize=lambda dt: dt)
sys.modules['pytz'] = pytz
np.float = float
sys.path.insert(0, '/root/reference')
import warnings
warnings.filterwarnings('ignore')
from bop_toolkit_lib import misc as ref_misc            # noqa: E402
from bop_toolkit_lib import pose_error as ref_pe        # noqa: E402
from bop_toolkit_lib import visibility as ref_vis       # noqa: E402

sys.path.remove('/root/reference')
from bop_toolkit_b200 import synth                      # noqa: E402
from oracle import renderer_oracle                      # noqa: E402


def main():
  rng = np.random.default_rng(20240501)
  out = {}

  # ---- per-vertex errors: 3 models x pairs, with several symmetry sets ---------------
  pts_small, faces_small = synth.make_mesh(12, 24, diameter=150.0, seed=3)        # 266 pts, f32-exact
  pts_f64 = pts_small * 1.0000001234 + 1e-7                                       # not f32-exact
  models_info = [
    {'diameter': 150.0},
    {'diameter': 150.0, 'symmetries_discrete': synth.discrete_z_symmetries(4)},
    {'diameter': 150.0, 'symmetries_continuous': [{'axis': [0, 0, 1], 'offset': [0, 0, 0]}]},
    {'diameter': 150.0, 'symmetries_discrete': synth.discrete_z_symmetries(2),
     'symmetries_continuous': [{'axis': [0, 1, 0], 'offset': [1.5, 0, -2.0]}]},
  ]
  steps = [0.01, 0.01, 0.085, 0.2]
  sym_sets = []
  for mi, st in zip(models_info, steps):
    syms = ref_misc.get_symmetry_transformations(mi, st)
    arr = np.array([np.hstack([s['R'].reshape(9), s['t'].reshape(3)]) for s in syms])
    sym_sets.append((syms, arr))
  for i, (_, arr) in enumerate(sym_sets):
    out['sym_%d' % i] = arr
  out['sym_steps'] = np.array(steps)

  n = 24
  R_e, t_e, R_g, t_g = synth.make_pose_pairs(n, seed=11, outlier_frac=0.15)
  # a few special pairs: identical, symmetric-equivalent
  R_e[0], t_e[0] = R_g[0], t_g[0]
  s = sym_sets[1][0][2]
  R_e[1], t_e[1] = R_g[1] @ s['R'], (R_g[1] @ s['t']).reshape(3) + t_g[1]
  K = synth.LM_K.copy()
  K_skew = K.copy(); K_skew[0, 1] = 0.7
  out.update(pts_small=pts_small, faces_small=faces_small, pts_f64=pts_f64, R_e=R_e, t_e=t_e,
             R_g=R_g, t_g=t_g, K=K, K_skew=K_skew)
  for mname, pts in (('f32', pts_small), ('f64', pts_f64)):
    for si, (syms, _) in enumerate(sym_sets):
      if mname == 'f64' and si not in (0, 1):
        continue
      ms = [ref_pe.mssd(R_e[i], t_e[i].reshape(3, 1), R_g[i], t_g[i].reshape(3, 1), pts, syms) for i in range(n)]
      mp = [ref_pe.mspd(R_e[i], t_e[i].reshape(3, 1), R_g[i], t_g[i].reshape(3, 1), K, pts, syms) for i in range(n)]
      out['mssd_%s_%d' % (mname, si)] = np.array(ms)
      out['mspd_%s_%d' % (mname, si)] = np.array(mp)
    out['mspd_skew_' + mname] = np.array(
      [ref_pe.mspd(R_e[i], t_e[i].reshape(3, 1), R_g[i], t_g[i].reshape(3, 1), K_skew, pts, sym_sets[1][0]) for i in range(n)])
    out['add_' + mname] = np.array([ref_pe.add(R_e[i], t_e[i].reshape(3, 1), R_g[i], t_g[i].reshape(3, 1), pts) for i in range(n)])
    out['adi_' + mname] = np.array([ref_pe.adi(R_e[i], t_e[i].reshape(3, 1), R_g[i], t_g[i].reshape(3, 1), pts) for i in range(n)])
    out['proj_' + mname] = np.array([ref_pe.proj(R_e[i], t_e[i].reshape(3, 1), R_g[i], t_g[i].reshape(3, 1), K, pts) for i in range(n)])
  out['re'] = np.array([ref_pe.re(R_e[i], R_g[i]) for i in range(n)])
  out['te'] = np.array([ref_pe.te(t_e[i].reshape(3, 1), t_g[i].reshape(3, 1)) for i in range(n)])
  out['sphere_proj_overlap'] = np.array(
    [ref_misc.overlapping_sphere_projections(75.0, t_e[i], t_g[i]) for i in range(n)])

  # ---- VSD on a small image (160x120, K scaled by 1/4) -------------------------------
  H, W = 120, 160
  Kv = K.copy(); Kv[:2] *= 0.25
  pts_v, faces_v = synth.make_mesh(16, 32, diameter=180.0, seed=5)
  ren = renderer_oracle.RendererOracle(W, H)
  ren.add_object(7, pts=pts_v, faces=faces_v)
  nv = 8
  Rv_e, tv_e, Rv_g, tv_g = synth.make_pose_pairs(nv, seed=21, K=Kv, width=W, height=H,
                                                 z_range=(500.0, 900.0), outlier_frac=0.0)
  Rv_e[0], tv_e[0] = Rv_g[0], tv_g[0]
  taus = list(np.arange(0.05, 0.51, 0.05))
  depth_tests, vsd_step, vsd_tlin, vsd_unnorm, ren_e, ren_g = [], [], [], [], [], []
  for i in range(nv):
    dg = ren.render_object(7, Rv_g[i], tv_g[i].reshape(3, 1), Kv[0, 0], Kv[1, 1], Kv[0, 2], Kv[1, 2])['depth']
    de = ren.render_object(7, Rv_e[i], tv_e[i].reshape(3, 1), Kv[0, 0], Kv[1, 1], Kv[0, 2], Kv[1, 2])['depth']
    dt = synth.degrade_depth(dg, seed=100 + i, K=Kv, occluder_prob=0.5)
    depth_tests.append(dt); ren_e.append(de); ren_g.append(dg)
    args = (Rv_e[i], tv_e[i].reshape(3, 1), Rv_g[i], tv_g[i].reshape(3, 1), dt, Kv, 15, taus)
    vsd_step.append(ref_pe.vsd(*args, True, 180.0, ren, 7, 'step'))
    vsd_tlin.append(ref_pe.vsd(*args, True, 180.0, ren, 7, 'tlinear'))
    vsd_unnorm.append(ref_pe.vsd(*args[:6], 15, [5.0, 20.0], False, 180.0, ren, 7, 'step'))
  out.update(vsd_pts=pts_v, vsd_faces=faces_v, vsd_K=Kv, vsd_R_e=Rv_e, vsd_t_e=tv_e, vsd_R_g=Rv_g,
             vsd_t_g=tv_g, vsd_taus=np.array(taus), vsd_depth_test=np.array(depth_tests),
             vsd_render_est=np.array(ren_e), vsd_render_gt=np.array(ren_g),
             vsd_step=np.array(vsd_step), vsd_tlinear=np.array(vsd_tlin),
             vsd_unnorm=np.array(vsd_unnorm))
  # ---- mask-overlap errors on the same renders (reference functions) -------------------
  cam = (Kv[0, 0], Kv[1, 1], Kv[0, 2], Kv[1, 2])
  out['cus'] = np.array([ref_pe.cus(Rv_e[i], tv_e[i].reshape(3, 1), Rv_g[i], tv_g[i].reshape(3, 1), Kv, ren, 7)
                         for i in range(nv)])
  out['cou_bb_proj'] = np.array([ref_pe.cou_bb_proj(Rv_e[i], tv_e[i].reshape(3, 1), Rv_g[i],
                                                    tv_g[i].reshape(3, 1), Kv, ren, 7) for i in range(nv)])
  out['cou_mask'] = np.array([ref_pe.cou_mask(ren_e[i] > 0, ren_g[i] > 0) for i in range(nv)])
  bbs = rng.integers(0, 60, size=(12, 2, 4))
  bbs[0, 1] = bbs[0, 0]                      # identical boxes
  bbs[1, 1, :2] = bbs[1, 0, :2] + 500        # disjoint boxes
  out['cou_bb_in'] = bbs
  out['cou_bb'] = np.array([ref_pe.cou_bb(list(map(int, b[0])), list(map(int, b[1]))) for b in bbs])
  out['bbox2d'] = np.array([ref_misc.calc_2d_bbox(*np.nonzero(ren_e[i] > 0)[::-1]) for i in range(nv)])

  # dist image + visibility masks of pair 1 (reference functions, called directly)
  d_test = ref_misc.depth_im_to_dist_im_fast(depth_tests[1], Kv)
  d_gt = ref_misc.depth_im_to_dist_im_fast(ren_g[1], Kv)
  d_est = ref_misc.depth_im_to_dist_im_fast(ren_e[1], Kv)
  out['dist_test_1'] = d_test
  for mode in ('bop19', 'bop18'):
    vg = ref_vis.estimate_visib_mask_gt(d_test, d_gt, 15, visib_mode=mode)
    ve = ref_vis.estimate_visib_mask_est(d_test, d_est, vg, 15, visib_mode=mode)
    out['visib_gt_1_' + mode] = np.packbits(vg)
    out['visib_est_1_' + mode] = np.packbits(ve)

  np.savez_compressed(os.path.join(HERE, 'pose_error_golden.npz'), **out)
  print('wrote', os.path.join(HERE, 'pose_error_golden.npz'),
        os.path.getsize(os.path.join(HERE, 'pose_error_golden.npz')) // 1024, 'KiB')


if __name__ == '__main__':
  main()

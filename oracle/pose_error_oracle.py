"""CPU oracle for the BOP pose-error hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A numpy restatement of the reference's algorithms (gist-ailab/bop_toolkit), each function
citing the reference file:line it follows.  Pinned against the reference itself: the script
tests/golden/make_golden.py imports the real `bop_toolkit_lib` from /root/reference, runs it
on seeded inputs and stores inputs + outputs under tests/golden/; tests/test_oracle_golden.py
checks this file against those vectors.  The rasterisation step has no runnable reference (no
graphics stack): for rendered pixels parity is UNPINNED (see oracle/raster_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may
import this module.
"""
import math

import numpy as np
from scipy import spatial


# --------------------------------------------------------------------------- misc.py
def transform_pts_Rt(pts, R, t):
  """bop_toolkit_lib/misc.py:266-276."""
  assert pts.shape[1] == 3
  pts_t = R.dot(pts.T) + t.reshape((3, 1))
  return pts_t.T


def project_pts(pts, K, R, t):
  """bop_toolkit_lib/misc.py:93-107."""
  assert pts.shape[1] == 3
  P = K.dot(np.hstack((R, t.reshape((3, 1)))))
  pts_h = np.hstack((pts, np.ones((pts.shape[0], 1))))
  pts_im = P.dot(pts_h.T)
  pts_im /= pts_im[2, :]
  return pts_im[:2, :].T


class Precomputer(object):
  """bop_toolkit_lib/misc.py:110-140: memoises the per-pixel ray factors across calls; they are
  recomputed only when the image shape or K changes (the reference's class-level cache)."""
  xs, ys = None, None
  pre_Xs, pre_Ys = None, None
  depth_im_shape = None
  K = None

  @staticmethod
  def precompute_lazy(depth_im, K):
    if depth_im.shape != Precomputer.depth_im_shape:
      Precomputer.depth_im_shape = depth_im.shape
      Precomputer.xs, Precomputer.ys = np.meshgrid(
        np.arange(depth_im.shape[1]), np.arange(depth_im.shape[0]))
      Precomputer.K = None   # the reference's second test is dead code; a new shape must refresh
    if Precomputer.K is None or not np.all(K == Precomputer.K):
      Precomputer.K = np.array(K, copy=True)
      Precomputer.pre_Xs = (Precomputer.xs - K[0, 2]) / np.float64(K[0, 0])
      Precomputer.pre_Ys = (Precomputer.ys - K[1, 2]) / np.float64(K[1, 1])
    return Precomputer.pre_Xs, Precomputer.pre_Ys


def depth_im_to_dist_im(depth_im, K):
  """bop_toolkit_lib/misc.py:143-163 (depth_im_to_dist_im_fast, with the Precomputer cache of
  :110-140 so that a call costs what the reference's does)."""
  pre_Xs, pre_Ys = Precomputer.precompute_lazy(depth_im, K)
  return np.sqrt(
    np.multiply(pre_Xs, depth_im) ** 2 +
    np.multiply(pre_Ys, depth_im) ** 2 +
    depth_im.astype(np.float64) ** 2)


def rotation_matrix(angle, direction):
  """bop_toolkit_lib/transform.py:302-343 (Gohlke), 3x3 part, no `point`."""
  sina, cosa = math.sin(angle), math.cos(angle)
  d = np.array(direction[:3], dtype=np.float64, copy=True)
  d /= math.sqrt(np.dot(d, d))
  R = np.diag([cosa, cosa, cosa])
  R += np.outer(d, d) * (1.0 - cosa)
  d = d * sina
  R += np.array([[0.0, -d[2], d[1]],
                 [d[2], 0.0, -d[0]],
                 [-d[1], d[0], 0.0]])
  return R


def get_symmetry_transformations(model_info, max_sym_disc_step):
  """bop_toolkit_lib/misc.py:43-90."""
  trans_disc = [{'R': np.eye(3), 't': np.array([[0, 0, 0]]).T}]
  if 'symmetries_discrete' in model_info:
    for sym in model_info['symmetries_discrete']:
      sym_4x4 = np.reshape(sym, (4, 4))
      trans_disc.append({'R': sym_4x4[:3, :3], 't': sym_4x4[:3, 3].reshape((3, 1))})
  trans_cont = []
  if 'symmetries_continuous' in model_info:
    for sym in model_info['symmetries_continuous']:
      axis = np.array(sym['axis'])
      offset = np.array(sym['offset']).reshape((3, 1))
      discrete_steps_count = int(np.ceil(np.pi / max_sym_disc_step))
      discrete_step = 2.0 * np.pi / discrete_steps_count
      for i in range(1, discrete_steps_count):
        R = rotation_matrix(i * discrete_step, axis)
        t = -R.dot(offset) + offset
        trans_cont.append({'R': R, 't': t})
  trans = []
  for tran_disc in trans_disc:
    if len(trans_cont):
      for tran_cont in trans_cont:
        R = tran_cont['R'].dot(tran_disc['R'])
        t = tran_cont['R'].dot(tran_disc['t']) + tran_cont['t']
        trans.append({'R': R, 't': t})
    else:
      trans.append(tran_disc)
  return trans


def overlapping_sphere_projections(radius, p1, p2):
  """bop_toolkit_lib/misc.py:309-331."""
  if p1[2] == 0 or p2[2] == 0:
    return False
  proj1 = (p1 / p1[2])[:2]
  proj2 = (p2 / p2[2])[:2]
  proj_dist = np.linalg.norm(proj1 - proj2)
  proj_dist_thresh = radius * (1.0 / p1[2] + 1.0 / p2[2])
  return bool(proj_dist < proj_dist_thresh)


# --------------------------------------------------------------------- visibility.py
def _estimate_visib_mask(d_test, d_model, delta, visib_mode='bop19'):
  """bop_toolkit_lib/visibility.py:9-42."""
  assert d_test.shape == d_model.shape
  d_diff = d_model.astype(np.float32) - d_test.astype(np.float32)
  if visib_mode == 'bop18':
    mask_valid = np.logical_and(d_test > 0, d_model > 0)
    return np.logical_and(d_diff <= delta, mask_valid)
  elif visib_mode == 'bop19':
    return np.logical_and(np.logical_or(d_diff <= delta, d_test == 0), d_model > 0)
  raise ValueError('Unknown visibility mode.')


def estimate_visib_mask_gt(d_test, d_gt, delta, visib_mode='bop19'):
  """bop_toolkit_lib/visibility.py:45-55."""
  return _estimate_visib_mask(d_test, d_gt, delta, visib_mode)


def estimate_visib_mask_est(d_test, d_est, visib_gt, delta, visib_mode='bop19'):
  """bop_toolkit_lib/visibility.py:58-75."""
  visib_est = _estimate_visib_mask(d_test, d_est, delta, visib_mode)
  return np.logical_or(visib_est, np.logical_and(visib_gt, d_est > 0))


# --------------------------------------------------------------------- pose_error.py
def vsd_from_depth(depth_est, depth_gt, depth_test, K, delta, taus, normalized_by_diameter,
                   diameter, cost_type='step', visib_mode='bop19', return_counts=False):
  """bop_toolkit_lib/pose_error.py:47-91 (everything after the two render calls)."""
  dist_test = depth_im_to_dist_im(depth_test, K)
  dist_gt = depth_im_to_dist_im(depth_gt, K)
  dist_est = depth_im_to_dist_im(depth_est, K)
  visib_gt = estimate_visib_mask_gt(dist_test, dist_gt, delta, visib_mode)
  visib_est = estimate_visib_mask_est(dist_test, dist_est, visib_gt, delta, visib_mode)
  visib_inter = np.logical_and(visib_gt, visib_est)
  visib_union = np.logical_or(visib_gt, visib_est)
  visib_union_count = visib_union.sum()
  visib_comp_count = visib_union_count - visib_inter.sum()
  dists = np.abs(dist_gt[visib_inter] - dist_est[visib_inter])
  if normalized_by_diameter:
    dists /= diameter
  counts = [int(visib_union_count), int(visib_inter.sum())]
  if visib_union_count == 0:
    errors = [1.0] * len(taus)
    counts += [0] * len(taus)
  else:
    errors = []
    for tau in taus:
      if cost_type == 'step':
        costs = dists >= tau
      elif cost_type == 'tlinear':
        costs = dists / tau
        costs[costs > 1.0] = 1.0
      else:
        raise ValueError('Unknown pixel matching cost.')
      counts.append(float(np.sum(costs)))
      errors.append(float((np.sum(costs) + visib_comp_count) / float(visib_union_count)))
  if return_counts:
    return errors, counts
  return errors


def vsd(R_est, t_est, R_gt, t_gt, depth_test, K, delta, taus, normalized_by_diameter, diameter,
        renderer, obj_id, cost_type='step'):
  """bop_toolkit_lib/pose_error.py:17-93."""
  fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
  depth_est = renderer.render_object(obj_id, R_est, t_est, fx, fy, cx, cy)['depth']
  depth_gt = renderer.render_object(obj_id, R_gt, t_gt, fx, fy, cx, cy)['depth']
  return vsd_from_depth(depth_est, depth_gt, depth_test, K, delta, taus, normalized_by_diameter,
                        diameter, cost_type)


def mssd(R_est, t_est, R_gt, t_gt, pts, syms):
  """bop_toolkit_lib/pose_error.py:96-118."""
  pts_est = transform_pts_Rt(pts, R_est, t_est)
  es = []
  for sym in syms:
    R_gt_sym = R_gt.dot(sym['R'])
    t_gt_sym = R_gt.dot(sym['t']) + t_gt
    pts_gt_sym = transform_pts_Rt(pts, R_gt_sym, t_gt_sym)
    es.append(np.linalg.norm(pts_est - pts_gt_sym, axis=1).max())
  return min(es)


def mspd(R_est, t_est, R_gt, t_gt, K, pts, syms):
  """bop_toolkit_lib/pose_error.py:121-144."""
  proj_est = project_pts(pts, K, R_est, t_est)
  es = []
  for sym in syms:
    R_gt_sym = R_gt.dot(sym['R'])
    t_gt_sym = R_gt.dot(sym['t']) + t_gt
    proj_gt_sym = project_pts(pts, K, R_gt_sym, t_gt_sym)
    es.append(np.linalg.norm(proj_est - proj_gt_sym, axis=1).max())
  return min(es)


def add(R_est, t_est, R_gt, t_gt, pts):
  """bop_toolkit_lib/pose_error.py:147-161."""
  pts_est = transform_pts_Rt(pts, R_est, t_est)
  pts_gt = transform_pts_Rt(pts, R_gt, t_gt)
  return np.linalg.norm(pts_est - pts_gt, axis=1).mean()


def adi(R_est, t_est, R_gt, t_gt, pts):
  """bop_toolkit_lib/pose_error.py:164-184.  Third-party: scipy.spatial.cKDTree (reference pins
  scipy==1.5.1, requirements.txt:1; exact k=1 Euclidean NN, so the version does not matter)."""
  pts_est = transform_pts_Rt(pts, R_est, t_est)
  pts_gt = transform_pts_Rt(pts, R_gt, t_gt)
  nn_index = spatial.cKDTree(pts_est)
  nn_dists, _ = nn_index.query(pts_gt, k=1)
  return nn_dists.mean()


def proj(R_est, t_est, R_gt, t_gt, K, pts):
  """bop_toolkit_lib/pose_error.py:217-232."""
  proj_est = project_pts(pts, K, R_est, t_est)
  proj_gt = project_pts(pts, K, R_gt, t_gt)
  return np.linalg.norm(proj_est - proj_gt, axis=1).mean()


def re(R_est, R_gt):
  """bop_toolkit_lib/pose_error.py:187-202."""
  assert R_est.shape == R_gt.shape == (3, 3)
  error_cos = float(0.5 * (np.trace(R_est.dot(np.linalg.inv(R_gt))) - 1.0))
  error_cos = min(1.0, max(-1.0, error_cos))
  return 180.0 * math.acos(error_cos) / np.pi


def te(t_est, t_gt):
  """bop_toolkit_lib/pose_error.py:205-214."""
  assert t_est.size == t_gt.size == 3
  return np.linalg.norm(t_gt - t_est)


def calc_2d_bbox(xs, ys):
  """misc.calc_2d_bbox without clipping (bop_toolkit_lib/misc.py:202-218): (x, y, w, h) with
  w = xmax - xmin, h = ymax - ymin."""
  return [xs.min(), ys.min(), xs.max() - xs.min(), ys.max() - ys.min()]


def iou(bb_a, bb_b):
  """misc.iou (bop_toolkit_lib/misc.py:236-263)."""
  tl_a, br_a = (bb_a[0], bb_a[1]), (bb_a[0] + bb_a[2], bb_a[1] + bb_a[3])
  tl_b, br_b = (bb_b[0], bb_b[1]), (bb_b[0] + bb_b[2], bb_b[1] + bb_b[3])
  w_inter = min(br_a[0], br_b[0]) - max(tl_a[0], tl_b[0])
  h_inter = min(br_a[1], br_b[1]) - max(tl_a[1], tl_b[1])
  if w_inter > 0 and h_inter > 0:
    area_inter = w_inter * h_inter
    return area_inter / float(bb_a[2] * bb_a[3] + bb_b[2] * bb_b[3] - area_inter)
  return 0.0


def cou_mask(mask_est, mask_gt):
  """pose_error.cou_mask (bop_toolkit_lib/pose_error.py:235-253)."""
  me, mg = mask_est.astype(bool), mask_gt.astype(bool)
  union_count = float(np.logical_or(mg, me).sum())
  return 1.0 - np.logical_and(mg, me).sum() / union_count if union_count > 0 else 1.0


def cus(R_est, t_est, R_gt, t_gt, K, renderer, obj_id):
  """pose_error.cus (bop_toolkit_lib/pose_error.py:256-286)."""
  fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
  depth_est = renderer.render_object(obj_id, R_est, t_est, fx, fy, cx, cy)['depth']
  depth_gt = renderer.render_object(obj_id, R_gt, t_gt, fx, fy, cx, cy)['depth']
  return cou_mask(depth_est > 0, depth_gt > 0)


def cou_bb(bb_est, bb_gt):
  """pose_error.cou_bb (bop_toolkit_lib/pose_error.py:289-297)."""
  return 1.0 - iou(bb_est, bb_gt)


def cou_bb_proj(R_est, t_est, R_gt, t_gt, K, renderer, obj_id):
  """pose_error.cou_bb_proj (bop_toolkit_lib/pose_error.py:300-330)."""
  fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
  depth_est = renderer.render_object(obj_id, R_est, t_est, fx, fy, cx, cy)['depth']
  depth_gt = renderer.render_object(obj_id, R_gt, t_gt, fx, fy, cx, cy)['depth']
  ys_est, xs_est = (depth_est > 0).nonzero()
  ys_gt, xs_gt = (depth_gt > 0).nonzero()
  return 1.0 - iou(calc_2d_bbox(xs_est, ys_est), calc_2d_bbox(xs_gt, ys_gt))

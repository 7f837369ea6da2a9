"""All correctness thresholds of an error type in one pass.

eval_bop19_pose.py scores every (tau, threshold) setting with a separate run of
eval_calc_scores.py: 120 times the same grouping, sorting and validity bookkeeping, and a pure
Python greedy matching per (image, object).  Only the thresholds differ between those runs.
ThresholdSweep builds the threshold-independent part once per error type --

  * the GT rows in the order pose_matching.match_poses_scene emits them (:111-160) with their
    validity (eval_calc_scores.py:198-226),
  * per (scene, image, object) the estimates ranked by score (stable, cut at n_top) and, per
    estimate, its candidate GT rows in dict order with the normalised errors (:235-247; invalid
    GT poses carry +inf and can never be picked),

-- and then runs the greedy matching of pose_matching.match_poses (:39-83) for ALL thresholds
(and, for VSD, any tau column) at once: rank by rank, each estimate takes the still-unmatched
candidate with the lowest error if that is below the threshold (numpy argmin returns the first
minimum, which is what the reference's strict `<` against the best-so-far keeps).  Estimates of
one rank belong to different (image, object) groups, so their candidate rows are disjoint and a
rank is one vectorised step.  Counts, scores_*.json and matches_*.json come out identical to
eval_calc_scores.score_errors (tests/test_scores.py compares the two on the reference's own
error files); error lists with more than one element ('ad', 'rete') are not handled here --
use score_errors for those.
"""
import numpy as np

from . import eval_calc_scores as ecs
from . import score


class ThresholdSweep(object):

  def __init__(self, p, err_type, n_top, dataset, split, split_type, scene_errs_by_scene, ctx=None):
    """p: eval_calc_scores parameters (visib_gt_min, normalized_by_*, datasets_path, targets_filename).
    scene_errs_by_scene: {scene_id: records}; every errors list must have the same length C
    (C = number of VSD taus for records straight out of calc_errors, else 1)."""
    ctx = {} if ctx is None else ctx
    ecs.load_context(p, dataset, split, split_type, ctx)
    self.n_top = n_top
    self.scene_ids = ctx['dp_split']['scene_ids']
    self.obj_ids = ctx['dp_model']['obj_ids']
    by_diam = err_type in p['normalized_by_diameter']
    by_width = err_type in p['normalized_by_im_width']
    factor = 640.0 / float(ctx['dp_split']['im_size'][0])
    models_info = ctx['models_info']

    row_scene, row_im, row_obj, row_gt, row_valid = [], [], [], [], []
    est_rank, est_id, est_score = [], [], []
    # candidates (estimate, GT row) as flat Python lists, turned into arrays once at the end: a
    # YCB-V-size run has 1.6e5 estimates and a numpy call per estimate was most of the score stage
    c_est, c_slot, c_row, c_vals, c_diam, c_valid = [], [], [], [], [], []
    for scene_id, scene_targets in ctx['targets_org'].items():
      scene_gt, scene_gt_info = ecs.scene_gt_and_info(ctx, scene_id)
      by_im_obj = {}
      for e in scene_errs_by_scene[scene_id]:
        by_im_obj.setdefault(e['im_id'], {}).setdefault(e['obj_id'], []).append(e)
      for im_id, im_targets in scene_targets.items():
        im_gts = scene_gt[im_id]
        valid = [bool(v) for v in
                 ecs.gt_valid_mask(im_gts, scene_gt_info[im_id], im_targets, p['visib_gt_min'])]
        row0 = len(row_gt)
        for gt_id, gt in enumerate(im_gts):
          row_scene.append(scene_id); row_im.append(im_id); row_obj.append(gt['obj_id'])
          row_gt.append(gt_id); row_valid.append(valid[gt_id])
        im_errs = by_im_obj.get(im_id)
        if not im_errs:
          continue
        for obj_id in set(gt['obj_id'] for gt in im_gts):
          if obj_id not in im_errs:
            continue
          errs = im_errs[obj_id]
          if len(errs) == 1:
            order = (0,)
          else:   # descending score, stable: what np.argsort(-scores, kind='stable') gives
            order = np.argsort(-np.array([e['score'] for e in errs], dtype=np.float64), kind='stable')
          if n_top > 0:
            order = order[:n_top]
          diameter = float(models_info[obj_id]['diameter']) if by_diam else 1.0
          for rank, k in enumerate(order):
            e = errs[k]
            i = len(est_rank)
            est_rank.append(rank); est_id.append(e['est_id']); est_score.append(e['score'])
            for j, (g, ev) in enumerate(e['errors'].items()):
              c_est.append(i); c_slot.append(j); c_row.append(row0 + g); c_vals.append(ev)
              c_diam.append(diameter); c_valid.append(valid[g])
    n_est = len(est_rank)
    try:
      vals = np.array(c_vals, dtype=np.float64).reshape(len(c_vals), -1)
    except ValueError:
      raise ValueError('ThresholdSweep needs error lists of one length')
    self.n_cols = vals.shape[1] if len(c_vals) else 1
    if by_diam:
      vals = vals / np.array(c_diam, dtype=np.float64)[:, None]
    if by_width:
      vals = factor * vals
    if len(c_vals):
      vals[~np.array(c_valid, dtype=bool)] = np.inf
      # match_poses compares with a strict '<' (pose_matching.py:62-70): a NaN error never beats
      # the best so far, i.e. such a candidate is skipped -- argmin would pick it first instead
      vals[np.isnan(vals)] = np.inf
    self.rows = dict(scene_id=row_scene, im_id=row_im, obj_id=row_obj, gt_id=row_gt, valid=row_valid)
    self.n_rows = len(row_gt)
    c_est = np.array(c_est, dtype=np.int64)
    c_slot = np.array(c_slot, dtype=np.int64)
    g_max = int(c_slot.max()) + 1 if len(c_slot) else 1
    # padded candidate tables; the pad row (index n_rows) is permanently taken
    self.cand_rows = np.full((n_est, g_max), self.n_rows, dtype=np.int64)
    self.cand_err = np.full((self.n_cols, n_est, g_max), np.inf, dtype=np.float64)
    if len(c_slot):
      self.cand_rows[c_est, c_slot] = np.array(c_row, dtype=np.int64)
      self.cand_err[:, c_est, c_slot] = vals.T
    self.est_id, self.est_score = est_id, est_score
    rank = np.array(est_rank, dtype=np.int64)
    self.by_rank = [np.flatnonzero(rank == r) for r in range(int(rank.max()) + 1 if n_est else 0)]
    # threshold-independent counters of score.count_targets_and_tps (score.py:77-96)
    insts = {}
    for o, s, im, v in zip(row_obj, row_scene, row_im, row_valid):
      if v:
        insts[(o, s, im)] = insts.get((o, s, im), 0) + 1
    self.obj_tars = {i: 0 for i in self.obj_ids}
    self.scene_tars = {i: 0 for i in self.scene_ids}
    for (o, s, _), c in insts.items():
      c = min(n_top, c) if n_top > 0 else c
      self.obj_tars[o] += c
      self.scene_tars[s] += c
    self._obj_index = {o: k for k, o in enumerate(self.obj_ids)}
    self._scene_index = {s: k for k, s in enumerate(self.scene_ids)}
    self._row_obj_k = np.array([self._obj_index.get(o, -1) for o in row_obj], dtype=np.int64)
    self._row_scene_k = np.array([self._scene_index.get(s, -1) for s in row_scene], dtype=np.int64)

  def run(self, thresholds, col=0, want_matches=True):
    """thresholds: T floats (one single-element correct_th each); col: which element of the
    error lists (the VSD tau).  -> list of T (scores, matches-or-None)."""
    ths = np.array([float(t) for t in thresholds], dtype=np.float64)
    T = len(ths)
    taken = np.zeros((T, self.n_rows + 1), dtype=bool)
    taken[:, self.n_rows] = True
    m_est = np.full((T, self.n_rows), -1, dtype=np.int64)
    m_err = np.zeros((T, self.n_rows), dtype=np.float64)
    t_ix = np.arange(T)[:, None]
    for idx in self.by_rank:
      rows = self.cand_rows[idx]                              # [n, G]
      err = self.cand_err[col][idx]                           # [n, G]
      e = np.where(taken[:, rows], np.inf, err[None])         # [T, n, G]
      j = np.argmin(e, axis=2)                                # first minimum, like the strict '<'
      v = np.take_along_axis(e, j[:, :, None], axis=2)[:, :, 0]
      ok = v < ths[:, None]
      tt, ii = np.nonzero(ok)
      rr = rows[ii, j[tt, ii]]
      taken[tt, rr] = True
      m_est[tt, rr] = idx[ii]
      m_err[tt, rr] = v[tt, ii]
    out = []
    n_obj, n_scene = len(self.obj_ids), len(self.scene_ids)
    for t in range(T):
      hit = m_est[t] >= 0
      if hit.any() and (self._row_obj_k[hit].min() < 0 or self._row_scene_k[hit].min() < 0):
        raise KeyError('a matched GT pose has an object / scene id outside the dataset tables')
      obj_tps = dict(zip(self.obj_ids, np.bincount(self._row_obj_k[hit], minlength=n_obj).tolist()))
      scene_tps = dict(zip(self.scene_ids, np.bincount(self._row_scene_k[hit], minlength=n_scene).tolist()))
      scores = score.scores_from_counts(self.obj_tars, self.scene_tars, obj_tps, scene_tps, self.n_rows)
      out.append((scores, self._matches(m_est[t], m_err[t], float(ths[t])) if want_matches else None))
    return out

  def _matches(self, m_est, m_err, th):
    r = self.rows
    out = []
    for k in range(self.n_rows):
      i = int(m_est[k])
      if i >= 0:
        e = float(m_err[k])
        out.append({'scene_id': r['scene_id'][k], 'im_id': r['im_id'][k], 'obj_id': r['obj_id'][k],
                    'gt_id': r['gt_id'][k], 'est_id': self.est_id[i], 'score': self.est_score[i],
                    'error': [e], 'error_norm': [e / th], 'valid': r['valid'][k]})
      else:
        out.append({'scene_id': r['scene_id'][k], 'im_id': r['im_id'][k], 'obj_id': r['obj_id'][k],
                    'gt_id': r['gt_id'][k], 'est_id': -1, 'score': -1, 'error': -1, 'error_norm': -1,
                    'valid': r['valid'][k]})
    return out

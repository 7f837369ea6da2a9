"""Greedy matching of pose estimates to ground-truth poses -- same contracts as
bop_toolkit_lib/pose_matching.py (match_poses :9-85, match_poses_scene :88-160).

The score stage is host code in the reference and stays host code here: it consumes a few
numbers per pair.  Results are bit-identical to the reference's because the
same comparisons run on the same float64 values in the same order.
"""


def match_poses(errs, error_ths, max_ests_count=0, gt_valid_mask=None):
  """Estimates are visited by decreasing score (stable); each takes the still-unmatched, valid
  GT whose error is element-wise below the best seen so far, starting from the thresholds
  (pose_matching.py:39-83).  Returns one dict per matched estimate."""
  ranked = sorted(errs, key=lambda e: e['score'], reverse=True)
  if max_ests_count > 0:
    ranked = ranked[:max_ests_count]
  ths = list(error_ths)
  n_el = len(ths)
  taken = []
  matches = []
  for e in ranked:
    best_gt, best_err = -1, ths
    for gt_id, err in e['errors'].items():
      if gt_valid_mask and not gt_valid_mask[gt_id]:
        continue
      if gt_id in taken:
        continue
      better = True
      for i in range(n_el):
        if not err[i] < best_err[i]:
          better = False
          break
      if better:
        best_gt, best_err = gt_id, err
    if best_gt >= 0:
      taken.append(best_gt)
      matches.append({
        'est_id': e['est_id'], 'gt_id': best_gt, 'score': e['score'], 'error': best_err,
        'error_norm': [best_err[i] / float(ths[i]) for i in range(n_el)]})
  return matches


def match_poses_scene(scene_id, scene_gt, scene_gt_valid, scene_errs, correct_th, n_top):
  """One entry per GT pose of the scene (in scene_gt order), filled with the matched estimate
  if any (pose_matching.py:111-160)."""
  by_im_obj = {}
  for e in scene_errs:
    by_im_obj.setdefault(e['im_id'], {}).setdefault(e['obj_id'], []).append(e)
  out = []
  for im_id, im_gts in scene_gt.items():
    rows = [{'scene_id': scene_id, 'im_id': im_id, 'obj_id': gt['obj_id'], 'gt_id': gt_id,
             'est_id': -1, 'score': -1, 'error': -1, 'error_norm': -1,
             'valid': scene_gt_valid[im_id][gt_id]} for gt_id, gt in enumerate(im_gts)]
    im_errs = by_im_obj.get(im_id)
    if im_errs:
      for obj_id in set(gt['obj_id'] for gt in im_gts):
        if obj_id not in im_errs:
          continue
        for m in match_poses(im_errs[obj_id], correct_th, n_top, scene_gt_valid[im_id]):
          g = rows[m['gt_id']]
          g['est_id'], g['score'] = m['est_id'], m['score']
          g['error'], g['error_norm'] = m['error'], m['error_norm']
    out += rows
  return out

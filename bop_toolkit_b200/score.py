"""Performance scores -- same contracts as bop_toolkit_lib/score.py (calc_ap :12-44,
calc_recall :47-58, calc_localization_scores :62-157).  Counting is integer work, so the
recall values are bit-identical to the reference's for the same matches."""
import numpy as np

from . import misc


def calc_ap(rec, pre):
  """Area under the monotonically decreasing precision/recall curve (PASCAL VOC 2010+,
  score.py:12-44)."""
  order = np.argsort(rec)
  mrec = np.concatenate(([0], np.array(rec)[order], [1]))
  mpre = np.concatenate(([0], np.array(pre)[order], [0]))
  assert mrec.shape == mpre.shape
  # running maximum from the right over mpre[0 .. size-2]
  head = np.maximum.accumulate(mpre[:-1][::-1])[::-1]
  mpre = np.concatenate((head, mpre[-1:]))
  idx = np.nonzero(mrec[1:] != mrec[:-1])[0] + 1
  return np.sum((mrec[idx] - mrec[idx - 1]) * mpre[idx])


def calc_recall(tp_count, targets_count):
  return 0.0 if targets_count == 0 else tp_count / float(targets_count)


def count_targets_and_tps(scene_ids, obj_ids, matches, n_top):
  """Integer counters behind the recall (score.py:77-110): per-(object, scene) targets =
  sum over images of min(n_top, #valid instances) (all instances when n_top <= 0), and true
  positives = valid GT poses with a matched estimate.  Returns (obj_tars, scene_tars, obj_tps,
  scene_tps) as dicts keyed like the reference's.  These counters are what a multi-GPU run
  all-reduces (sharding.allreduce_counts)."""
  insts = {}
  for m in matches:
    if m['valid']:
      key = (m['obj_id'], m['scene_id'], m['im_id'])
      insts[key] = insts.get(key, 0) + 1
  obj_tars = {i: 0 for i in obj_ids}
  scene_tars = {i: 0 for i in scene_ids}
  for (obj_id, scene_id, _), c in insts.items():
    c = min(n_top, c) if n_top > 0 else c
    obj_tars[obj_id] += c      # KeyError for ids outside obj_ids / scene_ids, as in the reference
    scene_tars[scene_id] += c
  obj_tps = {i: 0 for i in obj_ids}
  scene_tps = {i: 0 for i in scene_ids}
  for m in matches:
    if m['valid'] and m['est_id'] != -1:
      obj_tps[m['obj_id']] += 1
      scene_tps[m['scene_id']] += 1
  return obj_tars, scene_tars, obj_tps, scene_tps


def scores_from_counts(obj_tars, scene_tars, obj_tps, scene_tps, gt_count):
  tars, tps = sum(obj_tars.values()), sum(obj_tps.values())
  obj_recalls = {i: calc_recall(obj_tps[i], obj_tars[i]) for i in obj_tars}
  scene_recalls = {i: float(calc_recall(scene_tps[i], scene_tars[i])) for i in scene_tars}
  return {
    'recall': float(calc_recall(tps, tars)),
    'obj_recalls': obj_recalls,
    'mean_obj_recall': float(np.mean(list(obj_recalls.values())).squeeze()),
    'scene_recalls': scene_recalls,
    'mean_scene_recall': float(np.mean(list(scene_recalls.values())).squeeze()),
    'gt_count': int(gt_count),
    'targets_count': int(tars),
    'tp_count': int(tps),
  }


def calc_localization_scores(scene_ids, obj_ids, matches, n_top, do_print=True):
  """Scores of the 6D localization task (score.py:62-157)."""
  scores = scores_from_counts(*count_targets_and_tps(scene_ids, obj_ids, matches, n_top),
                              gt_count=len(matches))
  if do_print:
    misc.log('')
    misc.log('GT count:           {:d}'.format(scores['gt_count']))
    misc.log('Target count:       {:d}'.format(scores['targets_count']))
    misc.log('TP count:           {:d}'.format(scores['tp_count']))
    misc.log('Recall:             {:.4f}'.format(scores['recall']))
    misc.log('Mean object recall: {:.4f}'.format(scores['mean_obj_recall']))
    misc.log('Mean scene recall:  {:.4f}'.format(scores['mean_scene_recall']))
    misc.log('Object recalls:\n{}'.format(
      ', '.join('{}: {:.3f}'.format(i, s) for i, s in scores['obj_recalls'].items())))
    misc.log('Scene recalls:\n{}'.format(
      ', '.join('{}: {:.3f}'.format(i, s) for i, s in scores['scene_recalls'].items())))
    misc.log('')
  return scores

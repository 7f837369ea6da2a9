"""Calculates performance scores for 6D object localization -- drop-in for
scripts/eval_calc_scores.py of the reference: same flags (:22-131), reads the errors_*.json
files written by eval_calc_errors (ours or the reference's), writes the same
scores_<sign>.json / matches_<sign>.json next to them (:245-262).

`score_errors` is the in-process entry: it takes the error records of every scene (as returned
by eval_calc_errors.calc_errors) and needs no files, so a whole BOP19 evaluation -- errors for
three error types, then scores at every (tau x threshold) setting -- runs in one process instead
of the reference's 120+ subprocesses that re-read JSON (scripts/eval_bop19_pose.py:147-230).
"""
import argparse
import copy
import os
import time

from . import dataset_params, inout, misc, pose_matching, score

CORRECT_TH = {'vsd': [0.3], 'mssd': [0.2], 'mspd': [10], 'cus': [0.5], 'rete': [5.0, 5.0],
              're': [5.0], 'te': [5.0], 'proj': [5.0], 'ad': [0.1], 'add': [0.1], 'adi': [0.1]}


def default_params():
  return {
    'correct_th': copy.deepcopy(CORRECT_TH),
    'normalized_by_diameter': ['ad', 'add', 'adi', 'mssd'],
    'normalized_by_im_width': ['mspd'],
    'visib_gt_min': -1,
    'error_dir_paths': [r'/path/to/calculated/errors'],
    'eval_path': os.environ.get('BOP_EVAL_PATH', '/path/to/eval/folder'),
    'datasets_path': os.environ.get('BOP_PATH', '/path/to/bop/datasets'),
    'targets_filename': 'test_targets_bop19.json',
    'error_tpath': os.path.join('{eval_path}', '{error_dir_path}', 'errors_{scene_id:06d}.json'),
    'out_matches_tpath': os.path.join('{eval_path}', '{error_dir_path}', 'matches_{score_sign}.json'),
    'out_scores_tpath': os.path.join('{eval_path}', '{error_dir_path}', 'scores_{score_sign}.json'),
  }


def parse_args(argv=None):
  p = default_params()
  ap = argparse.ArgumentParser()
  for et in p['correct_th']:
    ap.add_argument('--correct_th_' + et, default=','.join(map(str, p['correct_th'][et])))
  ap.add_argument('--normalized_by_diameter', default=','.join(p['normalized_by_diameter']))
  ap.add_argument('--normalized_by_im_width', default=','.join(p['normalized_by_im_width']))
  ap.add_argument('--visib_gt_min', default=p['visib_gt_min'])
  ap.add_argument('--error_dir_paths', default=','.join(p['error_dir_paths']),
                  help='Comma-sep. paths to errors from eval_calc_errors.py.')
  for k in ('eval_path', 'datasets_path', 'targets_filename', 'error_tpath', 'out_matches_tpath',
            'out_scores_tpath'):
    ap.add_argument('--' + k, default=p[k])
  a = ap.parse_args(argv)
  for et in p['correct_th']:
    p['correct_th'][et] = list(map(float, getattr(a, 'correct_th_' + et).split(',')))
  p['normalized_by_diameter'] = a.normalized_by_diameter.split(',')
  p['normalized_by_im_width'] = a.normalized_by_im_width.split(',')
  p['visib_gt_min'] = float(a.visib_gt_min)
  p['error_dir_paths'] = a.error_dir_paths.split(',')
  for k in ('eval_path', 'datasets_path', 'targets_filename', 'error_tpath', 'out_matches_tpath',
            'out_scores_tpath'):
    p[k] = str(getattr(a, k))
  return p


def parse_error_dir(error_dir_path):
  """'<method>_<dataset>-<split>[-<type>]/error=<type>_ntop=<n>[...]' (eval_calc_scores.py:142-152)."""
  sign = os.path.basename(error_dir_path)
  err_type = str(sign.split('_')[0].split('=')[1])
  n_top = int(sign.split('_')[1].split('=')[1])
  info = os.path.basename(os.path.dirname(error_dir_path)).split('_')
  ds = info[1].split('-')
  return err_type, n_top, info[0], ds[0], ds[1], (ds[2] if len(ds) > 2 else None)


def gt_valid_mask(im_gt, im_gt_info, im_targets, visib_gt_min):
  """Which GT poses of an image count (eval_calc_scores.py:198-226): with visib_gt_min >= 0 the
  targets visible from at least that fraction; otherwise the `inst_count` most visible
  instances of every target object (stable sort by decreasing visib_fract)."""
  n = len(im_gt)
  if visib_gt_min >= 0:
    return [(im_gt[g]['obj_id'] in im_targets) and (im_gt_info[g]['visib_fract'] >= visib_gt_min)
            for g in range(n)]
  valid = [False] * n
  left = {obj_id: t['inst_count'] for obj_id, t in im_targets.items()}
  for g in sorted(range(n), key=lambda g: im_gt_info[g]['visib_fract'], reverse=True):
    obj_id = im_gt[g]['obj_id']
    if left.get(obj_id, 0) > 0:
      valid[g] = True
      left[obj_id] -= 1
  return valid


def normalize_errors(scene_errs, err_type, p, models_info, im_width):
  """Per-object diameter / image-width normalisation before thresholding (:235-247).  Same
  arithmetic as the reference (e / diameter, then factor * e) but into NEW records: the input
  (e.g. the records calc_errors returned) is left untouched and nothing is deep-copied."""
  by_diam = err_type in p['normalized_by_diameter']
  by_width = err_type in p['normalized_by_im_width']
  if not (by_diam or by_width):
    return scene_errs
  factor = 640.0 / float(im_width)
  out = []
  for err in scene_errs:
    errors = err['errors']
    if by_diam:
      diameter = float(models_info[err['obj_id']]['diameter'])
      errors = {g: [e / diameter for e in es] for g, es in errors.items()}
    if by_width:
      errors = {g: [factor * e for e in es] for g, es in errors.items()}
    rec = dict(err)
    rec['errors'] = errors
    out.append(rec)
  return out


def load_context(p, dataset, split, split_type, ctx):
  """Fills the cache dict `ctx` with what every score setting of a dataset split shares: dataset
  tables, models info, the targets organised by scene / image / object."""
  key = (dataset, split, split_type, p['targets_filename'], p['datasets_path'])
  if ctx.get('key') != key:
    ctx.clear()
    ctx['key'] = key
    ctx['dp_split'] = dataset_params.get_split_params(p['datasets_path'], dataset, split, split_type)
    ctx['dp_model'] = dataset_params.get_model_params(p['datasets_path'], dataset, 'eval')
    ctx['models_info'] = inout.load_json(ctx['dp_model']['models_info_path'], keys_to_int=True)
    targets = inout.load_json(os.path.join(ctx['dp_split']['base_path'], p['targets_filename']))
    torg = {}
    for t in targets:
      torg.setdefault(t['scene_id'], {}).setdefault(t['im_id'], {})[t['obj_id']] = t
    ctx['targets_org'] = torg
    ctx['scene_gt'], ctx['scene_gt_info'] = {}, {}
  return ctx


def scene_gt_and_info(ctx, scene_id):
  if scene_id not in ctx['scene_gt']:
    dp_split = ctx['dp_split']
    # scene_gt through the error stage's parse cache: a one-process evaluation reads it once
    from .eval_calc_errors import _memo
    ctx['scene_gt'][scene_id] = _memo(inout.load_scene_gt, dp_split['scene_gt_tpath'].format(scene_id=scene_id))
    ctx['scene_gt_info'][scene_id] = inout.load_json_scene(
      dp_split['scene_gt_info_tpath'].format(scene_id=scene_id))
  return ctx['scene_gt'][scene_id], ctx['scene_gt_info'][scene_id]


def score_errors(p, err_type, n_top, dataset, split, split_type, scene_errs_by_scene,
                 do_print=True, ctx=None):
  """Matches + scores for one error setting from in-memory records.
  scene_errs_by_scene: {scene_id: records} (never mutated).
  ctx: optional cache dict shared between calls (targets, GT, GT info, models info).
  Returns (scores, matches).  (score_sweep.ThresholdSweep does many thresholds in one pass.)"""
  ctx = load_context(p, dataset, split, split_type, {} if ctx is None else ctx)
  dp_split, dp_model = ctx['dp_split'], ctx['dp_model']
  matches = []
  for scene_id, scene_targets in ctx['targets_org'].items():
    scene_gt, scene_gt_info = scene_gt_and_info(ctx, scene_id)
    gt_curr, valid = {}, {}
    for im_id, im_targets in scene_targets.items():
      gt_curr[im_id] = scene_gt[im_id]
      valid[im_id] = gt_valid_mask(scene_gt[im_id], scene_gt_info[im_id], im_targets, p['visib_gt_min'])
    errs = normalize_errors(scene_errs_by_scene[scene_id], err_type, p, ctx['models_info'],
                            dp_split['im_size'][0])
    matches += pose_matching.match_poses_scene(scene_id, gt_curr, valid, errs,
                                               p['correct_th'][err_type], n_top)
  scores = score.calc_localization_scores(dp_split['scene_ids'], dp_model['obj_ids'],
                                          matches, n_top, do_print)
  return scores, matches


def calc_scores(p, log=misc.log):
  """File-based driver loop of eval_calc_scores.py:136-266."""
  out = {}
  ctx = {}
  for error_dir_path in p['error_dir_paths']:
    log('Processing: {}'.format(error_dir_path))
    t0 = time.time()
    err_type, n_top, method, dataset, split, split_type = parse_error_dir(error_dir_path)
    score_sign = misc.get_score_signature(p['correct_th'][err_type], p['visib_gt_min'])
    log('Calculating score - error: {}, method: {}, dataset: {}.'.format(err_type, method, dataset))
    targets = inout.load_json(os.path.join(
      dataset_params.get_split_params(p['datasets_path'], dataset, split, split_type)['base_path'],
      p['targets_filename']))
    scene_ids = []
    for t in targets:
      if t['scene_id'] not in scene_ids:
        scene_ids.append(t['scene_id'])
    errs = {s: inout.load_json(p['error_tpath'].format(
      eval_path=p['eval_path'], error_dir_path=error_dir_path, scene_id=s), keys_to_int=True)
      for s in scene_ids}
    scores, matches = score_errors(p, err_type, n_top, dataset, split, split_type, errs,
                                   do_print=log is misc.log, ctx=ctx)
    fmt = dict(eval_path=p['eval_path'], error_dir_path=error_dir_path, score_sign=score_sign)
    inout.save_json(p['out_scores_tpath'].format(**fmt), scores)
    inout.save_json(p['out_matches_tpath'].format(**fmt), matches)
    out[error_dir_path] = scores
    log('Matching and score calculation took {}s.'.format(time.time() - t0))
  log('Done.')
  return out


def main(argv=None):
  p = parse_args(argv)
  misc.log('-----------')
  misc.log('Parameters:')
  for k, v in p.items():
    misc.log('- {}: {}'.format(k, v))
  misc.log('-----------')
  calc_scores(p)


if __name__ == '__main__':
  main()

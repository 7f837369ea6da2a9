"""Score stage (pose_matching / score / eval_calc_scores), dataset tables, GT info and GT masks
against outputs of the REAL reference scripts on the synthetic dataset
(tests/golden/driver_golden.json + gt_masks_golden.npz, made by tests/golden/make_golden_driver.py).

CPU: fed the reference's own error files and scene_gt_info.json, our eval_calc_scores writes
scores_*.json and matches_*.json that are byte-identical to the reference's (BASELINE.json:
"score.py recall is bit-identical when fed the reference's error files").
GPU: our errors -> our scores within 0.1 pp of the reference's recall; calc_gt_info /
calc_gt_masks against the reference's files.
"""
import json
import os

import numpy as np
import pytest

from test_driver import dataset, driver_golden, _golden_files, _params  # noqa: F401 (fixtures)


def _install_gt_info(info, driver_golden):
  base = os.path.join(info['datasets_path'], 'lm', 'test')
  for scene_id, text in driver_golden['gt_info'].items():
    with open(os.path.join(base, '%06d' % int(scene_id), 'scene_gt_info.json'), 'w') as f:
      f.write(text)


def test_dataset_tables_match_reference(driver_golden):
  from bop_toolkit_b200 import dataset_params as dp
  n = 0
  for key, ref in driver_golden['tables'].items():
    parts = key.split('|')
    if parts[0] == 'model':
      got = dp.get_model_params('/d', parts[1], 'eval')
    else:
      got = dp.get_split_params('/d', parts[1], parts[2], None if parts[3] == 'None' else parts[3])
      got = {k: (list(v) if isinstance(v, tuple) else v) for k, v in got.items()}
    for k, v in ref.items():
      assert got[k] == v, (key, k, got[k], v)
      n += 1
  assert n > 2000


def test_match_poses_semantics():
  from bop_toolkit_b200 import pose_matching as pm
  errs = [{'est_id': 0, 'score': 0.9, 'errors': {0: [0.1], 1: [0.05]}},
          {'est_id': 1, 'score': 0.9, 'errors': {0: [0.2], 1: [0.01]}},     # tie: CSV order wins
          {'est_id': 2, 'score': 0.5, 'errors': {0: [0.25], 1: [0.0]}}]
  m = pm.match_poses(errs, [0.3])
  assert [(x['est_id'], x['gt_id']) for x in m] == [(0, 1), (1, 0)]
  assert m[0]['error_norm'] == [0.05 / 0.3]
  assert pm.match_poses(errs, [0.3], max_ests_count=1)[0]['est_id'] == 0
  assert [(x['est_id'], x['gt_id']) for x in pm.match_poses(errs, [0.3], 0, [True, False])] == [(0, 0)]
  # strict '<' against the threshold: est 1 (0.01 vs 0.01) does not match, est 2 (0.0) does
  assert [(x['est_id'], x['gt_id']) for x in pm.match_poses(errs, [0.01])] == [(2, 1)]
  two = [{'est_id': 0, 'score': 1.0, 'errors': {0: [4.0, 6.0], 1: [4.9, 4.9]}}]
  assert pm.match_poses(two, [5.0, 5.0])[0]['gt_id'] == 1


def test_calc_ap_and_recall():
  from bop_toolkit_b200 import score
  tp = np.array([False, True, True, False, True, False])
  tp_c, fp_c = np.cumsum(tp).astype(float), np.cumsum(~tp).astype(float)
  ap = score.calc_ap(tp_c / tp.size, tp_c / (fp_c + tp_c))
  assert abs(ap - (1 / 6 * 2 / 3 + 1 / 6 * 2 / 3 + 1 / 6 * 0.6)) < 1e-15   # score.py:160-168 example
  assert score.calc_recall(0, 0) == 0.0 and score.calc_recall(1, 4) == 0.25


def test_scores_bit_identical_on_reference_error_files(dataset, driver_golden, tmp_path):
  from bop_toolkit_b200 import eval_calc_scores as ecs
  _install_gt_info(dataset, driver_golden)
  n_files = 0
  for error_type, n_top in driver_golden['score_runs']:
    eval_path = str(tmp_path / ('eval_%s_%d' % (error_type, n_top)))
    for rel, text in _golden_files(driver_golden, error_type, n_top).items():
      os.makedirs(os.path.dirname(os.path.join(eval_path, rel)), exist_ok=True)
      with open(os.path.join(eval_path, rel), 'w') as f:
        f.write(text)
    for si, extra in enumerate(driver_golden['score_settings']):
      pre = '%s|%d|%d|' % (error_type, n_top, si)
      gold = {k[len(pre):]: v for k, v in driver_golden['scores'].items() if k.startswith(pre)}
      assert gold
      dirs = sorted(set(os.path.dirname(rel) for rel in gold))
      p = ecs.parse_args(['--error_dir_paths=' + ','.join(dirs), '--eval_path=' + eval_path,
                          '--datasets_path=' + dataset['datasets_path']] + extra)
      ecs.calc_scores(p, log=lambda s: None)
      for rel, text in gold.items():
        assert open(os.path.join(eval_path, rel)).read() == text, (error_type, n_top, si, rel)
        n_files += 1
  assert n_files == len(driver_golden['scores'])


@pytest.mark.gpu
def test_recall_from_our_errors_within_0p1pp(dataset, driver_golden, tmp_path):
  """Errors by our CUDA driver, scores by our score stage (in-process, no files), against the
  reference's scores on the reference's errors."""
  from bop_toolkit_b200 import eval_calc_errors as ece, eval_calc_scores as ecs, misc
  _install_gt_info(dataset, driver_golden)
  checked = 0
  for error_type, n_top in driver_golden['score_runs']:
    p = _params(dataset, error_type, n_top, str(tmp_path / 'unused'))
    records = ece.calc_errors(p, log=lambda s: None, write=False)['synth_lm-test']
    ctx = {}
    for si, extra in enumerate(driver_golden['score_settings']):
      ps = ecs.parse_args(['--datasets_path=' + dataset['datasets_path']] + extra)
      sign = misc.get_score_signature(ps['correct_th'][error_type], ps['visib_gt_min'])
      taus = p['vsd_taus'] if error_type == 'vsd' else [None]
      for tau_id, tau in enumerate(taus):
        esign = misc.get_error_signature(error_type, n_top, vsd_delta=15.0, vsd_tau=tau) \
          if error_type == 'vsd' else misc.get_error_signature(error_type, n_top)
        key = '%s|%d|%d|synth_lm-test/%s/scores_%s.json' % (error_type, n_top, si, esign, sign)
        if key not in driver_golden['scores']:
          continue
        recs = {s: (ece.records_for_tau(r, tau_id) if error_type == 'vsd' else r) for s, r in records.items()}
        scores, _ = ecs.score_errors(ps, error_type, n_top, 'lm', 'test', None, recs, do_print=False, ctx=ctx)
        ref = json.loads(driver_golden['scores'][key])
        assert abs(scores['recall'] - ref['recall']) <= 1e-3, (key, scores['recall'], ref['recall'])
        assert scores['targets_count'] == ref['targets_count'] and scores['gt_count'] == ref['gt_count']
        checked += 1
  assert checked >= 20


@pytest.mark.gpu
def test_calc_gt_info_matches_reference(dataset, driver_golden):
  from bop_toolkit_b200 import calc_gt_info as cgi
  p = cgi.parse_args(['--dataset=lm', '--dataset_split=test', '--datasets_path=' + dataset['datasets_path']])
  ours = cgi.calc_gt_info(p, log=lambda s: None, write=False)
  exact = total = 0
  for scene_id, text in driver_golden['gt_info'].items():
    ref = json.loads(text)
    got = ours[int(scene_id)]
    assert sorted(int(k) for k in ref) == sorted(got)
    for im_id, infos in ref.items():
      assert len(infos) == len(got[int(im_id)])
      for a, b in zip(got[int(im_id)], infos):
        total += 1
        exact += int(a == b)
        for k in ('px_count_all', 'px_count_valid', 'px_count_visib'):
          assert abs(a[k] - b[k]) <= 3, (scene_id, im_id, k, a, b)        # silhouette-edge budget
        assert abs(a['visib_fract'] - b['visib_fract']) <= 2e-3
        assert np.all(np.abs(np.array(a['bbox_obj']) - np.array(b['bbox_obj'])) <= 1)
        assert np.all(np.abs(np.array(a['bbox_visib']) - np.array(b['bbox_visib'])) <= 1)
  print('gt info records identical to the reference: %d of %d' % (exact, total))
  assert exact >= 0.9 * total


@pytest.mark.gpu
def test_calc_gt_masks_match_reference(dataset, driver_golden):
  import cv2
  from bop_toolkit_b200 import calc_gt_info as cgi
  gold = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'gt_masks_golden.npz'))
  p = cgi.parse_args(['--dataset=lm', '--dataset_split=test', '--datasets_path=' + dataset['datasets_path']])
  assert cgi.calc_gt_masks(p, log=lambda s: None) > 0
  scene_dir = os.path.join(dataset['datasets_path'], 'lm', 'test', '%06d' % int(gold['scene_id']))
  n = bad = 0
  for key in gold.files:
    if key == 'scene_id':
      continue
    m = cv2.imread(os.path.join(scene_dir, key), cv2.IMREAD_UNCHANGED)
    assert m is not None and m.dtype == np.uint8 and m.shape == (480, 640)
    assert set(np.unique(m)) <= {0, 255}
    diff = int(np.count_nonzero(np.packbits(m > 0) != gold[key]))
    bad += diff
    n += 1
    assert diff <= 3, (key, diff)
  print('gt masks: %d files, %d differing bytes in total' % (n, bad))
  assert n == 30


@pytest.mark.gpu
def test_eval_bop19_pose_matches_reference(dataset, driver_golden, tmp_path):
  """The whole BOP19 evaluation in one process (3 error types, 120 score settings) against the
  files the reference's eval_bop19_pose.py + 123 child scripts left behind."""
  import hashlib
  from bop_toolkit_b200 import eval_bop19_pose as e19
  _install_gt_info(dataset, driver_golden)
  eval_path = str(tmp_path / 'eval_bop19')
  p = e19.parse_args(['--result_filenames=' + dataset['result_filename'],
                      '--results_path=' + dataset['results_path'], '--eval_path=' + eval_path,
                      '--datasets_path=' + dataset['datasets_path']])
  final = e19.evaluate(p, log=lambda s: None)['synth_lm-test']
  gold = driver_golden['bop19']
  ref_final = json.loads(gold['scores']['synth_lm-test/scores_bop19.json'])
  for k, v in ref_final.items():
    assert abs(float(final[k]) - v) <= 1e-3, (k, final[k], v)          # 0.1 pp
  same_scores = same_matches = 0
  for rel, text in gold['scores'].items():
    ours = open(os.path.join(eval_path, rel)).read()
    same_scores += int(ours == text)
    a, b = json.loads(ours), json.loads(text)
    assert set(a) == set(b)
    if 'recall' in b:
      assert abs(a['recall'] - b['recall']) <= 1e-3, rel
      assert a['targets_count'] == b['targets_count'] and a['gt_count'] == b['gt_count']
  for rel, sha in gold['matches_sha256'].items():
    ours = open(os.path.join(eval_path, rel)).read()
    same_matches += int(hashlib.sha256(ours.encode()).hexdigest() == sha)
  n_err = sum(1 for d, _, names in os.walk(eval_path) for n in names if n.startswith('errors_'))
  assert n_err == gold['n_error_files']
  print('bop19: %d of %d score files and %d of %d match files byte-identical to the reference; AR %s'
        % (same_scores, len(gold['scores']), same_matches, len(gold['matches_sha256']),
           {k: round(float(v), 4) for k, v in final.items()}))


def test_threshold_sweep_equals_per_setting_scoring(dataset, driver_golden):
  """score_sweep.ThresholdSweep (all thresholds of an error type in one vectorised pass) against
  eval_calc_scores.score_errors on the REFERENCE's error files: identical scores and matches for
  every (error type, n_top, tau, threshold, visib_gt_min) combination tried, including
  thresholds equal to actual error values (strict '<') and an n_top cut."""
  import copy
  from bop_toolkit_b200 import eval_calc_scores as ecs, inout, score_sweep
  _install_gt_info(dataset, driver_golden)
  n_cmp = 0
  for error_type, n_top in driver_golden['score_runs']:
    files = _golden_files(driver_golden, error_type, n_top)
    by_dir = {}
    for rel, text in files.items():
      scene_id = int(os.path.basename(rel).split('_')[1].split('.')[0])
      recs = json.loads(text, parse_constant=lambda c: {'Infinity': float('inf'), '-Infinity': -float('inf'),
                                                         'NaN': float('nan')}[c])
      for r in recs:
        r['errors'] = {int(k): v for k, v in r['errors'].items()}
      by_dir.setdefault(os.path.dirname(rel), {})[scene_id] = recs
    for error_dir, scene_errs in by_dir.items():
      if any(len(v) != 1 for recs in scene_errs.values() for r in recs for v in r['errors'].values()):
        continue                                           # multi-element errors: not the sweep's job
      vals = sorted({v[0] for recs in scene_errs.values() for r in recs for v in r['errors'].values()
                     if np.isfinite(v[0])})
      for visib in (0.1, -1.0, 0.6):
        ps = ecs.parse_args(['--datasets_path=' + dataset['datasets_path'], '--visib_gt_min=%s' % visib])
        base = {'vsd': [0.05, 0.3, 0.5], 'mssd': [0.05, 0.3], 'mspd': [5.0, 20.0]}.get(error_type, [0.1, 10.0])
        diam_like = [v for v in vals[::max(1, len(vals) // 4)]][:4]     # thresholds ON error values
        ths = base + [float(v) for v in diam_like] + [1e9, 0.0]
        ctx = {}
        sweep = score_sweep.ThresholdSweep(ps, error_type, n_top, 'lm', 'test', None, scene_errs, ctx)
        fast = sweep.run(ths)
        for th, (sc_f, m_f) in zip(ths, fast):
          p1 = copy.deepcopy(ps)
          p1['correct_th'][error_type] = [th]
          sc_s, m_s = ecs.score_errors(p1, error_type, n_top, 'lm', 'test', None, scene_errs, do_print=False, ctx=ctx)
          assert inout.dumps_errors_layout(m_f) == inout.dumps_errors_layout(m_s), (error_type, n_top, th, visib)
          assert json.dumps(sc_f, sort_keys=True) == json.dumps(sc_s, sort_keys=True), (error_type, n_top, th)
          n_cmp += 1
  assert n_cmp >= 100

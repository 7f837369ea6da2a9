"""Driver (eval_calc_errors drop-in) against the error files written by the REAL reference
script on the same synthetic BOP-layout dataset (tests/golden/driver_golden.json, produced by
tests/golden/make_golden_driver.py).

CPU part: the dataset regenerates byte-identically; the flattened record skeleton (order,
im_id / obj_id / est_id / score, GT ids) equals the reference's for every n_top mode.
GPU part: the files our driver writes carry the same records with errors within tolerance
(1e-5 relative for the point errors, 1e-3 absolute per tau for VSD, bars of BASELINE.json).
"""
import hashlib
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _oracle_render_fn(pts, faces, R, t, K):
  from oracle import renderer_oracle
  from bop_toolkit_b200 import synth
  return renderer_oracle.render_depth(pts, faces, R, t, K[0, 0], K[1, 1], K[0, 2], K[1, 2],
                                      synth.IM_H, synth.IM_W)


def _checksum(info):
  import cv2
  h = hashlib.sha256()
  base = os.path.join(info['datasets_path'], 'lm')
  for scene_id in info['scene_ids']:
    ddir = os.path.join(base, 'test', '%06d' % scene_id, 'depth')
    for name in sorted(os.listdir(ddir)):
      h.update(cv2.imread(os.path.join(ddir, name), cv2.IMREAD_UNCHANGED).tobytes())
  for path in (os.path.join(info['results_path'], info['result_filename']),
               os.path.join(base, 'test_targets_bop19.json')):
    h.update(open(path, 'rb').read())
  return h.hexdigest()


@pytest.fixture(scope='module')
def driver_golden():
  return json.load(open(os.path.join(ROOT, 'tests', 'golden', 'driver_golden.json')))


@pytest.fixture(scope='module')
def dataset(tmp_path_factory, driver_golden, oracle_built):
  from bop_toolkit_b200 import synth
  root = str(tmp_path_factory.mktemp('bop_ds'))
  kw = dict(driver_golden['dataset_kw'])
  kw['scene_ids'], kw['obj_ids'] = tuple(kw['scene_ids']), tuple(kw['obj_ids'])
  info = synth.write_bop_dataset(root, _oracle_render_fn, **kw)
  assert _checksum(info) == driver_golden['dataset_sha256'], \
    'the synthetic dataset did not regenerate byte-identically; goldens do not apply'
  return info


def _params(info, error_type, n_top, eval_path):
  from bop_toolkit_b200 import eval_calc_errors as ece
  return ece.parse_args([
    '--error_type=' + error_type, '--n_top=%d' % n_top,
    '--result_filenames=' + info['result_filename'], '--results_path=' + info['results_path'],
    '--eval_path=' + eval_path, '--datasets_path=' + info['datasets_path']])


def _golden_files(driver_golden, error_type, n_top):
  pre = '%s|%d|' % (error_type, n_top)
  return {k[len(pre):]: v for k, v in driver_golden['files'].items() if k.startswith(pre)}


def _load_text(text):
  return json.loads(text, parse_constant=lambda c: {'Infinity': float('inf'), '-Infinity': -float('inf'),
                                                     'NaN': float('nan')}[c])


def test_parse_args_quirks():
  from bop_toolkit_b200 import eval_calc_errors as ece
  p = ece.parse_args(['--skip_missing=False', '--vsd_normalized_by_diameter=', '--n_top=-1',
                      '--vsd_deltas=lmo:7.5,xx:3', '--vsd_taus=0.1,0.3'])
  assert p['skip_missing'] is True            # bool('False') -- eval_calc_errors.py:120
  assert p['vsd_normalized_by_diameter'] is False
  assert p['n_top'] == -1 and p['vsd_deltas'] == {'lmo': 7.5, 'xx': 3.0} and p['vsd_taus'] == [0.1, 0.3]
  d = ece.default_params()
  assert d['vsd_taus'][2] == 0.15000000000000002 and len(d['vsd_taus']) == 10
  assert ece.parse_result_filename('a/b/meth-x_lmo-test-pbr.csv') == ('meth-x_lmo-test-pbr', 'meth-x', 'lmo', 'test', 'pbr')


@pytest.mark.parametrize('n_top', [1, -1, 0])
def test_flatten_matches_reference_records(dataset, driver_golden, n_top):
  """Host logic only: record order, est_id, scores and GT ids equal the reference's."""
  from bop_toolkit_b200 import dataset_params, eval_calc_errors as ece, inout
  et = {1: 'mssd', -1: 'mspd', 0: 'mssd'}[n_top]
  gold = _golden_files(driver_golden, et, n_top)
  assert gold
  p = _params(dataset, et, n_top, '/nonexistent')
  dp = dataset_params.get_split_params(p['datasets_path'], 'lm', 'test')
  targets = inout.load_json(os.path.join(dp['base_path'], p['targets_filename']))
  torg, eorg = {}, {}
  for t in targets:
    torg.setdefault(t['scene_id'], {}).setdefault(t['im_id'], {})[t['obj_id']] = t
  for e in inout.load_bop_results(os.path.join(p['results_path'], p['result_filenames'][0])):
    eorg.setdefault(e['scene_id'], {}).setdefault(e['im_id'], {}).setdefault(e['obj_id'], []).append(e)
  for scene_id, st in torg.items():
    gt = inout.load_scene_gt(dp['scene_gt_tpath'].format(scene_id=scene_id))
    sp = ece.flatten_scene(p, scene_id, st, eorg.get(scene_id, {}), gt)
    ref = _load_text(gold['synth_lm-test/error=%s_ntop=%d/errors_%06d.json' % (et, n_top, scene_id)])
    assert len(ref) == len(sp.records)
    for a, b in zip(sp.records, ref):
      assert (a['im_id'], a['obj_id'], a['est_id'], a['score']) == (b['im_id'], b['obj_id'], b['est_id'], b['score'])
      assert [str(g) for g in a['errors']] == list(b['errors'])
    assert len(sp.rec) == sum(len(r['errors']) for r in ref)


def test_writer_layout_roundtrip(driver_golden):
  """inout.save_json reproduces the reference's text byte for byte from parsed records."""
  from bop_toolkit_b200 import inout
  n = 0
  for key, text in driver_golden['files'].items():
    recs = _load_text(text)
    for r in recs:
      r['errors'] = {int(k): v for k, v in r['errors'].items()}
    assert inout.dumps_errors_layout(recs) == text, key
    assert inout.dumps_error_records(recs) == text, key       # the driver's fast writer
    n += 1
  assert n >= 60


def test_fast_writer_equals_json_dumps():
  """inout.dumps_error_records == the json.dumps-based layout, also for per-tau views, several GT
  instances per record, records without GT, inf / nan / tiny / huge / integer-valued errors."""
  from bop_toolkit_b200 import eval_calc_errors as ece, inout
  rng = np.random.default_rng(0)
  special = [float('inf'), -float('inf'), float('nan'), 0.0, -0.0, 1.0, 1e16, 1e-5, 1.5e-7, 123456789012345678.0,
             0.1 + 0.2, 5e-324, 1.7976931348623157e308, 1e22, 1e15, 9999999999999998.0]
  recs = []
  for i in range(400):
    n_gt = int(rng.integers(0, 4))
    gts = sorted(rng.choice(30, n_gt, replace=False).tolist(), reverse=bool(i & 1))
    errs = {}
    for g in gts:
      e = (rng.random(10) * 10.0 ** rng.integers(-8, 8, 10)).tolist()
      if i % 7 == 0:
        e[int(rng.integers(0, 10))] = special[int(rng.integers(0, len(special)))]
      errs[int(g)] = e
    recs.append({'im_id': int(rng.integers(0, 2000)), 'obj_id': int(rng.integers(1, 30)), 'est_id': i % 5,
                 'score': [float(rng.random()), 1.0, -1.0, 3][i % 4], 'errors': errs})
  for e in special:
    recs.append({'im_id': 1, 'obj_id': 2, 'est_id': 0, 'score': e, 'errors': {4: [e] * 10}})
  assert inout.dumps_error_records(recs) == inout.dumps_errors_layout(recs)
  tails = inout.error_record_tails(recs)
  for tau_id in (0, 3, 9):
    assert inout.dumps_error_records(recs, tau_id, tails) == \
      inout.dumps_errors_layout(ece.records_for_tau(recs, tau_id))
  assert inout.dumps_error_records([]) == inout.dumps_errors_layout([])


TOL = {'vsd': (0.0, 1e-3), 'cus': (0.0, 2e-3), 'adi': (1e-5, 1e-6), 'ad': (1e-5, 1e-6), 're': (1e-5, 1e-6), 'rete': (1e-5, 1e-6)}


@pytest.mark.gpu
def test_driver_files_match_reference(dataset, driver_golden, tmp_path):
  from bop_toolkit_b200 import eval_calc_errors as ece
  identical = total = 0
  for error_type, n_top in driver_golden['runs']:
    gold = _golden_files(driver_golden, error_type, n_top)
    eval_path = str(tmp_path / ('eval_%s_%d' % (error_type, n_top)))
    ece.calc_errors(_params(dataset, error_type, n_top, eval_path), log=lambda s: None)
    ours = {}
    for dirpath, _, names in os.walk(eval_path):
      for name in names:
        full = os.path.join(dirpath, name)
        ours[os.path.relpath(full, eval_path)] = open(full).read()
    assert sorted(ours) == sorted(gold), (error_type, n_top)
    rtol, atol = TOL.get(error_type, (1e-5, 1e-9))
    for rel, text in gold.items():
      total += 1
      identical += int(ours[rel] == text)
      a, b = _load_text(ours[rel]), _load_text(text)
      assert len(a) == len(b)
      for ra, rb in zip(a, b):
        assert {k: ra[k] for k in ('im_id', 'obj_id', 'est_id', 'score')} == \
               {k: rb[k] for k in ('im_id', 'obj_id', 'est_id', 'score')}
        assert list(ra['errors']) == list(rb['errors'])
        for g in rb['errors']:
          ea, eb = np.array(ra['errors'][g], float), np.array(rb['errors'][g], float)
          assert ea.shape == eb.shape
          fin = np.isfinite(eb)
          assert np.array_equal(np.isfinite(ea), fin) and np.array_equal(ea[~fin], eb[~fin])
          assert np.all(np.abs(ea[fin] - eb[fin]) <= atol + rtol * np.abs(eb[fin])), \
            (error_type, n_top, rel, ra, rb)
  print('driver files: %d of %d byte-identical to the reference' % (identical, total))


def test_native_error_file_writer_equals_python(tmp_path):
  """bop_write_error_file_host (flat arrays -> file, no GIL) writes the bytes of the json.dumps
  layout: VSD tau files, single-value files, two-element lists ('rete'), records without GT and
  with several GT instances, inf / nan / tiny / huge values; odd records fall back to Python."""
  from bop_toolkit_b200 import _lib, eval_calc_errors as ece, misc
  _lib.build()
  rng = np.random.default_rng(4)
  special = [float('inf'), float('nan'), 0.0, 1.0, 1e16, 1e-5, 1.5e-7, 0.1 + 0.2, 5e-324, 1e22, 1e15, 123456.0]
  sp = ece.ScenePairs()
  rec, gt = [], []
  for i in range(500):
    n_gt = int(rng.integers(0, 4))
    gts = sorted(rng.choice(30, n_gt, replace=False).tolist())
    sp.records.append({'im_id': int(rng.integers(0, 2000)), 'obj_id': int(rng.integers(1, 30)),
                       'est_id': i % 5, 'score': float(rng.random()) if i % 9 else -1.0,
                       'errors': {int(g): None for g in gts}})
    rec += [i] * n_gt
    gt += gts
  sp.rec, sp.gt_id = np.array(rec, np.int64), np.array(gt, np.int64)
  for et, T in (('vsd', 10), ('mssd', 1), ('rete', 2)):
    errs = rng.random((len(rec), T)) * 10.0 ** rng.integers(-8, 8, (len(rec), T))
    errs[rng.random(errs.shape) < 0.05] = rng.choice(special, int((rng.random(errs.shape) < 0.05).sum()) or 1)[0]
    for k, v in enumerate(special):
      errs[k % len(rec), k % T] = v
    records = ece.fill_records(sp, errs)
    outs = {}
    for mode in ('native', 'python'):
      p = ece.default_params()
      p.update({'error_type': et, 'n_top': -1, 'eval_path': str(tmp_path / (mode + et)), 'vsd_taus': list(p['vsd_taus'])})
      kw = dict(sp=sp, errs=errs) if mode == 'native' else {}
      ece.save_scene_errors(p, 'lm', 'res', 7, records, log=lambda s: None, **kw)
      files = {}
      for dirpath, _, names in os.walk(p['eval_path']):
        for name in names:
          files[os.path.relpath(os.path.join(dirpath, name), p['eval_path'])] = \
            open(os.path.join(dirpath, name), 'rb').read()
      outs[mode] = files
    assert len(outs['native']) == (10 if et == 'vsd' else 1)
    assert outs['native'] == outs['python'], et
  assert ece._record_arrays(sp, errs) is not None
  sp.records[3]['score'] = 1          # an int score prints as `1`: not the native writer's job
  assert ece._record_arrays(sp, errs) is None


def test_native_results_reader_equals_reference_parse(tmp_path):
  """inout.load_bop_results through bop_parse_results_csv_host == the reference's line-by-line
  parse (types, shapes, values bit for bit); irregular files take the Python path and raise the
  reference's errors."""
  from bop_toolkit_b200 import _lib, inout
  _lib.build()
  rng = np.random.default_rng(2)
  res = [{'scene_id': int(rng.integers(1, 60)), 'im_id': int(rng.integers(0, 3000)), 'obj_id': int(rng.integers(1, 31)),
          'score': float(rng.random()) if i % 5 else 1.0, 'R': rng.normal(size=(3, 3)),
          't': rng.normal(size=(3, 1)) * 10.0 ** rng.integers(-3, 5), 'time': [0.25, -1, 1e-5][i % 3]}
         for i in range(3000)]
  path = str(tmp_path / 'm_lm-test.csv')
  inout.save_bop_results(path, res)
  a, b = inout.load_bop_results(path), inout._load_bop_results_py(path)
  assert len(a) == len(b) == 3000
  for x, y in zip(a, b):
    assert list(x) == list(y)
    for k in ('scene_id', 'im_id', 'obj_id', 'score', 'time'):
      assert x[k] == y[k] and type(x[k]) is type(y[k])
    assert x['R'].shape == (3, 3) and x['t'].shape == (3, 1) and x['R'].dtype == np.float64
    assert np.array_equal(x['R'], y['R']) and np.array_equal(x['t'], y['t'])
  # forms only Python reads (underscores, leading '+') fall back and still load
  open(path, 'w').write('scene_id,im_id,obj_id,score,R,t,time\r\n 1, 2 ,3,+0.5,1 0 0 0 1 0 0 0 1e0,1 2 3,-1\r\n'
                        '4,5,6,1_0,1 0 0 0 1 0 0 0 1, 1 2 3 ,inf')
  r = inout.load_bop_results(path)
  assert [e['score'] for e in r] == [0.5, 10.0] and r[1]['time'] == float('inf')
  for bad in ('1,2,3,0.5,1 0 0 0 1 0 0 0 1,1 2 3\n', '1,2,3,0.5,1 0 0 0 1 0 0 0 1,1 2 3,0.1\n\n',
              '1,2,3,0.5,1 0 0 0 1 0 0 0,1 2 3,0.1', '1,2,x,0.5,1 0 0 0 1 0 0 0 1,1 2 3,0.1'):
    open(path, 'w').write(bad)
    with pytest.raises(ValueError):
      inout.load_bop_results(path)
  open(path, 'w').write('scene_id,im_id,obj_id,score,R,t,time')
  assert inout.load_bop_results(path) == []


def test_flatten_cache_hands_out_fresh_records(dataset):
  """flatten_scene_cached: same arrays and skeletons as flatten_scene, but records that a second
  pass can fill without touching what the first pass returned."""
  from bop_toolkit_b200 import dataset_params, eval_calc_errors as ece, inout
  p = _params(dataset, 'mssd', -1, '/nonexistent')
  dp = dataset_params.get_split_params(p['datasets_path'], 'lm', 'test')
  tpath = os.path.join(dp['base_path'], p['targets_filename'])
  rpath = os.path.join(p['results_path'], p['result_filenames'][0])
  torg, eorg = {}, {}
  for t in inout.load_json(tpath):
    torg.setdefault(t['scene_id'], {}).setdefault(t['im_id'], {})[t['obj_id']] = t
  for e in inout.load_bop_results(rpath):
    eorg.setdefault(e['scene_id'], {}).setdefault(e['im_id'], {}).setdefault(e['obj_id'], []).append(e)
  scene_id = sorted(torg)[0]
  gpath = dp['scene_gt_tpath'].format(scene_id=scene_id)
  gt = inout.load_scene_gt(gpath)
  ref = ece.flatten_scene(p, scene_id, torg[scene_id], eorg.get(scene_id, {}), gt)
  a = ece.flatten_scene_cached(p, scene_id, torg[scene_id], eorg.get(scene_id, {}), gt, (rpath, tpath, gpath))
  ra = ece.fill_records(a, np.full((len(a.rec), 1), 1.0))
  b = ece.flatten_scene_cached(p, scene_id, torg[scene_id], eorg.get(scene_id, {}), gt, (rpath, tpath, gpath))
  assert b.records == ref.records and b.records is not a.records          # unfilled skeletons again
  rb = ece.fill_records(b, np.full((len(b.rec), 1), 2.0))
  assert all(v == [1.0] for r in ra for v in r['errors'].values())
  assert all(v == [2.0] for r in rb for v in r['errors'].values())
  for k in ('rec', 'gt_id', 'obj_id', 'im_pos', 'pose_est', 'pose_gt'):
    assert np.array_equal(getattr(b, k), getattr(ref, k))
  assert b.im_ids == ref.im_ids and b.n_ests == ref.n_ests

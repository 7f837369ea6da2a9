"""Host-side pieces of the cfg5 strong-scaling leg (tools/cfg5_leg.py) and of the GT grouping,
on CPU: the table every rank builds is deterministic, the record hash does not depend on the
order the ranks delivered the records in, the in-memory score context gives the recalls the
per-setting scoring gives, and group_pairs_by_gt groups exactly the pairs that share a GT pose."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'tools'))


def _fake_records(c, tab, seed=0):
  n = len(tab['p_im'])
  inst = tab['p_inst']
  sc = np.stack(c.scene_of(tab['p_im']), axis=1)
  ids = np.stack([sc[:, 0], sc[:, 1], tab['inst_obj'][inst] + 1, tab['p_est_id'], tab['inst_gt'][inst]],
                 axis=1).astype(np.int32)
  rng = np.random.default_rng(seed)
  errs = np.concatenate([rng.random((n, 10)), rng.random((n, 1)) * 60, rng.random((n, 1)) * 40], axis=1)
  key_all = (tab['p_im'].astype(np.int64) << 8) | tab['p_est_id'].astype(np.int64)
  srt = np.argsort(key_all)

  def score_of(idr):
    im = (idr[:, 0].astype(np.int64) - 48) + idr[:, 1].astype(np.int64) * c.N_SCENES
    return tab['p_score'][srt[np.searchsorted(key_all[srt], (im << 8) | idr[:, 3].astype(np.int64))]]
  tab['score_of'] = score_of
  return ids, errs


def test_table_is_deterministic_and_consistent():
  import cfg5_leg as c
  a, b = c.build_table(300), c.build_table(300)
  for k in a:
    assert np.array_equal(a[k], b[k]), k
  assert a['n_inst'].min() >= 3 and a['n_inst'].max() <= 7
  assert len(a['p_im']) == c.ESTS_PER_TARGET * len(a['inst_im'])
  # distinct objects inside an image; est ids count up inside an image; pairs sorted by instance
  first = np.concatenate([[0], np.cumsum(a['n_inst'])])
  for im in range(0, 300, 37):
    objs = a['inst_obj'][first[im]:first[im + 1]]
    assert len(set(objs.tolist())) == len(objs)
    ids = a['p_est_id'][a['p_im'] == im]
    assert np.array_equal(ids, np.arange(len(ids)))
  assert (np.diff(a['p_inst']) >= 0).all()


def test_record_hash_is_order_independent_and_recalls_match_per_setting_scoring():
  import cfg5_leg as c
  from bop_toolkit_b200 import eval_calc_scores as ecs
  tab = c.build_table(120)
  models = [dict(obj_id=k + 1, diameter=100.0 + 5 * k) for k in range(c.N_OBJ)]
  ids, errs = _fake_records(c, tab)
  perm = np.random.default_rng(1).permutation(len(ids))
  assert c.records_sha256(ids, errs) == c.records_sha256(ids[perm], errs[perm])
  errs2 = errs.copy(); errs2[5, 3] = np.nextafter(errs2[5, 3], 1.0)
  assert c.records_sha256(ids, errs) != c.records_sha256(ids, errs2)
  ar, flat = c.recalls(tab, models, ids[perm], errs[perm])
  assert len(flat) == 120 and all(0.0 <= r <= 1.0 for r in flat)
  assert abs(ar['vsd'] - float(np.mean(flat[:100]))) < 1e-15
  # spot-check one setting against the plain greedy matching on the same records (mssd, threshold 0.2)
  from bop_toolkit_b200 import pose_matching
  T = len(c.TAUS)
  order = np.lexsort((ids[:, 4], ids[:, 3], ids[:, 1], ids[:, 0]))
  sid, serr = ids[order], errs[order]
  score = tab['score_of'](sid)
  tp = n_gt = 0
  first = np.concatenate([[0], np.cumsum(tab['n_inst'])])
  for im in range(120):
    s, im_id = c.scene_of(im)
    rows = np.flatnonzero((sid[:, 0] == s) & (sid[:, 1] == im_id))
    objs = tab['inst_obj'][first[im]:first[im + 1]] + 1
    n_gt += len(objs)
    for o in objs:
      r = rows[sid[rows, 2] == o]
      es = [{'est_id': int(sid[k, 3]), 'score': float(score[k]),
             'errors': {int(sid[k, 4]): [float(serr[k, T]) / models[o - 1]['diameter']]}} for k in r]
      tp += len(pose_matching.match_poses(es, [0.2], 0, None))
  assert abs(tp / n_gt - flat[100 + 3]) < 1e-12


def test_group_pairs_by_gt():
  from bop_toolkit_b200 import batch
  rng = np.random.default_rng(0)
  gt = rng.normal(size=(7, 12))
  which = rng.integers(0, 7, size=200)
  obj = (which % 3).astype(np.int32)
  im = (which // 2).astype(np.int32)
  order, off = batch.group_pairs_by_gt(obj, gt[which], im)
  assert sorted(order.tolist()) == list(range(200)) and off[0] == 0 and off[-1] == 200
  assert (np.diff(off) >= 1).all() and (np.diff(off) <= batch.MAX_GROUP).all()
  w = which[order]
  for g in range(len(off) - 1):
    assert len(set(w[off[g]:off[g + 1]].tolist())) == 1          # one GT pose per group
  # a GT pose is split only when it has more than MAX_GROUP pairs
  n_groups_of = np.bincount([w[off[g]] for g in range(len(off) - 1)], minlength=7)
  need = -(-np.bincount(which, minlength=7) // batch.MAX_GROUP)
  assert np.array_equal(n_groups_of, need)
  # same pose but another image or diameter: different groups
  o2, off2 = batch.group_pairs_by_gt(np.zeros(4, np.int32), np.tile(gt[:1], (4, 1)), np.array([0, 0, 1, 1]))
  assert len(off2) - 1 == 2
  o3, off3 = batch.group_pairs_by_gt(np.zeros(4, np.int32), np.tile(gt[:1], (4, 1)), np.zeros(4, np.int32),
                                     diameter=np.array([1.0, 2.0, 1.0, 2.0]))
  assert len(off3) - 1 == 2

"""CPU, world_size 2 over gloo: the N>1 host logic (shard assignment + record gather)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bop_toolkit_b200 import sharding


def _free_port():
  s = socket.socket()
  s.bind(('127.0.0.1', 0))
  p = s.getsockname()[1]
  s.close()
  return p


def _worker(rank, ws, port, out_dir):
  os.environ['MASTER_ADDR'] = '127.0.0.1'
  os.environ['MASTER_PORT'] = str(port)
  dist.init_process_group('gloo', rank=rank, world_size=ws)
  rng = np.random.default_rng(0)
  n_im = 37
  pairs_per_im = rng.integers(1, 6, size=n_im)
  owner = sharding.assign_images(pairs_per_im, ws)
  ids, errs = [], []
  for im in range(n_im):
    if owner[im] != rank:
      continue
    for k in range(pairs_per_im[im]):
      ids.append([1, im, 3, k, 0])
      errs.append([im + 0.01 * k, float('inf') if k == 2 else 0.5])
  ids = torch.tensor(ids, dtype=torch.int32).reshape(-1, 5)
  errs = torch.tensor(errs, dtype=torch.float64).reshape(-1, 2)
  ids_all, errs_all = sharding.gather_records(ids, errs)
  cnt = sharding.allreduce_counts(torch.tensor([ids.shape[0]], dtype=torch.int64))
  # counts known up front (deterministic assignment): same result without the count exchange
  known = [int(pairs_per_im[owner == r].sum()) for r in range(ws)]
  ids_k, errs_k = sharding.gather_records(ids, errs, counts=known)
  assert torch.equal(ids_k, ids_all) and torch.equal(errs_k, errs_all)
  # the driver's shard / gather helpers (eval_calc_errors.py) on the same data
  from bop_toolkit_b200 import eval_calc_errors as ece
  sp = ece.ScenePairs()
  sp.im_ids = list(range(n_im))
  sp.im_pos = np.repeat(np.arange(n_im), pairs_per_im).astype(np.int32)
  sp.rec = np.arange(len(sp.im_pos))
  sel = ece._shard_pairs(sp, rank, ws)
  mine = np.stack([sel.astype(np.float64), 2.0 * sel], axis=1)
  full = ece._gather_errors(mine, sel, len(sp.rec), torch.device('cpu'), sp=sp, world=ws)
  assert np.array_equal(full[:, 0], np.arange(len(sp.rec))) and np.array_equal(full[:, 1], 2.0 * full[:, 0])
  np.savez(os.path.join(out_dir, 'r%d.npz' % rank), ids=ids_all.numpy(), errs=errs_all.numpy(),
           cnt=cnt.numpy(), own=owner)
  dist.destroy_process_group()


def test_assign_images_balanced_and_deterministic():
  costs = np.array([5, 1, 1, 1, 4, 4, 2, 2], dtype=float)
  a = sharding.assign_images(costs, 2)
  b = sharding.assign_images(costs, 2)
  assert np.array_equal(a, b)
  loads = [costs[a == r].sum() for r in range(2)]
  assert abs(loads[0] - loads[1]) <= 1
  assert np.array_equal(sharding.round_robin(5, 2), [0, 1, 0, 1, 0])
  assert sharding.assign_images(costs, 1).max() == 0


def test_gather_records_world2(tmp_path):
  ws = 2
  port = _free_port()
  mp.spawn(_worker, args=(ws, port, str(tmp_path)), nprocs=ws, join=True)
  r0 = np.load(str(tmp_path / 'r0.npz'))
  r1 = np.load(str(tmp_path / 'r1.npz'))
  assert np.array_equal(r0['ids'], r1['ids']) and np.array_equal(r0['errs'], r1['errs'])
  # the gathered set equals the world_size=1 set, order-normalised
  rng = np.random.default_rng(0)
  pairs_per_im = rng.integers(1, 6, size=37)
  total = int(pairs_per_im.sum())
  assert r0['ids'].shape == (total, 5) and int(r0['cnt'][0]) == total
  key = r0['ids'][:, 1] * 100 + r0['ids'][:, 3]
  assert len(np.unique(key)) == total
  order = np.argsort(key)
  e = r0['errs'][order]
  i = r0['ids'][order]
  assert np.allclose(e[:, 0], i[:, 1] + 0.01 * i[:, 3])
  assert np.all(np.isinf(e[i[:, 3] == 2, 1]))


def test_gpu_depth_source_grouping_on_cpu():
  """Host logic of eval_calc_errors.GpuDepthSource with a stand-in decoder (no GPU): images are
  decoded in groups that cut across scenes, requests are served as views or concatenations in
  request order, groups behind the reader are dropped, late registrations re-stage a group."""
  import torch
  from bop_toolkit_b200 import eval_calc_errors as ece

  class FakeBatch(object):
    def __init__(self):
      self.staged, self.decoded = [], []

    def png_stage(self, paths, threads, slot):
      self.staged.append((list(paths), slot))
      return (list(paths),)

    def decode_depth_staged(self, staged, h, w, device):
      paths = staged[0]
      self.decoded.append(list(paths))
      ids = torch.tensor([int(p.split('_')[-1]) for p in paths], dtype=torch.int16)
      return ids.view(-1, 1, 1).expand(len(paths), h, w).contiguous()

  fb = FakeBatch()
  src = ece.GpuDepthSource(fb, 'cpu', 2, 3, group=5)
  uid = {}
  for scene, n in ((1, 3), (2, 4), (3, 6)):          # 13 images -> groups of 5, 5, 3
    ims = list(range(10, 10 + n))
    for i in ims:
      uid[(scene, i)] = len(uid)
    src.register(scene, ims, ['s%d_%d' % (scene, uid[(scene, i)]) for i in ims])
  src.register(1, [10], ['ignored_99'])               # already known: not added twice
  assert len(src.paths) == 13

  def ids(t):
    return t[:, 0, 0].tolist()
  a = src.get(1, [10, 11, 12])
  assert ids(a) == [0, 1, 2] and a.shape == (3, 2, 3)
  assert a.data_ptr() == src.live[0].data_ptr()       # consecutive inside one group: a view
  assert [len(p) for p, _ in fb.staged] == [5, 5]     # group 1 is staged while group 0 is in use
  b = src.get(2, [10, 11, 12, 13])                    # spans groups 0 and 1
  assert ids(b) == [3, 4, 5, 6] and sorted(src.live) == [0, 1]
  c = src.get(3, [12, 10])                            # not consecutive: gathered in request order
  assert ids(c) == [9, 7] and sorted(src.live) == [1]        # group 0 dropped
  d = src.get(3, [11, 13, 14, 15])                    # groups 1 and 2
  assert ids(d) == [8, 10, 11, 12] and sorted(src.live) == [1, 2]
  assert [len(p) for p in fb.decoded] == [5, 5, 3]
  assert [slot for _, slot in fb.staged] == [0, 1, 0]
  import pytest
  with pytest.raises(ValueError):
    src.get(3, [15, 11])                              # across groups, backwards
  # a scene announced late lands in the last, partly filled group: that group is decoded again
  src.register(4, [1, 2], ['s4_13', 's4_14'])
  e = src.get(4, [1, 2])
  assert ids(e) == [13, 14] and [len(p) for p in fb.decoded][-1] == 5
  assert src.get(4, []).shape == (0, 2, 3)
  src.close()
  assert src.live == {} and src.pool is None

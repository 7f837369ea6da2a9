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

"""Multi-GPU sharding of the error calculation: one process per GPU, the unit of
work is an image (keeps `depth_test` local), models are replicated per GPU, and the only
exchange is one all-gather of fixed-width error records at the end (NCCL over NVLink on the
GPU box, gloo in the CPU tests).  There is no data-path collective: pairs are independent
(scripts/eval_calc_errors.py:281-389 has no cross-pair state)."""
import numpy as np
import torch
import torch.distributed as dist


def assign_images(costs, world_size):
  """Greedy longest-processing-time assignment of images to ranks.
  costs: [n_images] estimated cost (e.g. pairs x ROI px + N x S).  Returns int32 [n_images]
  rank of every image; deterministic (ties broken by image index)."""
  costs = np.asarray(costs, dtype=np.float64)
  order = np.lexsort((np.arange(len(costs)), -costs))
  load = np.zeros(world_size, dtype=np.float64)
  owner = np.empty(len(costs), dtype=np.int32)
  for i in order:
    r = int(np.argmin(load))
    owner[i] = r
    load[r] += costs[i]
  return owner


def round_robin(n_images, world_size):
  return (np.arange(n_images) % world_size).astype(np.int32)


def gather_records(ids, errs, group=None):
  """All-gathers per-rank error records.
  ids: int32 [n, 5] = scene_id, im_id, obj_id, est_id, gt_id; errs: float64 [n, T] (score may be
  carried as an extra column).  Shards differ in size: counts are exchanged first, payloads are
  padded to the maximum.  Returns (ids_all [N,5], errs_all [N,T]) ordered by rank."""
  if not dist.is_initialized() or dist.get_world_size(group) == 1:
    return ids, errs
  ws = dist.get_world_size(group)
  dev = ids.device
  n = torch.tensor([ids.shape[0]], dtype=torch.int64, device=dev)
  counts = [torch.zeros_like(n) for _ in range(ws)]
  dist.all_gather(counts, n, group=group)
  counts = [int(c.item()) for c in counts]
  m = max(counts)
  T = errs.shape[1]
  # one payload: ids bit-cast into float64 lanes would lose nothing, but two gathers keep dtypes
  ids_p = torch.zeros((m, ids.shape[1]), dtype=ids.dtype, device=dev)
  err_p = torch.zeros((m, T), dtype=errs.dtype, device=dev)
  ids_p[:ids.shape[0]] = ids
  err_p[:errs.shape[0]] = errs
  ids_all = torch.empty((ws * m, ids.shape[1]), dtype=ids.dtype, device=dev)
  err_all = torch.empty((ws * m, T), dtype=errs.dtype, device=dev)
  if dev.type == 'cuda':
    dist.all_gather_into_tensor(ids_all, ids_p, group=group)
    dist.all_gather_into_tensor(err_all, err_p, group=group)
  else:
    l1 = [torch.empty_like(ids_p) for _ in range(ws)]
    l2 = [torch.empty_like(err_p) for _ in range(ws)]
    dist.all_gather(l1, ids_p, group=group)
    dist.all_gather(l2, err_p, group=group)
    ids_all, err_all = torch.cat(l1), torch.cat(l2)
  keep = torch.cat([torch.arange(c, device=dev) + r * m for r, c in enumerate(counts)])
  return ids_all[keep], err_all[keep]


def allreduce_counts(counts, group=None):
  """Sum of integer counters (tp, targets, ...) over ranks, consumed by score.py-style recall."""
  if dist.is_initialized() and dist.get_world_size(group) > 1:
    dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
  return counts

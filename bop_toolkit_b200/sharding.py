"""Multi-GPU sharding of the error calculation: one process per GPU, the unit of
work is an image (keeps `depth_test` local), models are replicated per GPU, and the only
exchange is one all-gather of fixed-width error records at the end (NCCL over NVLink on the
GPU box, gloo in the CPU tests).  There is no data-path collective: pairs are independent
(scripts/eval_calc_errors.py:281-389 has no cross-pair state)."""
import numpy as np
import torch
import torch.distributed as dist


def init_from_env():
  """One process per GPU under torchrun: reads RANK / LOCAL_RANK / WORLD_SIZE, binds the process
  to its GPU and initialises the process group (NCCL).  -> (rank, world, torch.device).
  BOP_B200_DIST_BACKEND=gloo is for boxes with fewer GPUs than ranks (tests): ranks then share
  GPUs (local_rank % device_count) and the record gather goes through gloo."""
  import os
  world = int(os.environ.get('WORLD_SIZE', '1'))
  rank = int(os.environ.get('RANK', '0'))
  local = int(os.environ.get('LOCAL_RANK', '0'))
  backend = os.environ.get('BOP_B200_DIST_BACKEND', 'nccl')
  n_dev = torch.cuda.device_count()
  dev_index = local % max(1, n_dev) if backend != 'nccl' else local
  torch.cuda.set_device(dev_index)
  device = torch.device('cuda', dev_index)
  if world > 1 and not dist.is_initialized():
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    if backend == 'nccl':
      dist.init_process_group('nccl', device_id=device)
    else:
      dist.init_process_group(backend)
  return rank, world, device


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


def gather_records(ids, errs, group=None, counts=None):
  """All-gathers per-rank error records in ONE collective.
  ids: int32 [n, 5] = scene_id, im_id, obj_id, est_id, gt_id; errs: float64 [n, T] (score may be
  carried as an extra column).  The ids are bit-cast into three float64 lanes in front of the
  errors, so a record is one fixed-width row of (3 + T) float64 and a single
  all_gather_into_tensor moves everything (NCCL copies bytes; no arithmetic touches the lanes).
  Shards differ in size: `counts` (rows per rank, known to every rank when the image assignment
  is deterministic) skips the count exchange; otherwise counts are all-gathered first.
  Payloads are padded to the maximum.  Returns (ids_all [N,5], errs_all [N,T]) ordered by rank."""
  if not dist.is_initialized() or dist.get_world_size(group) == 1:
    return ids, errs
  ws = dist.get_world_size(group)
  dev = ids.device
  n, T = ids.shape[0], errs.shape[1]
  if counts is None:
    c = torch.tensor([n], dtype=torch.int64, device=dev)
    cl = [torch.zeros_like(c) for _ in range(ws)]
    dist.all_gather(cl, c, group=group)
    counts = [int(x.item()) for x in cl]
  else:
    counts = [int(x) for x in counts]
    assert counts[dist.get_rank(group)] == n
  m = max(counts)
  row = torch.zeros((m, 3 + T), dtype=torch.float64, device=dev)
  if n:
    row[:n, :3].view(torch.int32)[:, :5] = ids
    row[:n, 3:] = errs
  out = torch.empty((ws * m, 3 + T), dtype=torch.float64, device=dev)
  if dev.type == 'cuda' and dist.get_backend(group) == 'nccl':
    dist.all_gather_into_tensor(out, row, group=group)
  else:
    parts = [torch.empty_like(row) for _ in range(ws)]
    dist.all_gather(parts, row, group=group)
    out = torch.cat(parts)
  if any(c != m for c in counts):
    keep = torch.cat([torch.arange(c, device=dev) + r * m for r, c in enumerate(counts)])
    out = out[keep]
  ids_all = out[:, :3].contiguous().view(torch.int32)[:, :5]
  return ids_all, out[:, 3:]


def allreduce_counts(counts, group=None):
  """Sum of integer counters (tp, targets, ...) over ranks, consumed by score.py-style recall."""
  if dist.is_initialized() and dist.get_world_size(group) > 1:
    dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
  return counts

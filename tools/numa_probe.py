#!/usr/bin/env python
"""Where do pinned host pages live relative to each GPU?  (diagnostic for the end-to-end rate at N>1)

  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/numa_probe.py

Per rank: the GPU's NUMA node (sysfs), NVML's ideal CPU set, the CPUs / memory nodes this process
may use, then the host->device rate of a 1 GiB pinned buffer (a) one rank at a time and (b) all
ranks at once, first with the default memory policy and then -- when the kernel lets us -- with
the pages bound to the GPU's own node (set_mempolicy(MPOL_PREFERRED)).  One JSON line per rank.
"""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def read(path):
  try:
    with open(path) as f:
      return f.read().strip()
  except OSError as e:
    return 'ERR %s' % e.__class__.__name__


def status_field(name):
  for line in read('/proc/self/status').splitlines():
    if line.startswith(name):
      return line.split(':', 1)[1].strip()
  return None


def set_mempolicy(mode, node):
  libc = ctypes.CDLL(None, use_errno=True)
  if node is None:
    rc = libc.syscall(238, 0, None, 0)
  else:
    mask = ctypes.c_ulong(1 << node)
    rc = libc.syscall(238, mode, ctypes.byref(mask), 64)
  return rc if rc == 0 else -ctypes.get_errno()


def main():
  import torch
  import torch.distributed as dist
  world = int(os.environ.get('WORLD_SIZE', '1'))
  rank = int(os.environ.get('RANK', '0'))
  local = int(os.environ.get('LOCAL_RANK', '0'))
  torch.cuda.set_device(local)
  if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
  barrier = (lambda: dist.barrier()) if world > 1 else (lambda: None)
  props = torch.cuda.get_device_properties(local)
  pci = '%04x:%02x:%02x.0' % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
  info = {'rank': rank, 'pci': pci, 'gpu_numa_node': read('/sys/bus/pci/devices/%s/numa_node' % pci),
          'gpu_local_cpulist': read('/sys/bus/pci/devices/%s/local_cpulist' % pci),
          'nodes_online': read('/sys/devices/system/node/online'),
          'cpus_allowed': status_field('Cpus_allowed_list'), 'mems_allowed': status_field('Mems_allowed_list')}
  n_bytes = 1 << 30
  dst = torch.empty(n_bytes, dtype=torch.uint8, device='cuda')

  def rate(buf):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
      dst.copy_(buf, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    return round(3 * n_bytes / (e0.elapsed_time(e1) * 1e-3) / 1e9, 2)

  def measure(tag, register=False):
    if register:            # pages first-touched by this process under its policy, then registered
      import numpy as np
      arr = np.empty(n_bytes, np.uint8)
      arr.fill(1)
      rc = torch.cuda.cudart().cudaHostRegister(arr.ctypes.data, n_bytes, 0)
      info[tag + '_register_rc'] = int(rc)
      buf = torch.from_numpy(arr)
      info[tag + '_is_pinned'] = bool(buf.is_pinned())
    else:
      buf = torch.empty(n_bytes, dtype=torch.uint8).pin_memory()
      buf.fill_(1)
    rate(buf)
    for r in range(world):
      barrier()
      if r == rank:
        info[tag + '_alone_GBps'] = rate(buf)
    barrier()
    info[tag + '_all_ranks_GBps'] = rate(buf)
    barrier()
    if register:
      torch.cuda.cudart().cudaHostUnregister(arr.ctypes.data)
    del buf

  measure('default')
  node = info['gpu_numa_node']
  if node.lstrip('-').isdigit() and int(node) >= 0:
    rc = set_mempolicy(1, int(node))          # MPOL_PREFERRED
    info['set_mempolicy_rc'] = rc
    if rc == 0:
      measure('gpu_node')
      measure('gpu_node_reg', register=True)
      set_mempolicy(0, None)
  for r in range(world):
    barrier()
    if r == rank:
      print(json.dumps(info), flush=True)
  if world > 1:
    dist.destroy_process_group()


if __name__ == '__main__':
  main()

"""CPU / memory placement next to a GPU.

The end-to-end path hands PINNED host buffers to the kernels (depth images are read in place over
PCIe).  On a multi-socket host those pages should live on the NUMA node the GPU hangs off;
bind_to_gpu() pins the calling process to the CPUs local to the GPU (NVML's ideal affinity, sysfs
as fallback) BEFORE the buffers are allocated, so first-touch places them there.  On the r01
measurement boxes (one NUMA node, 16 CPUs) it is a no-op and is recorded as such in the bench line.
(Run-to-run swings of the zero-copy rate seen at 4000 pairs were an L2 artefact -- 110 MB of
fetched depth fits the 126 MB L2, so some runs never touched PCIe after warm-up; at the 10,000-pair
bench size, 280 MB per step, six fresh processes gave 1.029-1.032 M pairs/s.)"""
import os


def _sysfs_cpulist(index):
  import torch
  bus = torch.cuda.get_device_properties(index)
  pci = '%04x:%02x:%02x.0' % (bus.pci_domain_id, bus.pci_bus_id, bus.pci_device_id)
  with open('/sys/bus/pci/devices/%s/local_cpulist' % pci) as f:
    cpus = set()
    for part in f.read().strip().split(','):
      if '-' in part:
        a, b = part.split('-')
        cpus.update(range(int(a), int(b) + 1))
      elif part:
        cpus.add(int(part))
  return cpus


def bind_to_gpu(index=0):
  """Restricts this process to the CPUs local to GPU `index`.  Returns a short description
  (for logs) or None when nothing could be done; never raises."""
  try:
    allowed = os.sched_getaffinity(0)
  except AttributeError:
    return None
  cpus, how = None, None
  try:
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(index)
    words = (max(allowed) + 64) // 64
    mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
    cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
    how = 'nvml'
  except Exception:
    try:
      cpus, how = _sysfs_cpulist(index), 'sysfs'
    except Exception:
      return None
  cpus = set(cpus) & set(allowed)
  if not cpus:
    return None
  try:
    os.sched_setaffinity(0, cpus)
  except OSError:
    return None
  return '%s: %d CPUs local to GPU %d (of %d allowed)' % (how, len(cpus), index, len(allowed))

"""Two disjoint SM partitions of one GPU, a stream in each (CUDA green contexts, driver API through
cuda-python: host plumbing like torch's streams, no arithmetic).

Why: with HOST-resident depth images the fused VSD kernel is bound by the PCIe link, not by the
SMs, and the geometry / MSSD / MSPD kernels of a batch cannot overlap it from a second stream by
themselves -- every kernel of this library fills the register file of the SMs it runs on, so the
block scheduler serialises them.  With the SMs split, two independent streams of batches run side
by side: while one partition's pixel pass waits on the link, the other one's geometry pass
computes.  Calls on the two streams need a batch.Workspace each (batch.private_workspace).
No counterpart in the reference (its evaluation is one Python process per core).
"""
import torch


def _ck(ret):
  from cuda.bindings import driver as cu
  if ret[0] != cu.CUresult.CUDA_SUCCESS:
    raise RuntimeError('CUDA driver call failed: %s' % (ret[0],))
  return ret[1] if len(ret) == 2 else ret[1:]


class SmPartition(object):
  """streams[0] runs on `sms_a` SMs (rounded up by the driver to its partition granularity),
  streams[1] on the rest; sm_counts holds what the driver gave."""

  def __init__(self, device_index, sms_a):
    from cuda.bindings import driver as cu
    self.device = torch.device('cuda', device_index)
    with torch.cuda.device(self.device):
      torch.zeros(1, device=self.device)           # primary context up before the green contexts
      _ck(cu.cuInit(0))
      dev = _ck(cu.cuDeviceGet(device_index))
      res = _ck(cu.cuDeviceGetDevResource(dev, cu.CUdevResourceType.CU_DEV_RESOURCE_TYPE_SM))
      total = int(res.sm.smCount)
      if not 0 < sms_a < total:
        raise ValueError('sms_a must be in (0, %d)' % total)
      groups, n_groups, rest = _ck(cu.cuDevSmResourceSplitByCount(1, res, 0, int(sms_a)))
      if int(n_groups) != 1 or int(rest.sm.smCount) == 0:
        raise RuntimeError('the driver could not split %d SMs into %d + rest' % (total, sms_a))
      self.sm_counts = (int(groups[0].sm.smCount), int(rest.sm.smCount))
      self._ctx, self._raw, self.streams = [], [], []
      for r in (groups[0], rest):
        desc = _ck(cu.cuDevResourceGenerateDesc([r], 1))
        g = _ck(cu.cuGreenCtxCreate(desc, dev, cu.CUgreenCtxCreate_flags.CU_GREEN_CTX_DEFAULT_STREAM))
        st = _ck(cu.cuGreenCtxStreamCreate(g, cu.CUstream_flags.CU_STREAM_NON_BLOCKING, 0))
        self._ctx.append(g)
        self._raw.append(st)
        self.streams.append(torch.cuda.ExternalStream(int(st), device=self.device))

  def close(self):
    from cuda.bindings import driver as cu
    torch.cuda.synchronize(self.device)
    for st in self._raw:
      cu.cuStreamDestroy(st)
    for g in self._ctx:
      cu.cuGreenCtxDestroy(g)
    self._raw, self._ctx, self.streams = [], [], []

#!/bin/bash
# Resident-step timing for a list of BOP_TUNE settings (no culling, 10k pairs).  usage: gpu_tune.sh tune...
for t in "$@"; do
  echo "TUNE $t"
  BOP_TUNE=$t timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --cfg5-images 0 2>&1 | python -c "
import sys,json
for l in sys.stdin:
  if l.startswith('{'):
    d=json.loads(l); r=d['roofline']; print('  pairs/s %.0f ms %.3f fused %.3f geometry %.3f | cull_auto %.0f fused %.3f' % (d['value'], d['ms_per_step'], r['kernel_ms'], r['other_kernels_ms']['geometry'], d['cull_auto']['value'], d['cull_auto']['kernels_ms']['fused']))
  elif 'rror' in l: print(l.strip()[:300])
"
done

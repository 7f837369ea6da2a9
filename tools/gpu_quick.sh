#!/bin/bash
# Quick check on the GPU box: parity tests + resident/e2e throughput at 4000 pairs (+ optional BOP_TUNE list).
python -m pytest tests -m gpu -q -x 2>&1 | tail -2
for t in "${@:-0,0,0,0}"; do
  echo "TUNE $t"
  BOP_TUNE=$t python bench.py --pairs 4000 --steps 5 --warmup 3 --no-cpu-baseline | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('  pairs/s %.0f fused_ms %.3f other %s e2e %.0f' % (d['value'], d['roofline']['kernel_ms'], {k: round(v,3) for k,v in d['roofline']['other_kernels_ms'].items()}, d['e2e']['value']))"
done

#!/bin/bash
# Run-to-run variance of the end-to-end (pinned host, zero-copy) path: 8 fresh processes.
for i in 1 2 3 4 5 6; do
  python bench.py --pairs ${PAIRS:-10000} --steps 5 --warmup 3 --no-cpu-baseline | python -c "
import sys,json,os
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']
print('pairs/s %.0f e2e %.0f (%.2f ms) full-copy e2e %.0f  h2d %.1f MB/step' % (d['value'], e['value'], e['ms_per_step'], e['full_copy']['value'], e['h2d_bytes_per_step']/1e6))"
done

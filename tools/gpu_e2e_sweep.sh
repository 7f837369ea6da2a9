#!/bin/bash
# End-to-end leg on two SM partitions: sweep "SMS:PIECES" settings.  usage: gpu_e2e_sweep.sh TAG 88:1,1 88:1,2 ...
TAG=${1:-e2e}; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
for cfg in "$@"; do
  sms=${cfg%%:*}; pieces=${cfg#*:}
  timeout 600 python -X faulthandler bench.py --steps 10 --warmup 3 --no-cpu-baseline --cfg5-images 0 --e2e-sms $sms --e2e-pieces $pieces $E2E_EXTRA > $OUT/bench_$cfg.json 2> $OUT/bench_$cfg.err
  python - $OUT/bench_$cfg.json $cfg <<'PY'
import sys, json
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); e = d['e2e']
print(sys.argv[2], 'resident %.0f' % d['value'], 'e2e %.0f (%.2f ms)' % (e['value'], e['ms_per_step']),
      'serial %.0f (%.2f ms)' % (e['serial']['value'], e['serial']['ms_per_step']), 'partition', e['sm_partition'])
PY
  tail -2 $OUT/bench_$cfg.err
done

#!/bin/bash
# Quick check on the GPU box: parity tests + resident/e2e throughput without and with the opt-in
# back-face culling (+ optional BOP_TUNE list).   usage: gpu_quick.sh TAG [PAIRS] [tune ...]
TAG=${1:-quick}; PAIRS=${2:-10000}; shift; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu.log
for t in "${@:-0,0,0,0}"; do
 for cull in off auto; do
  echo "TUNE $t cull=$cull"
  BOP_B200_CULL=$cull BOP_TUNE=$t timeout 600 python bench.py --pairs $PAIRS --steps 10 --warmup 3 --no-cpu-baseline 2> $OUT/bench_${t}_$cull.err | tee $OUT/bench_${t}_$cull.json | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('  pairs/s %.0f ms %.3f fused_ms %.3f other %s e2e %.0f e2e_kernels %s' % (d['value'], d['ms_per_step'], r['kernel_ms'], {k: round(v,3) for k,v in r['other_kernels_ms'].items()}, d['e2e']['value'], {k: round(v,3) for k,v in d['e2e']['kernels_ms'].items()}))"
  tail -3 $OUT/bench_${t}_$cull.err
 done
done

#!/bin/bash
# Full ncu captures (one launch each) of the fused STEP kernel and the assemble kernel at PAIRS pairs.
# usage: gpu_ncu2.sh TAG [PAIRS]
TAG=${1:-ncu}; PAIRS=${2:-2000}
OUT=gpurun_out/$TAG; mkdir -p $OUT
CMD="python bench.py --pairs $PAIRS --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 $CMD > $OUT/bench_small.json 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
  -k 'regex:fused_kernel<\(int\)0' --launch-skip 3 -c 1 -f -o $OUT/fused $CMD > $OUT/ncu_fused.log 2>&1; echo "ncu fused rc=$?"
# the assemble launches of model set-up and depth synthesis come first (grid <= 800); skip to the bench steps
N_SETUP=$(python - <<'PY'
print(0)
PY
)
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
  -k 'regex:geometry_kernel' --launch-skip ${ASM_SKIP:-8} -c 1 -f -o $OUT/assemble $CMD > $OUT/ncu_asm.log 2>&1; echo "ncu assemble rc=$?"
ls -la $OUT

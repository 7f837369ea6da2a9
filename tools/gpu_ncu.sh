#!/bin/bash
# ncu evidence only: launch list + one full capture of the fused VSD kernel (STEP cost).
TAG=${1:-ncu}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python bench.py --pairs 2000 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/bench_small.json 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches.csv \
  python bench.py --pairs 2000 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
  -k 'regex:fused_kernel<\(int\)0' --launch-skip 3 -c 1 -f -o $OUT/fused \
  python bench.py --pairs 2000 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"
grep -c "PROF" $OUT/ncu_full.log; ls -la $OUT

#!/bin/bash
# ncu launch list (per-kernel gpu time) of a short bench run.  usage: gpu_launches.sh TAG [PAIRS]
TAG=${1:-launches}; PAIRS=${2:-10000}
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $OUT/launches.csv \
  python bench.py --pairs $PAIRS --steps 2 --warmup 3 --no-cpu-baseline > $OUT/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python tools/launch_summary.py $OUT/launches.csv "bench.py --pairs $PAIRS --steps 2 --warmup 3" | tee $OUT/launch_summary.txt

#!/bin/bash
# One full ncu capture of the geometry kernel at PAIRS pairs.  usage: gpu_ncu_geom.sh TAG [PAIRS] [SKIP]
TAG=${1:-ncu}; PAIRS=${2:-2000}; SKIP=${3:-31}
OUT=gpurun_out/$TAG; mkdir -p $OUT
CMD="python bench.py --pairs $PAIRS --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
  -k 'regex:geometry_kernel' --launch-skip $SKIP -c 1 -f -o $OUT/geometry $CMD > $OUT/ncu_geom.log 2>&1; echo "ncu geometry rc=$?"
ls -la $OUT

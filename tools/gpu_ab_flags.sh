#!/bin/bash
# A/B of nvcc flag variants on the GPU box: builds each, runs the tuned bench line
for f in "$@"; do
  echo "FLAGS $f"
  BOP_NVCC_FLAGS="$f" python -c "from bop_toolkit_b200 import _lib; _lib.build()" > /dev/null 2>&1
  BOP_NVCC_FLAGS="$f" bash tools/gpu_tune.sh 0,0,0,0 | tail -1
done

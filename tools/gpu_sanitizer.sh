#!/bin/bash
# compute-sanitizer on the smoke workload (fused raster + pixel pass, geometry sort, MSSD / MSPD)
# and on the GPU PNG-decode tests.  usage: gpu_sanitizer.sh TAG
TAG=${1:-sanitizer}
OUT=gpurun_out/$TAG; mkdir -p $OUT
SMOKE='import __graft_entry__ as g; g.smoke()'
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python -c "$SMOKE" > $OUT/smoke_$tool.log 2>&1
  echo "smoke $tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|smoke ok|hazard" $OUT/smoke_$tool.log | tail -5
done
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python -m pytest tests/test_png_depth.py -m gpu -x -q > $OUT/png_$tool.log 2>&1
  echo "png $tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|hazard" $OUT/png_$tool.log | tail -5
done

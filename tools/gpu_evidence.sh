#!/bin/bash
# Round evidence in one visit: bench (both arms), ncu launch list, full captures of the fused,
# geometry and MSPD kernels (no culling = the default).  usage: gpu_evidence.sh TAG
TAG=${1:-evidence}
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --pairs 2000 --steps 2 --warmup 3 --no-cpu-baseline --cfg5-images 0 --e2e-sms 0"
timeout 900 $CMD > $OUT/bench_small.json 2> /dev/null
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
# launches before the bench steps: cull probes of the second ModelSet (3 per model) + depth synthesis (5 chunks)
for k in "fused:fused_kernel<\(int\)0:3" "geometry:geometry_kernel:31" "mspd:sym_max_kernel<\(int\)1:3" "mssd:sym_max_kernel<\(int\)0:3"; do
  name=${k%%:*}; rest=${k#*:}; re=${rest%:*}; skip=${rest##*:}
  timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k "regex:$re" --launch-skip $skip -c 1 -f -o $OUT/$name $CMD > $OUT/ncu_$name.log 2>&1; echo "ncu $name rc=$?"
done
ls -la $OUT

#!/bin/bash
# One full ncu capture of the first launch of a kernel whose grid is at least MIN_GRID CTAs
# (skips the small launches of model set-up).  usage: gpu_ncu_kernel.sh TAG KERNEL_REGEX MIN_GRID [PAIRS]
TAG=${1:-ncu}; KRE=${2:-assemble_kernel}; MING=${3:-1000}; PAIRS=${4:-2000}
OUT=gpurun_out/$TAG
mkdir -p $OUT
CMD="python bench.py --pairs $PAIRS --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k "regex:$KRE" \
  -c 400 --csv --log-file $OUT/list.csv $CMD > $OUT/list.log 2>&1; echo "list rc=$?"
SKIP=$(python - "$OUT/list.csv" "$MING" <<'PY'
import csv, sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]
gi = hdr.index('Grid Size')
n = 0
for r in rows[1:]:
  g = int(r[gi].strip('()').split(',')[0])
  if g >= int(sys.argv[2]):
    break
  n += 1
print(n + 2)   # two launches further: past the first (cold) large launch
PY
)
echo "skip=$SKIP"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
  -k "regex:$KRE" --launch-skip $SKIP -c 1 -f -o $OUT/kernel $CMD > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $OUT

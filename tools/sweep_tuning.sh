for t in "20480,1024,1,160" "16384,1024,1,160" "20480,1024,1,256" "16384,1024,1,256" "12288,1024,1,256" "20480,1024,1,1000" "18432,1024,1,400"; do
  echo "TUNE $t"; BOP_TUNE=$t python bench.py --pairs 4000 --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
  if l.startswith('{'):
    d=json.loads(l); print('  pairs/s %.0f  fused_ms %.3f  frac %.3f'%(d['value'], d['roofline']['kernel_ms'], d['roofline']['frac']))
  elif 'rror' in l: print(l.strip()[:200])
"
done

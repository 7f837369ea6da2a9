for t in "24576,512,1,36" "24576,1024,1,36" "12288,512,2,36" "24576,512,1,16" "24576,512,1,64" "16384,768,1,36" "12288,1024,1,36" "8192,512,2,36"; do
  echo "TUNE $t"; BOP_TUNE=$t python bench.py --pairs 4000 --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
  if l.startswith('{'):
    d=json.loads(l); print('  pairs/s %.0f  fused_ms %.3f  frac %.3f'%(d['value'], d['roofline']['kernel_ms'], d['roofline']['frac']))
  elif 'rror' in l: print(l.strip()[:200])
"
done

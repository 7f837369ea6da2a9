#!/usr/bin/env python
"""Summarises an ncu launch list (--metrics gpu__time_duration.sum --csv --log-file X): launches,
total time and share per kernel.  Usage: python tools/launch_summary.py launches.csv ["header"]"""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors='replace')) if len(r) > 10]
hdr = rows[0]
ik, iv, iu, im = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit'), hdr.index('Metric Name')
agg = {}
for r in rows[1:]:
  if r[im] != 'gpu__time_duration.sum':
    continue
  v = float(r[iv].replace(',', ''))
  v *= {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6}.get(r[iu], 1.0)
  name = r[ik]
  a = agg.setdefault(name, [0, 0.0])
  a[0] += 1; a[1] += v
ours = {k: v for k, v in agg.items() if ('<unnamed>::' in k or 'bop_' in k) and 'at::' not in k}
tot = sum(v[1] for v in ours.values()) or 1.0
if len(sys.argv) > 2:
  print('# ' + sys.argv[2])
print('# total launches captured: %d, of which this library: %d; library time %.1f us, other (torch) %.1f us'
      % (sum(v[0] for v in agg.values()), sum(v[0] for v in ours.values()), tot,
         sum(v[1] for k, v in agg.items() if k not in ours)))
for k, v in sorted(ours.items(), key=lambda kv: -kv[1][1]):
  print('%-72s launches %4d  total %10.1f us  share %5.1f%%' % (k[:72], v[0], v[1], 100 * v[1] / tot))

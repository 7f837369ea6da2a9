#!/usr/bin/env python
"""Instruction / sample / lane shares of the fused kernel by phase (source-line ranges of vsd.cu and
raster.cuh given on the command line as name:file:lo-hi,...).  usage: ncu_phases.py REP spec..."""
import csv, io, subprocess, sys
rep = sys.argv[1]
specs = []
for a in sys.argv[2:]:
  name, f, r = a.split(':'); lo, hi = r.split('-'); specs.append((name, f, int(lo), int(hi)))
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
cur, h2, agg = None, None, {}
for r in csv.reader(io.StringIO(src)):
  if len(r) >= 2 and r[0] == 'File Path':
    cur = r[1].split('/')[-1]; continue
  if r and r[0] == 'Line No':
    h2 = r; continue
  if h2 is None or len(r) < 10 or r[2] != '-':
    continue
  try:
    line, inst, samp, tinst = int(r[0]), int(r[7]), int(r[6]), int(r[8])
  except ValueError:
    continue
  name = 'other'
  for n, f, lo, hi in specs:
    if cur.startswith(f) and lo <= line <= hi:
      name = n; break
  a = agg.setdefault(name, [0, 0, 0]); a[0] += inst; a[1] += samp; a[2] += tinst
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
  print('%-16s warp-inst %5.1f%%  samples %5.1f%%  lanes %4.1f  (lane-slots lost: %4.1f%% of all)' % (
    k, 100.0 * a[0] / ti, 100.0 * a[1] / ts, a[2] / max(1, a[0]), 100.0 * (32 * a[0] - a[2]) / (32 * ti)))

#!/usr/bin/env python
"""Summarises an .ncu-rep: headline metrics (raw page) + instruction share per source line.
Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [top_n]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__block_size', 'launch__grid_size', 'launch__shared_mem_per_block_dynamic',
        'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
for w in want:
  for i, h in enumerate(hdr):
    if h == w:
      print('%-70s %-12s %s' % (w, units[i], [r[i][:60] for r in data]))
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
cur, h2, agg, stalls = None, None, {}, {}
for r in csv.reader(io.StringIO(src)):
  if len(r) >= 2 and r[0] == 'File Path':
    cur = r[1]; continue
  if r and r[0] == 'Line No':
    h2 = r; continue
  if h2 is None or len(r) < 10 or r[2] != '-':
    continue
  try:
    line, inst, samp, tinst = int(r[0]), int(r[7]), int(r[6]), int(r[8])
  except ValueError:
    continue
  a = agg.setdefault((cur.split('/')[-1], line), [0, 0, 0, r[1]])
  a[0] += inst; a[1] += samp; a[2] += tinst
  for i, name in enumerate(h2):
    if name.startswith('stall_') and 'Not Issued' not in name:
      try:
        stalls[name] = stalls.get(name, 0) + int(r[i])
      except (ValueError, IndexError):
        pass
tot = sum(a[0] for a in agg.values()) or 1
ts = sum(a[1] for a in agg.values()) or 1
print('\ntotal warp-inst %d, thread-inst %d, samples %d' % (tot, sum(a[2] for a in agg.values()), ts))
print('stall samples:', ', '.join('%s %.1f%%' % (k[6:], 100.0 * v / max(1, sum(stalls.values())))
                                  for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:9]))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
  print('%-14s %4d inst %5.1f%% samp %5.1f%% lanes %4.1f | %s' % (k[0][:14], k[1], 100.0 * a[0] / tot,
                                                                100.0 * a[1] / ts, a[2] / max(a[0], 1), a[3][:88]))

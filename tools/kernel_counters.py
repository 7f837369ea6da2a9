#!/usr/bin/env python
"""Per-pair instruction / DRAM counters of the dominant kernel from ONE ncu --set full capture
-> profiles/kernel_counters.json (read by bench.py's `roofline`).
usage: python tools/kernel_counters.py REP PAIRS_PAST_GATE "source note" [out.json]"""
import csv
import io
import json
import subprocess
import sys

rep, pairs, note = sys.argv[1], int(sys.argv[2]), sys.argv[3]
out = sys.argv[4] if len(sys.argv) > 4 else 'profiles/kernel_counters.json'
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2]


def val(name):
  i = hdr.index(name)
  v = float(data[i].replace(',', ''))
  u = units[i]
  return v * {'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'byte': 1.0, 'us': 1e-3, 'ms': 1.0, 'ns': 1e-6, 's': 1e3}.get(u, 1.0)


winst = val('smsp__inst_executed.sum')
lanes = val('smsp__thread_inst_executed_per_inst_executed.ratio')
dram = val('dram__bytes_read.sum') + val('dram__bytes_write.sum')
d = {
  'fused_warp_inst_per_pair': winst / pairs,
  'fused_thread_inst_per_pair': winst * lanes / pairs,
  'fused_dram_bytes_per_pair': dram / pairs,
  'pairs_in_launch': pairs,
  'capture': {'kernel': data[hdr.index('Kernel Name')], 'gpu_time_ms': val('gpu__time_duration.sum'),
              'issue_active_pct': val('smsp__issue_active.avg.pct_of_peak_sustained_active'),
              'lanes_per_inst': lanes, 'registers': val('launch__registers_per_thread'),
              'dram_pct_of_peak': val('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')},
  'source': note,
}
json.dump(d, open(out, 'w'), indent=1)
print(json.dumps(d, indent=1))

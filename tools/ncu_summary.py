#!/usr/bin/env python
"""Summarise an .ncu-rep: headline metrics per captured launch + top source lines per kernel.
usage: ncu_summary.py report.ncu-rep [kernel-regex ...]"""
import csv
import io
import os
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'smsp__inst_executed.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active']
rep = sys.argv[1]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print('## ' + r[hdr.index('Kernel Name')][:90])
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            print(w, units[i], r[i])
    # warp-state breakdown: stall cycles per issued instruction, largest first
    st = []
    for i, h in enumerate(hdr):
        if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio'):
            try:
                st.append((float(r[i].replace(',', '')), h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]))
            except ValueError:
                pass
    for v, h in sorted(st, reverse=True)[:8]:
        print('stall_per_issue.%s %.3f' % (h, v))
here = os.path.dirname(os.path.abspath(__file__))
for pat in sys.argv[2:]:
    src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass,cuda', '--kernel-name',
                          'regex:' + pat], capture_output=True, text=True).stdout
    open('/tmp/_ncu_src.csv', 'w').write(src)
    print('\n## top source lines by executed warp instructions: ' + pat)
    print(subprocess.run([sys.executable, os.path.join(here, 'ncu_lines.py'), '/tmp/_ncu_src.csv', '30'],
                         capture_output=True, text=True).stdout)

#!/usr/bin/env python3
"""ncu_metrics.py REPORT.ncu-rep -- the raw-page metrics the profiles/*_full_metrics.txt summaries quote (time, DRAM bytes,
shared-memory wavefronts and bank conflicts, instruction count, issue / pipe utilisation, stall reasons per issue)."""
import csv
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum', 'smsp__inst_executed.sum', 'smsp__inst_executed_op_shared_ld.sum',
        'sm__cycles_elapsed.max', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'smsp__warps_active.avg.per_cycle_active'] + [
        f'smsp__average_warps_issue_stalled_{k}_per_issue_active.ratio' for k in
        ('barrier', 'short_scoreboard', 'long_scoreboard', 'mio_throttle', 'wait', 'math_pipe_throttle', 'lg_throttle', 'not_selected')]

out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
for r in rows[2:]:
    print(r[hdr.index('Kernel Name')] if 'Kernel Name' in hdr else '')
    for w in WANT:
        if w in hdr:
            print(' ', w, rows[1][hdr.index(w)], r[hdr.index(w)])

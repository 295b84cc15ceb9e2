#!/usr/bin/env python
"""Key counters of the first kernel in an ncu report as CSV (metric,unit,value).
  python tools/ncu_summary.py <report.ncu-rep> [> profiles/x.csv]"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines()))
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__warps_active.avg.pct', 'launch__',
        'sm__inst_executed_pipe_fp64', 'sm__pipe_fp64_cycles_active.avg', 'smsp__issue_active.avg',
        'smsp__average_warps_issue_stalled', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared', 'sm__throughput.avg',
        'smsp__inst_executed.sum', 'smsp__warps_eligible.avg', 'lts__t_sector_hit_rate', 'dram__throughput.avg.pct',
        'smsp__average_warp_latency', 'sm__inst_executed_pipe_tensor', 'sm__pipe_tensor', 'sm__inst_executed_pipe_fmaheavy']
print('metric,unit,value')
for h, u, v in zip(r[0], r[1], r[2]):
    if any(k in h for k in keys):
        print(f'{h},{u},{v}')

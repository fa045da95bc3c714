#!/usr/bin/env python
"""Text summary of one `ncu --set full` report (the lines profiles/*.txt hold): usage
   tools/ncu_summary.py report.ncu-rep > profiles/rNN_xxx.txt"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
keep = ['Kernel Name', 'gpu__time_duration.sum', 'sm__pipe_tensor', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'launch__', 'sm__cycles_elapsed.max', 'smsp__average_warp', 'smsp__warp_issue_stalled', 'sm__inst_executed_pipe',
        'l1tex__data_pipe', 'sm__throughput', 'lts__throughput', 'smsp__inst_executed.sum', 'smsp__issue_active',
        'sm__warps_active', 'l1tex__data_bank', 'smsp__inst_executed_pipe_xu']
for h, u, v in zip(rows[0], rows[1], rows[2]):
    if any(h.startswith(k) or k in h for k in keep):
        print(f'{h} [{u}] = {v}')

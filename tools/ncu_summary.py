#!/usr/bin/env python3
"""Text summary of one `ncu --set full` report (the metrics DESIGN.md / bench.py quote).
usage: ncu_summary.py report.ncu-rep "title line" > profiles/...summary.txt"""
import csv, subprocess, sys
rows = list(csv.reader(subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
h = rows[0]
keep = ('dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__time_duration.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'launch__block_size', 'launch__grid_size', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio')
print(sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])
for i, n in enumerate(h):
    if n in keep or ('issue_stalled' in n and 'per_issue' in n):
        print('%-90s %-16s %s' % (n, rows[1][i], " ".join(r[i] for r in rows[2:])))

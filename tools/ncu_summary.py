#!/usr/bin/env python3
"""One-line-per-kernel summary of an ncu report's raw page:  ncu -i rep --page raw --csv | python tools/ncu_summary.py [out.csv]"""
import csv
import sys

WANT = ['Kernel Name', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread', 'gpu__time_duration.sum',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'launch__waves_per_multiprocessor',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__cycles_active.avg', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct']
rows = list(csv.reader(sys.stdin))
hdr, units = rows[0], rows[1]
idx = [hdr.index(x) for x in WANT if x in hdr]
out = csv.writer(open(sys.argv[1], "a") if len(sys.argv) > 1 else sys.stdout)
out.writerow([hdr[i] for i in idx]); out.writerow([units[i] for i in idx])
for r in rows[2:]:
    out.writerow([r[i] for i in idx])

#!/usr/bin/env python
"""print selected metrics of an ncu report: python profiles/ncu_extract.py report.ncu-rep [metric substrings...]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
default = ['Kernel Name', 'launch__grid_size', 'launch__registers_per_thread', 'gpu__time_duration.sum',
           'sm__ops_path_tensor_src_fp64.sum', 'sm__ops_path_tensor_src_fp64.sum.pct_of_peak_sustained_elapsed',
           'sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active',
           'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
           'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
           'sm__cycles_active.avg', 'sm__cycles_elapsed.max', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
           'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
           'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
           'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum']
want = sys.argv[2:] or default
for r in rows[2:]:
    print('----')
    for i, h in enumerate(hdr):
        if any((w == h) if not sys.argv[2:] else (w in h) for w in want):
            print("%-90s %s %s" % (h, r[i], units[i]))

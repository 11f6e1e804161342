#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key raw metrics per kernel -> CSV on stdout or a file."""
import csv, subprocess, sys, io
KEYS = ['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','lts__t_bytes.sum','smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.sum','sm__inst_executed_pipe_alu.sum',
        'sm__inst_executed_pipe_lsu.sum','sm__inst_executed_pipe_xu.sum','sm__inst_executed_pipe_fp64.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'launch__registers_per_thread','launch__grid_size','launch__block_size','sm__cycles_elapsed.max','smsp__cycles_active.avg',
        'smsp__average_warp_latency_issue_stalled_barrier.pct','smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio','smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio','smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio','smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio','smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_membar_per_issue_active.ratio','smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio']
rep = sys.argv[1]
raw = subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
out = csv.writer(sys.stdout)
out.writerow(['metric','unit']+[r[hdr.index('Kernel Name')][:40] for r in rows[2:]])
for k in KEYS:
    if k in hdr:
        i = hdr.index(k); out.writerow([k, units[i]]+[r[i] for r in rows[2:]])

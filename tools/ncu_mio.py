#!/usr/bin/env python
"""Shared-memory data-pipe summary of an .ncu-rep (read here, no GPU): wavefronts by kind, pipe utilisation, issue, stalls."""
import csv, subprocess, sys, io
rep = sys.argv[1]
steps = float(sys.argv[2]) if len(sys.argv) > 2 else None     # warp steps in the launch (frame pairs for n_fft 1024)
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
for r in rows[2:]:
    g = lambda k: float(r[hdr.index(k)].replace(',', '')) if k in hdr and r[hdr.index(k)] not in ('', 'n/a') else float('nan')
    print('kernel', r[hdr.index('Kernel Name')][:60])
    ms = g('gpu__time_duration.sum'); print(' time', ms, 'ms')
    tot = g('l1tex__data_pipe_lsu_wavefronts_mem_shared.sum'); ld = g('l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum'); st = g('l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum')
    print(' shared wavefronts total %.4g ld %.4g st %.4g other %.4g' % (tot, ld, st, tot - ld - st))
    if steps: print(' per step: total %.1f ld %.1f st %.1f other %.1f' % (tot / steps, ld / steps, st / steps, (tot - ld - st) / steps))
    print(' bank conflicts ld %.4g st %.4g' % (g('l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum'), g('l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum')))
    print(' data pipe %% of peak (elapsed) %.1f ; cycles elapsed max %.4g active avg %.4g' % (g('l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed'), g('sm__cycles_elapsed.max'), g('smsp__cycles_active.avg')))
    print(' inst %.4g  issue active %.1f%%  fma pipe %.1f%%  warps active %.1f%%' % (g('smsp__inst_executed.sum'), g('smsp__issue_active.avg.pct_of_peak_sustained_active'), g('sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active'), g('sm__warps_active.avg.pct_of_peak_sustained_active')))
    if steps: print(' inst per step %.0f' % (g('smsp__inst_executed.sum') / steps))
    for k in hdr:
        if k.startswith('smsp__average_warps_issue_stalled_') and k.endswith('_per_issue_active.ratio'):
            v = g(k)
            if v > 0.05: print('  stall %-22s %.2f' % (k[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')], v))

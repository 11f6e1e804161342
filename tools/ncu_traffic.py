#!/usr/bin/env python
"""Update profiles/ncu_traffic.json (DRAM bytes per audio-second of a kernel) and profiles/ncu_f0.json (issue / FP64 pipe
utilisation of the pYIN kernels) from .ncu-rep captures (read here, no GPU).
   python tools/ncu_traffic.py traffic rep.ncu-rep kernel_substr audio_seconds "capture label"
   python tools/ncu_traffic.py f0 rep.ncu-rep "capture label" """
import csv, io, json, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent


def rows_of(rep):
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(raw)))
    return r[0], r[1], r[2:]


def val(hdr, units, row, key):
    i = hdr.index(key)
    v = float(row[i].replace(',', ''))
    u = units[i].lower()
    return v * {'gbyte': 1e9, 'mbyte': 1e6, 'kbyte': 1e3, 'byte': 1.0}.get(u, 1.0)


mode = sys.argv[1]
if mode == 'traffic':
    rep, kern, audio_s, label = sys.argv[2], sys.argv[3], float(sys.argv[4]), sys.argv[5]
    hdr, units, rows = rows_of(rep)
    row = [r for r in rows if kern in r[hdr.index('Kernel Name')]][0]
    b = val(hdr, units, row, 'dram__bytes_read.sum') + val(hdr, units, row, 'dram__bytes_write.sum')
    p = ROOT / 'profiles' / 'ncu_traffic.json'
    d = json.loads(p.read_text()) if p.exists() else {}
    d[kern] = {'dram_bytes_per_audio_second': b / audio_s, 'dram_bytes_read': val(hdr, units, row, 'dram__bytes_read.sum'),
               'dram_bytes_write': val(hdr, units, row, 'dram__bytes_write.sum'), 'audio_seconds': audio_s, 'capture': label}
    p.write_text(json.dumps(d, indent=1) + '\n')
    print(kern, d[kern])
else:
    rep, label = sys.argv[2], sys.argv[3]
    hdr, units, rows = rows_of(rep)
    out = {'capture': label}
    for row in rows:
        name = row[hdr.index('Kernel Name')].split('(')[0].replace('void ', '').strip()
        g = lambda k: float(row[hdr.index(k)].replace(',', '')) if k in hdr and row[hdr.index(k)] not in ('', 'n/a') else None
        out[name] = {'ms': g('gpu__time_duration.sum'), 'issue_active_pct': g('smsp__issue_active.avg.pct_of_peak_sustained_active'),
                     'fp64_pipe_pct': g('sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active'),
                     'fma_pipe_pct': g('sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active'),
                     'alu_pipe_pct': g('sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active'),
                     'lsu_data_pipe_pct': g('l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed'),
                     'warps_active_pct': g('sm__warps_active.avg.pct_of_peak_sustained_active')}
    (ROOT / 'profiles' / 'ncu_f0.json').write_text(json.dumps(out, indent=1) + '\n')
    print(json.dumps(out, indent=1))

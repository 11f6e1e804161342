#!/usr/bin/env python
"""Shared-memory wavefronts per source line from an .ncu-rep captured with --import-source on (read here, no GPU):
   python tools/ncu_shared_lines.py rep.ncu-rep kernel_regex [units] [top_n]; units = divisor (e.g. warp steps in the launch)."""
import csv, subprocess, sys, io
rep, kern = sys.argv[1], sys.argv[2]
units = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass', '--kernel-name', 'regex:' + kern],
                     capture_output=True, text=True).stdout
cur, hdr, out, first = None, None, [], None
active = True
for r in csv.reader(io.StringIO(raw)):
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name':
        if first is None: first = r[1]
        active = (r[1] == first); continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr is None or not active or not r[0].isdigit(): continue
    iw, ie, ii, ix = hdr.index('L1 Wavefronts Shared'), hdr.index('L1 Wavefronts Shared Excessive'), hdr.index('L1 Wavefronts Shared Ideal'), hdr.index('Instructions Executed')
    try: out.append((float(r[iw] or 0), float(r[ie] or 0), float(r[ii] or 0), float(r[ix] or 0), cur, int(r[0]), r[1].strip()))
    except ValueError: pass
tot = sum(o[0] for o in out)
print(f'# {first}: shared wavefronts {tot:.4g} ({tot / units:.1f} per unit), excessive {sum(o[1] for o in out) / units:.1f} per unit')
for w, e, i, n, f, ln, src in sorted(out, reverse=True)[:top]:
    print(f'{f}:{ln:4d} wf {w / units:7.1f} excess {e / units:6.1f} inst {n / units:6.1f} | {src[:100]}')

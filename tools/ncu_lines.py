#!/usr/bin/env python
"""Per-source-line summary of an .ncu-rep captured with --import-source on (read here, no GPU):
   python tools/ncu_lines.py rep.ncu-rep kernel_regex [top_n] -> file:line, warp-instructions, share, stall samples, source."""
import csv, subprocess, sys, io
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass', '--kernel-name', 'regex:' + kern],
                     capture_output=True, text=True).stdout
cur_file, hdr, out, seen_kernel = None, None, [], None
for r in csv.reader(io.StringIO(raw)):
    if not r: continue
    if r[0] == 'File Path': cur_file = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name':
        if seen_kernel is None: seen_kernel = r[1]
        active = (r[1] == seen_kernel); continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr is None or not active or not r[0].isdigit(): continue
    ie = hdr.index('Instructions Executed'); ss = hdr.index('# Samples')
    try: out.append((int(r[ie]), int(r[ss]), cur_file, int(r[0]), r[1].strip()))
    except ValueError: pass
tot = sum(o[0] for o in out) or 1; tots = sum(o[1] for o in out) or 1
print(f'# {seen_kernel}: {tot} warp-instructions, {tots} samples')
for n, s, f, ln, src in sorted(out, reverse=True)[:top]:
    print(f'{f}:{ln:4d} {n:12d} {100*n/tot:5.1f}% inst {100*s/tots:5.1f}% samp | {src[:110]}')

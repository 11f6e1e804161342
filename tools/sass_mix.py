#!/usr/bin/env python
"""Instruction mix + stall samples per opcode from `ncu --page source --csv --print-source sass` output."""
import csv, sys, collections
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]
iS=hdr.index('Source'); iN=hdr.index('Instructions Executed'); iSm=hdr.index('# Samples')
mix=collections.defaultdict(lambda:[0,0,0])
tot=0; tots=0
for r in rows[2:]:
    if len(r)<=iSm: continue
    op=r[iS].split()
    if not op: continue
    o=op[0] if not op[0].startswith('@') else op[1]
    o=o.split('.')[0] + ('.'+'.'.join(o.split('.')[1:2]) if o.startswith(('LDS','STS','LDG','STG','SHFL','BAR','MUFU')) else '')
    try: n=int(r[iN]); s=int(r[iSm])
    except: continue
    mix[o][0]+=n; mix[o][1]+=s; mix[o][2]+=1; tot+=n; tots+=s
print(f"total warp-instr {tot}  samples {tots}  static {sum(v[2] for v in mix.values())}")
for k,v in sorted(mix.items(), key=lambda kv:-kv[1][0])[:40]:
    print(f"{k:14s} exec={v[0]:12d} {v[0]/tot:6.3f}  samples={v[1]:8d} {v[1]/max(tots,1):6.3f} static={v[2]}")

#!/bin/bash
set -u
mkdir -p gpurun_out
for mb in 4096 12288; do
AFX_F0_SCRATCH_MB=$mb timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-legs 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());f=d['f0_path'];print('scratch $mb MB: f0_path',round(f['value'],1),'audio-hours/s',round(f['ms'],2),'ms',{k:round(v,2) for k,v in f['phases_ms'].items()})"
done
timeout 600 python -m pytest tests -m gpu -x -q -k "pyin or f0" 2>&1 | tail -2

#!/bin/bash
set -u
mkdir -p gpurun_out
for v in 5 4 5 4; do   # 5 = the five-warp shape (half width 25 only), anything else = the default four-warp shape
AFX_VIT_WARPS=$v timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-legs 2>gpurun_out/ab_$v.err | python -c "
import json,sys;d=json.loads(sys.stdin.read());f=d['f0_path'];print('viterbi warps $v: f0_path',round(f['value'],1),'audio-hours/s',round(f['ms'],2),'ms',{k:round(v,2) for k,v in f['phases_ms'].items()})"
done
AFX_VIT_WARPS=4 timeout 900 python -m pytest tests -m gpu -x -q -k "hw25 or pyin_f0_track or parity_sweep" 2>&1 | tail -3

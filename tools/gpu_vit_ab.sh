#!/bin/bash
# A/B of the Viterbi scan variants (AFX_VITERBI=v1: single-pass scan) + the F0 parity tests
set -u
mkdir -p gpurun_out
for v in v1 v2 v1 v2; do
AFX_VITERBI=$v timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-legs 2>gpurun_out/ab_$v.err | python -c "
import json,sys;d=json.loads(sys.stdin.read());f=d['f0_path'];print('viterbi $v: f0_path',round(f['value'],1),'audio-hours/s',round(f['ms'],2),'ms',{k:round(v,2) for k,v in f['phases_ms'].items()})"
done
timeout 900 python -m pytest tests -m gpu -x -q -k "pyin or f0 or smoke or c3 or c4 or default" 2>&1 | tail -3

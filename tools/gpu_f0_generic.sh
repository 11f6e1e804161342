#!/bin/bash
# F0 path after the frame-mode (any hop) generalisation of yin_obs_kernel: parity tests + the default-path bench line
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "pyin or f0 or smoke or default" 2>&1 | tail -15
timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-legs 2>gpurun_out/f0.err | python -c "
import json,sys;d=json.loads(sys.stdin.read());f=d['f0_path'];print('f0_path',round(f['value'],1),'audio-hours/s',round(f['ms'],2),'ms',{k:round(v,2) for k,v in f['phases_ms'].items()}); print('value', d['value'], d['ms_per_step'])"

#!/bin/bash
# bench the headline config under a few environment settings (one line each)
set -u
mkdir -p gpurun_out
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-f0"
run() { env "$@" timeout 300 $B 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$*','ms_per_step',round(d['ms_per_step'],3),'clips_ok',d['config']['clips_ok'])"; }
run AFX_STREAM_WARPS=16
run AFX_STREAM_WARPS=20
run AFX_STREAM_WARPS=12
run AFX_STREAM_WARPS=16 AFX_ITEM_STEPS=16
run AFX_STREAM_WARPS=16 AFX_ITEM_STEPS=32

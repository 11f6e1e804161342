#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
B="python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu-baseline --no-f0 --no-legs"
run() { env "$@" timeout 300 $B $EXTRA 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$* $EXTRA','ms_per_step',round(d['ms_per_step'],3),'frac',round(d['roofline']['frac'],4),'clips_ok',d['clips_ok'])"; }
EXTRA="--config c2 --clips 10000"; run A=1
EXTRA="--config c3 --clips 8000"; run A=1
EXTRA="--config c4 --clips 4000"; run A=1
timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-f0 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print(json.dumps({k:{kk:vv for kk,vv in v.items() if kk in ('value','ms','hbm_frac','slowdown_vs_clean','clips')} for k,v in d['legs'].items() if isinstance(v,dict)},indent=0))"

#!/bin/bash
# bench the headline config under a few environment settings (one line each), then the full default bench line
set -u
mkdir -p gpurun_out
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-f0 --no-legs"
run() { env "$@" timeout 300 $B 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$*','ms_per_step',round(d['ms_per_step'],3),'clips_ok',d['clips_ok'])"; }
run AFX_SLAB_RING_MB=48
run AFX_SLAB_RING_MB=24
run AFX_SLAB_RING_MB=16
run AFX_SLAB_RING_MB=12
run AFX_SLAB_RING_MB=96
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "default bench exit $?"; tail -3 gpurun_out/bench_default.err; python -c "
import json;d=json.load(open('gpurun_out/bench_default.json'));print(json.dumps({k:d[k] for k in ('value','ms_per_step','e2e','f0_path','legs','cpu_baseline')},indent=1)[:6000])"

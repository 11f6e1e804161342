#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "native exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench_native.json'));print(json.dumps({k:d[k] for k in ('value','ms_per_step','clips_ok','gpu_launches')}));print('frac',d['roofline']['frac']);print('e2e',d['e2e']['value']);print('f0',d['f0_path']['value'],d['f0_path']['phases_ms']);print(json.dumps({k:(v.get('value'),v.get('ms')) for k,v in d['legs'].items()}))"
for c in c3 c4; do python bench.py --config $c --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-f0 --no-legs 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$c', round(d['ms_per_step'],3),'ms', round(d['value'],1), 'frac', round(d['roofline']['frac'],4), d['config']['workload'][:60])"; done

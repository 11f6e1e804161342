#!/bin/bash
# what the driver runs at round end: gpu tests, smoke, reference arm, native arm
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"; cut -c1-300 gpurun_out/bench_reference.json
timeout 900 python bench.py > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "native exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench_native.json'));print(json.dumps({k:d[k] for k in ('value','ms_per_step','clips_ok','gpu_launches')}));print('frac',d['roofline']['frac']);print('e2e',d['e2e']['value']);print('f0',d['f0_path']['value'],d['f0_path']['phases_ms']);print(json.dumps({k:(v.get('value'),v.get('ms')) for k,v in d['legs'].items()}))"

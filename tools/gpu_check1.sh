#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu.log
python tools/h2d_probe_ranks.py > gpurun_out/h2d_1.json 2>/dev/null; cat gpurun_out/h2d_1.json
timeout 600 python bench.py --no-legs --no-f0 > gpurun_out/bench_1.json 2> gpurun_out/bench_1.err; echo "bench exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench_1.json'));print({k:d[k] for k in ('value','ms_per_step','e2e')})"

#!/bin/bash
# quick GPU check: parity tests + headline bench (native arm)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 ${BENCH_ARGS:-} > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err
echo "bench exit $?"; cat gpurun_out/bench_native.json; tail -5 gpurun_out/bench_native.err

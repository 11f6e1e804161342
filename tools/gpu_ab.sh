#!/bin/bash
# GPU check of the streaming kernel: parity tests, then the headline bench for the old (cta) and new (stream) kernels
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-f0"
AFX_CLIP_IMPL=cta timeout 300 $B > gpurun_out/bench_cta.json 2> gpurun_out/bench_cta.err; echo "cta exit $?"; cat gpurun_out/bench_cta.json
timeout 300 $B > gpurun_out/bench_stream.json 2> gpurun_out/bench_stream.err; echo "stream exit $?"; cat gpurun_out/bench_stream.json; tail -3 gpurun_out/bench_stream.err
for k in 4 16 32; do AFX_ITEM_STEPS=$k timeout 300 $B > gpurun_out/bench_stream_k$k.json 2>/dev/null; echo "k=$k"; cat gpurun_out/bench_stream_k$k.json; done

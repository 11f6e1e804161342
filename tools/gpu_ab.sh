#!/bin/bash
# GPU check of the streaming kernel: parity tests, the headline bench, and (optional) an ncu capture
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-f0"
timeout 300 $B > gpurun_out/bench_stream.json 2> gpurun_out/bench_stream.err; echo "stream exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench_stream.json'));print('ms_per_step',d['ms_per_step'],'frac',d['roofline']['frac'],'clips_ok',d['clips_ok'])"; tail -3 gpurun_out/bench_stream.err
if [ "${NCU:-0}" = "1" ]; then
SMALL="python bench.py --clips ${NCU_CLIPS:-2368} --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-f0"
$SMALL > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 1 -c 1 -f -o gpurun_out/prof_stream_kernel $SMALL > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
fi

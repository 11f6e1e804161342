#!/bin/bash
# parity tests + F0-path bench (2000 clips): prints throughput and per-phase device times
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python bench.py --features f0,mfcc,energy --clips ${F0_CLIPS:-2000} --steps 3 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/bench_f0.json 2> gpurun_out/bench_f0.err
python -c "
import json; d=json.load(open('gpurun_out/bench_f0.json')); print(d['value'], d['ms_per_step'], d['roofline']['phases'])"; tail -3 gpurun_out/bench_f0.err

#!/bin/bash
# r02m: GPU test suite after the container restore + source-level ncu captures of both pYIN kernels
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/pytest_gpu.log
F0="python bench.py --clips 1480 --steps 1 --warmup 1 --features f0,mfcc,energy --no-e2e --no-cpu-baseline --no-legs"
$F0 > gpurun_out/plain_f0.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"viterbi_kernel" -s 0 -c 1 -f -o gpurun_out/prof_vit $F0 > gpurun_out/ncu_vit.log 2>&1; echo "ncu vit exit $?"
ncu --set full --clock-control none --import-source on -k regex:"yin_obs_kernel" -s 0 -c 1 -f -o gpurun_out/prof_yin $F0 > gpurun_out/ncu_yin.log 2>&1; echo "ncu yin exit $?"
tail -1 gpurun_out/plain_f0.log | cut -c1-600

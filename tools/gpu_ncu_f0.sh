#!/bin/bash
set -u
mkdir -p gpurun_out
F0="python bench.py --clips 1480 --steps 1 --warmup 1 --features f0,mfcc,energy --no-e2e --no-cpu-baseline --no-legs"
$F0 > gpurun_out/plain_f0.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"viterbi_kernel" -s 1 -c 1 -f -o gpurun_out/prof_vit $F0 > gpurun_out/ncu_vit.log 2>&1; echo "ncu vit exit $?"

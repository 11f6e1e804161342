#!/bin/bash
# ncu --set full capture of the two pYIN kernels on a small F0 workload
set -u
mkdir -p gpurun_out
SMALL="python bench.py --clips ${NCU_CLIPS:-296} --features f0,mfcc,energy --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
$SMALL > gpurun_out/plain_f0.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"yin_obs|viterbi" -s 2 -c 2 -f -o gpurun_out/prof_f0 $SMALL > gpurun_out/ncu_f0.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_f0.log

#!/bin/bash
# ncu --set full capture of the two DTW kernels on a small batch (after the same command ran clean without ncu)
set -u
mkdir -p gpurun_out
SMALL="python tools/bench_dtw.py ${NCU_PAIRS:-592} 862"
$SMALL > gpurun_out/plain_dtw.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"dtw_distance" -s 3 -c 1 -f -o gpurun_out/prof_dtw_dist $SMALL > gpurun_out/ncu_dtw.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"dtw_accumulate" -s 3 -c 1 -f -o gpurun_out/prof_dtw_acc $SMALL >> gpurun_out/ncu_dtw.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_dtw.log

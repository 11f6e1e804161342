#!/bin/bash
# final-round profile set: launch list + ncu --set full of the streaming kernel (C2), the pYIN kernels, plain bench lines
set -u
mkdir -p gpurun_out
SMALL="python bench.py --clips 2368 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-f0 --no-legs"
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1; echo "launch list exit $?"
$SMALL > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 1 -c 1 -f -o gpurun_out/prof_stream_kernel $SMALL > gpurun_out/ncu_full.log 2>&1; echo "ncu stream exit $?"

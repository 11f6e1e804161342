#!/bin/bash
# plain run, then launch list + one ncu --set full capture of the kernel matching $1 (default clip_kernel)
set -u
K=${1:-clip_kernel}
mkdir -p gpurun_out
SMALL="python bench.py --clips ${NCU_CLIPS:-1184} --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
$SMALL > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 -f -o gpurun_out/prof_$K $SMALL > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
cat gpurun_out/plain.log | tail -2

#!/bin/bash
# One gpurun call: GPU parity tests, the bench (both arms), an ncu launch list and one ncu --set full
# capture of the dominant kernel.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err
echo "bench exit $?"; cat gpurun_out/bench_native.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
echo "ref exit $?"; cat gpurun_out/bench_reference.json
SMALL="python bench.py --clips 1024 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
$SMALL > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:frames_kernel -s 1 -c 2 -f -o gpurun_out/prof_frames $SMALL > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la gpurun_out

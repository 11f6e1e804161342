#!/bin/bash
# final set (r02m, re-run as r02n after the last kernel changes): what the driver runs at round end (GPU tests, smoke, both bench arms) + launch list of the bench command
# + ncu --set full of the final pYIN kernels and of the streaming kernel
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"; cut -c1-200 gpurun_out/bench_reference.json
timeout 900 python bench.py > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "native exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench_native.json'));print(json.dumps({k:d[k] for k in ('value','ms_per_step','clips_ok','gpu_launches')}));print('frac',d['roofline']['frac']);print('e2e',d['e2e']['value']);print('f0',d['f0_path']['value'],d['f0_path']['phases_ms']);print(json.dumps({k:(v.get('value'),v.get('ms')) for k,v in d['legs'].items()}))"
SMALL="python bench.py --clips 2368 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-legs --f0-clips 740"
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
F0="python bench.py --clips 1480 --steps 1 --warmup 1 --features f0,mfcc,energy --no-e2e --no-cpu-baseline --no-legs"
$F0 > gpurun_out/plain_f0.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"yin_obs_kernel|viterbi_kernel" -s 0 -c 2 -f -o gpurun_out/prof_pyin $F0 > gpurun_out/ncu_f0.log 2>&1; echo "ncu pyin exit $?"
S2="python bench.py --clips 2368 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-f0 --no-legs"
$S2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none -k regex:stream_kernel -s 1 -c 1 -f -o gpurun_out/prof_stream_kernel $S2 > gpurun_out/ncu_full.log 2>&1; echo "ncu stream exit $?"

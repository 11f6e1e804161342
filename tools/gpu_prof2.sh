#!/bin/bash
set -u
mkdir -p gpurun_out
F0="python bench.py --clips 1480 --steps 1 --warmup 1 --features f0,mfcc,energy --no-e2e --no-cpu-baseline --no-legs"
$F0 > gpurun_out/plain_f0.log 2>&1 &&
ncu --set full --clock-control none -k regex:"yin_obs_kernel|viterbi_kernel" -s 2 -c 2 -f -o gpurun_out/prof_pyin $F0 > gpurun_out/ncu_f0.log 2>&1; echo "ncu pyin exit $?"
for c in c3 c4; do
  S="python bench.py --config $c --clips 1480 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-f0 --no-legs"
  $S > gpurun_out/plain_$c.log 2>&1 &&
  ncu --set full --clock-control none -k regex:stream_kernel -s 1 -c 1 -f -o gpurun_out/prof_stream_$c $S > gpurun_out/ncu_$c.log 2>&1; echo "ncu $c exit $?"
  python bench.py --config $c --clips 8000 --no-e2e --no-cpu-baseline --no-legs > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "bench $c exit $?"
done

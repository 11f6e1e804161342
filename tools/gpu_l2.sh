#!/bin/bash
# L2 eviction-hint experiment of the streaming kernel: time (plain bench) and DRAM bytes (ncu, three metrics) per mode
set -u
mkdir -p gpurun_out
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-f0 --no-legs"
for m in 0 1 2 3; do
  AFX_L2_MODE=$m timeout 300 $B 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('AFX_L2_MODE=$m ms_per_step',round(d['ms_per_step'],3))"
done
S="python bench.py --clips 2368 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-f0 --no-legs"
$S > gpurun_out/plain_l2.log 2>&1 || exit 1
for m in 0 1 2 3; do
  AFX_L2_MODE=$m ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:stream_kernel -s 1 -c 1 --csv --log-file gpurun_out/l2_mode_$m.csv $S > /dev/null 2>&1
  echo "mode $m"; grep -E "dram__bytes|gpu__time" gpurun_out/l2_mode_$m.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}'
done
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"stream_kernel|item_build|header_fetch" -c 12 --csv --log-file gpurun_out/launches.csv $S > /dev/null 2>&1; echo "launch list exit $?"

#!/bin/bash
# N-GPU check (N = number of visible GPUs): sharded batch_process test over NCCL, concurrent H2D probe, bench at N
set -u
mkdir -p gpurun_out
N=$(python -c "import torch;print(torch.cuda.device_count())")
echo "GPUs: $N"
timeout 600 python -m pytest tests -m gpu -x -q -k "sharded" > gpurun_out/pytest_sharded_${N}gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_sharded_${N}gpu.log
for w in 1 2 4 8; do
  if [ $w -le $N ]; then
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $w --master-addr 127.0.0.1 --master-port 29511 tools/h2d_probe_ranks.py 2>/dev/null | grep aggregate > gpurun_out/h2d_$w.json; cat gpurun_out/h2d_$w.json
  fi
done
for w in 2 4 8; do
  if [ $w -le $N ]; then
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $w --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $w --steps 5 --warmup 3 > gpurun_out/bench_${w}gpu.json 2> gpurun_out/bench_${w}gpu.err; echo "bench $w exit $?"
    python -c "
import json;d=json.load(open('gpurun_out/bench_${w}gpu.json'));print({k:d[k] for k in ('n_gpus','value','ms_per_step')}, d['e2e']['value'], d['e2e']['clips_per_gpu'])"
  fi
done
if [ $N -ge 8 ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --clips 45000 --steps 3 --warmup 3 --no-e2e --no-f0 --no-legs > gpurun_out/bench_c5_1000h_8gpu.json 2> gpurun_out/bench_c5_8gpu.err; echo "c5 exit $?"
  python -c "
import json;d=json.load(open('gpurun_out/bench_c5_1000h_8gpu.json'));print({k:d[k] for k in ('n_gpus','value','ms_per_step','clips_ok')})"
fi

#!/bin/bash
# 8-GPU record of the final round-2 state: the 2-rank NCCL test of batch_process_sharded, the bench at N = 8 and the
# 1000-audio-hour configuration (BASELINE.json configs[4])
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "sharded" > gpurun_out/pytest_sharded_8gpu.txt 2>&1; echo "sharded exit $?"; tail -2 gpurun_out/pytest_sharded_8gpu.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_8gpu.json 2> gpurun_out/bench_8gpu.err; echo "bench 8 exit $?"; cut -c1-250 gpurun_out/bench_8gpu.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --clips 45000 --steps 3 --warmup 3 --no-e2e --no-f0 --no-legs > gpurun_out/bench_c5_1000h_8gpu.json 2> gpurun_out/bench_c5_8gpu.err; echo "c5 exit $?"; cut -c1-250 gpurun_out/bench_c5_1000h_8gpu.json

"""Host->device copy rate with N ranks copying AT ONCE from pinned memory (torchrun, one rank per GPU): the ceiling of the
e2e leg of bench.py at N GPUs.  Rank 0 prints one JSON line; append it to profiles/h2d_probe.json under the key str(N).
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/h2d_probe_ranks.py"""
import json, os, time
import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); h.random_(0, 255)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for _ in range(2):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
reps = 6
t0 = time.perf_counter()
for _ in range(reps):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
t = torch.tensor([dt], dtype=torch.float64, device="cuda")
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    agg = world * reps * n / float(t.item()) / 1e9
    print(json.dumps({"ranks": world, "aggregate_gbs": agg, "per_rank_gbs": agg / world,
                      "source": f"tools/h2d_probe_ranks.py: {world} rank(s) x {reps} x 1 GiB pinned->device at once, slowest rank's wall time"}))
if world > 1:
    dist.destroy_process_group()

"""Multi-rank host logic on CPU: world_size-2 gloo job, the native extraction replaced by a deterministic fake
(the CUDA path itself is covered by the -m gpu tests; there is no collective inside the hot loop to test)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from audio_feature_extraction_b200.distributed import extract_sharded, gather_rows, shard_range


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 10000, 360000):
        for world in (1, 2, 3, 8):
            cuts = [shard_range(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _fake_rows(items):
    # "summary" of item i: [i, i^2, 3i]; meta: [status = i % 3 == 0, i, i + 1, 9]
    ids = np.asarray(items, dtype=np.int64)
    summ = torch.tensor(np.stack([ids, ids * ids, 3 * ids], axis=1).reshape(-1, 3), dtype=torch.float32)
    meta = torch.tensor(np.stack([(ids % 3 == 0).astype(np.int64), ids, ids + 1, np.full_like(ids, 9)], axis=1).reshape(-1, 4),
                        dtype=torch.int32)
    return summ, meta


def _worker(rank, world, port, n_items, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        items = list(range(n_items))
        summ, meta = extract_sharded(_fake_rows, items)
        if rank == 0:
            ref_s, ref_m = _fake_rows(items)
            assert torch.equal(summ, ref_s) and torch.equal(meta, ref_m)
            torch.save({"ok": True, "rows": summ.shape[0]}, out_path)
        else:
            assert summ is None and meta is None
        # uneven heights incl. an empty shard
        local = torch.full((rank * 2, 5), float(rank))
        g = gather_rows(local)
        if rank == 0:
            assert g.shape == (sum(r * 2 for r in range(world)), 5)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_items", [11, 1])
def test_two_rank_gloo_gather_matches_single_process(tmp_path, n_items):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = tmp_path / "r0.pt"
    mp.spawn(_worker, args=(2, port, n_items, str(out)), nprocs=2, join=True)
    res = torch.load(out)
    assert res["ok"] and res["rows"] == n_items


def test_parse_cpulist_and_numa_binding_is_harmless_without_a_gpu():
    from audio_feature_extraction_b200.distributed import bind_host_to_gpu_node, parse_cpulist
    assert parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    assert parse_cpulist("") == [] and parse_cpulist("5") == [5]
    import os
    before = os.sched_getaffinity(0)
    info = bind_host_to_gpu_node(0)          # no GPU / no sysfs entry here: reports, never raises, changes nothing
    assert isinstance(info, dict) and info["device"] == 0
    assert os.sched_getaffinity(0) == before or info.get("cpus_bound")

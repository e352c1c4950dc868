"""Multi-GPU host logic on CPU: shard arithmetic and the final gather over gloo, world_size 2."""
import os
import socket

import numpy as np
import pytest

from aslr_to_b200 import sharding as S
from aslr_to_b200 import workloads as W


def test_shard_bounds_partition_the_batch():
    for B in (0, 1, 7, 4096, 4097, 1023):
        for world in (1, 2, 3, 4, 8):
            blocks = [S.shard_bounds(B, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == B
            assert all(blocks[r][1] == blocks[r + 1][0] for r in range(world - 1))
            sizes = S.shard_sizes(B, world)
            assert max(sizes) - min(sizes) <= 1 and sum(sizes) == B
    with pytest.raises(ValueError):
        S.shard_bounds(8, 2, 2)


def test_shard_workload_slices_per_ocp_arrays_only():
    w = W.sea_batch(10)
    parts = [S.shard_workload(w, 3, r) for r in range(3)]
    assert np.array_equal(np.concatenate([p["x0"] for p in parts]), w["x0"])
    assert np.array_equal(np.concatenate([p["K_ocp"] for p in parts]), w["K_ocp"])
    assert all(p["T"] == w["T"] and p["running"] is w["running"] for p in parts)


def _worker(rank, world, port, B, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        w = S.shard_workload(W.sea_batch(B), world, rank)
        lo, hi = w["shard"]
        x0 = torch.from_numpy(np.ascontiguousarray(w["x0"]))
        # stand-in for the solve (the kernels need a GPU): per-OCP results that identify the OCP
        local = {"cost": x0[:, 0] * 3.0 + x0[:, 1], "iters": torch.arange(lo, hi, dtype=torch.int32),
                 "xs": x0[:, None, :].repeat(1, 4, 1)}
        full = S.gather_results(local, B)
        q.put((rank, {k: v.numpy() for k, v in full.items()}))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("B", [8, 9])
def test_gather_results_over_gloo_world_size_2(B):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, B, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    x0 = W.sea_batch(B)["x0"]
    for rank in (0, 1):
        full = res[rank]
        assert np.array_equal(full["iters"], np.arange(B))
        assert np.array_equal(full["cost"], x0[:, 0] * 3.0 + x0[:, 1])
        assert np.array_equal(full["xs"], np.repeat(x0[:, None, :], 4, axis=1))

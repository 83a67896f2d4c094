"""Host logic of the N>1 path, on CPU with the gloo backend (world size 2)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from xpano_b200 import sharding


def test_assign_groups_partitions_exactly():
    for n_groups in (0, 1, 7, 64):
        for world in (1, 2, 4, 8):
            seen = []
            for r in range(world):
                seen += sharding.assign_groups(n_groups, world, r)
            assert sorted(seen) == list(range(n_groups))
            assert sharding.groups_per_rank(n_groups, world) == [
                len(sharding.assign_groups(n_groups, world, r))
                for r in range(world)]
    with pytest.raises(ValueError):
        sharding.assign_groups(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        groups = sharding.assign_groups(5, world, rank)
        # stand-in for the per-group export: 10 MP per group, rank 1 slower
        mp_local = 10.0 * len(groups)
        ms_local = 100.0 * (rank + 1)
        r = sharding.job_throughput(mp_local, ms_local)
        all_groups = [None] * world
        dist.all_gather_object(all_groups, groups)
        if rank == 0:
            out.put((r, all_groups))
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_reduce():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q))
             for r in range(2)]
    for p in procs:
        p.start()
    r, groups = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert groups == [[0, 2, 4], [1, 3]]
    assert r["megapixels"] == 50.0 and r["ms"] == 200.0
    assert abs(r["mp_per_s"] - 250.0) < 1e-9

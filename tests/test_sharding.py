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


def test_strip_bounds_partition_the_roi():
    for x, w in ((-505, 1009), (0, 7), (-9303, 18606), (12, 100001)):
        for world in (1, 2, 3, 4, 8):
            edges = [sharding.strip_bounds(x, w, world, r)
                     for r in range(world)]
            assert edges[0][0] == x and edges[-1][1] == x + w
            for a, b in zip(edges, edges[1:]):
                assert a[1] == b[0] and a[0] <= a[1]
    with pytest.raises(ValueError):
        sharding.strip_bounds(0, 10, 2, 2)
    assert sharding.tile_halo(5) == 8 * 32 and sharding.tile_halo(7) == 1024


def _tile_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import numpy as np
        # the exchange of a canvas-tiled run: 8 partial sums per rank
        part = [float(rank + 1) * (k + 1) for k in range(8)]
        red = sharding.all_reduce_sum(part)
        x0, x1 = sharding.strip_bounds(-100, 333, world, rank)
        strip = np.full((4, x1 - x0, 3), rank + 1, np.uint8)
        mask = np.full((4, x1 - x0), 255, np.uint8)
        pano, full = sharding.gather_strips(strip, mask, x0 + 100, 333)
        if rank == 0:
            out.put((red, pano[0, :, 0].tolist(), int(full.sum())))
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_tile_exchange():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_tile_worker, args=(r, 2, port, q))
             for r in range(2)]
    for p in procs:
        p.start()
    red, row, msum = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert red == [3.0 * (k + 1) for k in range(8)]
    assert row == [1] * 166 + [2] * 167 and msum == 255 * 4 * 333

"""Multi-GPU sharding of the export path: by panorama group.

Panoramas are independent jobs in the reference (`StitcherData.panos`,
xpano/pipeline/stitcher_pipeline.h:48-52; one `Stitch()` per panorama), so
N GPUs take disjoint sets of groups and there is no data-path collective: the
only communication is the reduction of the per-rank counters that `bench.py`
reports (pixels produced, max elapsed time). One process per GPU, rendezvous
through `torch.distributed` (NCCL on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Sequence


def assign_groups(n_groups: int, world: int, rank: int) -> List[int]:
    """Round-robin assignment: group g goes to rank g % world."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} of {world}")
    return list(range(rank, n_groups, world))


def groups_per_rank(n_groups: int, world: int) -> List[int]:
    return [len(range(r, n_groups, world)) for r in range(world)]


def _tensor(values: Sequence[float], device):
    import torch
    return torch.tensor(list(values), dtype=torch.float64, device=device)


def _device():
    """Reduction buffers live where the backend wants them."""
    import torch
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def reduce_sum(value: float) -> float:
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = _tensor([value], _device())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def reduce_max(value: float) -> float:
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = _tensor([value], _device())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def job_throughput(local_megapixels: float, local_ms: float) -> dict:
    """Whole-job throughput: all pixels produced by all ranks over the slowest
    rank's time (the contract of bench.py)."""
    total_mp = reduce_sum(local_megapixels)
    ms = reduce_max(local_ms)
    return {"megapixels": total_mp, "ms": ms,
            "mp_per_s": total_mp / (ms / 1e3) if ms > 0 else 0.0}

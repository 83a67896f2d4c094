"""Multi-GPU sharding of the export path: by panorama group, and by canvas
strips of one panorama.

Panoramas are independent jobs in the reference (`StitcherData.panos`,
xpano/pipeline/stitcher_pipeline.h:48-52; one `Stitch()` per panorama), so
N GPUs take disjoint sets of groups and there is no data-path collective: the
only communication is the reduction of the per-rank counters that `bench.py`
reports (pixels produced, max elapsed time). One process per GPU, rendezvous
through `torch.distributed` (NCCL on the GPU box, gloo in the CPU tests).
"""
# Canvas strips (BASELINE.json configs[2]): rank r of N owns the
# columns [x + w*r/N, x + w*(r+1)/N) of the ROI rectangle (x, w) that
# ComputeRoi (stitcher.cc:144-169) gives; seams are computed for the whole
# canvas on every rank, the pyramids / composite / collapse for the strip plus
# a halo of tile_halo(levels) pixels that is recomputed locally, and the only
# exchange is an all-reduce of the 8 partial sums of the offset correction
# (`all_reduce_sum`; xp_composer_run_tile_begin / _end).
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


def strip_bounds(roi_x: int, roi_w: int, world: int, rank: int):
    """Columns of the ROI rectangle that rank `rank` of `world` owns (same
    integer arithmetic as the library, csrc/compose.cu make_plan)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} of {world}")
    return (roi_x + roi_w * rank // world, roi_x + roi_w * (rank + 1) // world)


def tile_halo(levels: int) -> int:
    """Halo (canvas pixels, each side) a strip recomputes: 8 * 2^levels."""
    from . import _lib
    return int(_lib.load().xp_tile_halo(int(levels)))


def all_reduce_sum(values: Sequence[float]) -> List[float]:
    """Sum of a short vector of doubles over the ranks (NCCL on the GPU box,
    gloo in the CPU tests); identity without a process group."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return [float(v) for v in values]
    t = _tensor(values, _device())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(v) for v in t.tolist()]


def gather_strips(strip, mask, x0: int, canvas_w: int):
    """Assemble the panorama on rank 0 from the strips of all ranks (host
    arrays): -> (pano, mask) on rank 0, (None, None) elsewhere."""
    import numpy as np
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return strip, mask
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, (int(x0), strip, mask))
    if dist.get_rank() != 0:
        return None, None
    h = strip.shape[0]
    pano = np.zeros((h, canvas_w, 3), strip.dtype)
    full = np.zeros((h, canvas_w), np.uint8)
    for px0, ps, pm in parts:
        pano[:, px0:px0 + ps.shape[1]] = ps
        full[:, px0:px0 + pm.shape[1]] = pm
    return pano, full

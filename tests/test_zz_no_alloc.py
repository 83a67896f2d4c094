"""GPU test of the allocation-free entry point xp_warper_warp_noalloc
(no counterpart in the reference): its
output equals xp_warper_warp's on the same device buffers, on two calls in a
row that share the static scratch array, and sizes beyond the stated maximum
are rejected with a clear error. Collected last on purpose (pytest -x)."""
import ctypes
import math
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import synth  # noqa: E402

pytestmark = pytest.mark.gpu


def _rot(yaw, pitch):
    cy, sy, cp, sp = math.cos(yaw), math.sin(yaw), math.cos(pitch), \
        math.sin(pitch)
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    return (Ry @ Rx).astype(np.float32)


def _f9(m):
    return np.ascontiguousarray(m, np.float32).reshape(9)


def _both(xb, torch, proj, scale, src, K, R, interp, border):
    """-> (existing entry point's result, no-alloc entry point's result,
    corner), both through the C ABI on device buffers."""
    lib = xb._lib.load()
    w = xb.RotationWarper(proj, scale)
    h_, w_ = src.shape[:2]
    cn = 1 if src.ndim == 2 else src.shape[2]
    depth = 8 * src.dtype.itemsize
    x, y, W, H = w.warpRoi((w_, h_), K, R)
    tdt = torch.uint8 if depth == 8 else torch.uint16
    d_src = torch.from_numpy(np.ascontiguousarray(src)).cuda()
    d_a = torch.zeros((H, W * cn), dtype=tdt, device="cuda")
    d_b = torch.zeros((H, W * cn), dtype=tdt, device="cuda")
    k, r = _f9(K), _f9(R)
    ca, cb = (ctypes.c_int * 2)(), (ctypes.c_int * 2)()
    torch.cuda.synchronize()
    xb._lib.check(lib.xp_warper_warp(
        w._h, d_src.data_ptr(), w_, h_, 0, depth, cn, k.ctypes.data,
        r.ctypes.data, interp, border, d_a.data_ptr(), 0, xb._lib.MEM_DEVICE,
        ca))
    stream = torch.cuda.current_stream().cuda_stream
    xb._lib.check(lib.xp_warper_warp_noalloc(
        w._h, d_src.data_ptr(), w_, h_, 0, depth, cn, k.ctypes.data,
        r.ctypes.data, interp, border, d_b.data_ptr(), 0, stream, cb))
    torch.cuda.synchronize()
    assert tuple(ca) == tuple(cb) == (x, y)
    return d_a.cpu().numpy(), d_b.cpu().numpy(), (x, y)


def test_noalloc_warp_equals_warp_twice_in_a_row():
    import torch
    import xpano_b200 as xb
    K = np.array([[420, 0, 250], [0, 415, 170], [0, 0, 1]], np.float32)
    # two calls in a row on the shared static scratch: different projection,
    # size, depth and rotation, so stale table entries would show
    src8 = synth.texture(340, 500, 71, 3, 8)
    a, b, _ = _both(xb, torch, xb.ProjectionType.kSpherical, 400.0, src8, K,
                    _rot(0.4, -0.1), xb.INTER_LINEAR, xb.BORDER_REFLECT)
    assert a.any() and np.array_equal(a, b)
    src16 = synth.texture(300, 260, 72, 3, 16)
    a, b, _ = _both(xb, torch, xb.ProjectionType.kCylindrical, 330.0, src16,
                    K, _rot(-0.3, 0.05), xb.INTER_LINEAR, xb.BORDER_REFLECT)
    assert a.any() and np.array_equal(a, b)
    # the mask warp of stitcher.cc:382-385 (single channel, nearest)
    mask = np.full((340, 500), 255, np.uint8)
    a, b, _ = _both(xb, torch, xb.ProjectionType.kSpherical, 400.0, mask, K,
                    _rot(0.4, -0.1), xb.INTER_NEAREST, xb.BORDER_CONSTANT)
    assert a.any() and np.array_equal(a, b)


def test_noalloc_warp_rejects_what_it_cannot_hold():
    import torch
    import xpano_b200 as xb
    lib = xb._lib.load()
    K = np.array([[420, 0, 250], [0, 415, 170], [0, 0, 1]], np.float32)
    k, r = _f9(K), _f9(np.eye(3))
    d = torch.zeros(500 * 340 * 3, dtype=torch.uint8, device="cuda")
    c = (ctypes.c_int * 2)()
    # scale 5e4: the ROI of a 500x340 frame is ~59000 x 40000 pixels
    big = xb.RotationWarper(xb.ProjectionType.kSpherical, 5.0e4)
    st = lib.xp_warper_warp_noalloc(big._h, d.data_ptr(), 500, 340, 0, 8, 3,
                                    k.ctypes.data, r.ctypes.data, 1, 2,
                                    d.data_ptr(), 0, None, c)
    assert st == 1 and b"XP_NOALLOC_MAX_ROI_PERIMETER" in lib.xp_last_error()
    merc = xb.RotationWarper(xb.ProjectionType.kMercator, 300.0)
    st = lib.xp_warper_warp_noalloc(merc._h, d.data_ptr(), 500, 340, 0, 8, 3,
                                    k.ctypes.data, r.ctypes.data, 1, 2,
                                    d.data_ptr(), 0, None, c)
    assert st == 3 and b"spherical and cylindrical" in lib.xp_last_error()

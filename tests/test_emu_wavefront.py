"""CPU-side verification of the wavefront kernels' source.

tests/emu/wavefront_emu.cpp compiles the SAME headers the GPU kernels are
built from (xpano_b200/csrc/wavefront.h, scan_problems.h) with g++ and runs
them on an emulated warp (32 OS threads, barrier shuffles). Results must be
bit-identical to the oracle's raster-order loops (oracle/mb_oracle.c, which is
pinned against the reference's goldens): seam labels and both distance maps,
and the inpainted pixels.
"""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import synth  # noqa: E402
from oracle import mb_oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "emu", "wavefront_emu.cpp")
LIB = os.path.join(HERE, "emu", "libwavefront_emu.so")
CSRC = os.path.join(os.path.dirname(HERE), "xpano_b200", "csrc")


@pytest.fixture(scope="module")
def emu():
    deps = [SRC] + [os.path.join(CSRC, f) for f in
                    ("wavefront.h", "scan_problems.h", "hd.h")]
    if (not os.path.exists(LIB)
            or os.path.getmtime(LIB) < max(os.path.getmtime(d) for d in deps)):
        subprocess.check_call(["g++", "-O2", "-std=c++20", "-pthread", "-fPIC",
                               "-shared", SRC, "-o", LIB])
    lib = ctypes.CDLL(LIB)
    vp, ci = ctypes.c_void_p, ctypes.c_int
    lib.emu_seams.argtypes = [ci, vp, vp, vp, vp, vp, ci, ci, ci, ci, vp, vp,
                              vp, ci]
    lib.emu_inpaint.argtypes = [ci, vp, vp, ci, vp, ci]
    return lib


def _seam_inputs(seed, n, mode, w=70, h=50):
    ims, pos = synth.blend_case(seed, n=n, w=w, h=h, mode=mode)
    ops, tpos = [], []
    for im, p in zip(ims, pos):
        bb = mb_oracle.trim(im)
        l, t, ww, hh = bb
        ops.append(np.ascontiguousarray(
            (im[t:t + hh, l:l + ww, 3] == 255).astype(np.uint8)))
        tpos.append((p[0] + l, p[1] + t))
    mx = min(p[0] for p in tpos)
    my = min(p[1] for p in tpos)
    tpos = [(p[0] - mx, p[1] - my) for p in tpos]
    W = max(p[0] + o.shape[1] for p, o in zip(tpos, ops))
    H = max(p[1] + o.shape[0] for p, o in zip(tpos, ops))
    return ops, tpos, W, H


def _emu_seams(lib, ops, tpos, W, H, reverse, wide, n_warps):
    n = len(ops)
    ia = lambda v: (ctypes.c_int * n)(*[int(x) for x in v])
    ptrs = (ctypes.c_void_p * n)(*[o.ctypes.data for o in ops])
    labels = np.zeros((H, W), np.int32)
    bwd = np.zeros((H, W), np.uint64)
    fwd = np.zeros((H, W), np.uint64)
    lib.emu_seams(n, ia(p[0] for p in tpos), ia(p[1] for p in tpos),
                  ia(o.shape[1] for o in ops), ia(o.shape[0] for o in ops),
                  ptrs, W, H, int(reverse), int(wide), labels.ctypes.data,
                  bwd.ctypes.data, fwd.ctypes.data, n_warps)
    return labels, bwd, fwd


@pytest.mark.parametrize("seed,n,mode,reverse,wide,n_warps", [
    (1, 2, "ellipse", 0, 0, 2),
    (2, 3, "blobs", 0, 0, 3),
    (3, 4, "ellipse", 1, 1, 2),
    (4, 3, "full", 0, 0, 1),
    (5, 5, "blobs", 1, 0, 4),
    (6, 4, "blobs", 0, 2, 2),
    (7, 3, "ellipse", 1, 2, 3),
])
def test_emulated_seams_match_oracle(emu, seed, n, mode, reverse, wide,
                                     n_warps):
    _check_seams(emu, seed, n, mode, reverse, wide, n_warps, 70, 50)


@pytest.mark.parametrize("seed,n,mode,reverse,wide,n_warps", [
    (21, 3, "full", 0, 0, 3),
    (22, 3, "ellipse", 1, 0, 2),
])
def test_emulated_seams_seed_only_chunks(emu, seed, n, mode, reverse, wide,
                                         n_warps):
    """Images wide enough that whole 32-step chunks consist of seed pixels and
    take the wavefront's parallel fast path (wavefront.h) between ordinary
    chunks."""
    _chunk_stats(emu)
    _check_seams(emu, seed, n, mode, reverse, wide, n_warps, 260, 100)
    fast, slow = _chunk_stats(emu)
    assert fast > 0 and slow > 0, (fast, slow)


@pytest.mark.parametrize("reverse,wide,n_warps", [(0, 0, 3), (1, 0, 2),
                                                  (0, 1, 2), (1, 2, 4)])
def test_emulated_seams_uncovered_tiles(emu, reverse, wide, n_warps):
    """Sparse placement: most 32x32 skewed tiles of the canvas are covered by
    no image and take the wavefront's uncovered step variant (first band,
    which keeps the border selects, and the bands below it). One image has a
    mostly transparent bounding box, so run-start bits fall into such tiles
    too."""
    rng = np.random.default_rng(77)
    W, H = 520, 150
    frame = np.zeros((H - 10, W - 20), np.uint8)   # opaque corners only
    frame[:12, :15] = 1
    frame[-9:, -14:] = 1
    frame[60:75, 240:262] = 1
    blob = (rng.random((40, 60)) < 0.85).astype(np.uint8)
    blob[0, :] = blob[-1, :] = blob[:, 0] = blob[:, -1] = 1
    small = np.ones((25, 30), np.uint8)
    ops = [frame, blob, small]
    tpos = [(0, 0), (W - 60, H - 40), (200, 50)]
    ref_l, ref_b, ref_f = mb_oracle.seams(ops, tpos, W, H,
                                          reverse=bool(reverse), debug=True)
    emu.emu_uncovered_chunks.restype = ctypes.c_long
    emu.emu_uncovered_chunks()
    l, b, f = _emu_seams(emu, ops, tpos, W, H, reverse, wide, n_warps)
    assert emu.emu_uncovered_chunks() >= 20
    assert np.array_equal(b, ref_b), "backward seam DT differs"
    assert np.array_equal(f, ref_f), "forward seam DT differs"
    assert np.array_equal(l, ref_l)


def _chunk_stats(emu):
    fast, slow = ctypes.c_long(), ctypes.c_long()
    emu.emu_chunk_stats(ctypes.byref(fast), ctypes.byref(slow))
    return fast.value, slow.value


def _check_seams(emu, seed, n, mode, reverse, wide, n_warps, w, h):
    ops, tpos, W, H = _seam_inputs(seed, n, mode, w=w, h=h)
    ref_l, ref_b, ref_f = mb_oracle.seams(ops, tpos, W, H, reverse=bool(reverse),
                                          debug=True)
    l, b, f = _emu_seams(emu, ops, tpos, W, H, reverse, wide, n_warps)
    assert np.array_equal(b, ref_b), "backward seam DT differs"
    assert np.array_equal(f, ref_f), "forward seam DT differs"
    assert np.array_equal(l, ref_l)


@pytest.mark.parametrize("seed,mode,bpp,n_warps,w,h", [
    (11, "ellipse", 8, 2, 60, 45), (12, "blobs", 8, 3, 60, 45),
    (13, "ellipse", 16, 2, 60, 45), (14, "ellipse", 8, 3, 280, 110),
])
def test_emulated_inpaint_matches_oracle(emu, seed, mode, bpp, n_warps, w, h):
    ims, _ = synth.blend_case(seed, n=3, w=w, h=h, mode=mode, bpp=bpp)
    bbs = [mb_oracle.trim(im) for im in ims]
    want = [mb_oracle.inpaint(im, bb)[0] for im, bb in zip(ims, bbs)]
    got = [np.ascontiguousarray(im).copy() for im in ims]
    n = len(got)
    ptrs = (ctypes.c_void_p * n)(*[g.ctypes.data for g in got])
    tw = (ctypes.c_int * n)(*[g.shape[1] for g in got])
    bb = (ctypes.c_int * (4 * n))(*[v for b in bbs for v in b])
    _chunk_stats(emu)
    emu.emu_inpaint(n, ptrs, tw, bpp, bb, n_warps)
    fast, slow = _chunk_stats(emu)
    assert slow > 0 and (w < 200 or fast > 0), (fast, slow)
    for g, w_, b in zip(got, want, bbs):
        l, t, ww, hh = b
        assert np.array_equal(g[t:t + hh, l:l + ww], w_[t:t + hh, l:l + ww])

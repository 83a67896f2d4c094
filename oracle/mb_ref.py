"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/_ref/libmb_ref.so.

libmb_ref.so is the UNMODIFIED reference blend library (external/multiblend)
compiled by oracle/build_ref.sh behind oracle/ref_driver.cpp. It is the
ground truth for the blend half of the hot path and the
timed CPU baseline (`cpu_baseline.kind == "reference"`).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may
import this module. The product (xpano_b200) never does.
"""
from __future__ import annotations

import ctypes
import os
import struct
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_ref", "libmb_ref.so")
_lib = None

TIMING_KEYS = ("images", "copy", "seam", "shrink_mask", "shrink", "laplace",
               "blend", "collapse", "wrap", "out")


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(_LIB_PATH)
        lib.mb_ref_blend.restype = ctypes.c_void_p
        lib.mb_ref_blend.argtypes = [
            ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
            ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
            ctypes.c_int, ctypes.c_int, ctypes.c_int]
        lib.mb_ref_info.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        lib.mb_ref_copy.argtypes = [ctypes.c_void_p] + [ctypes.c_void_p] * 5
        lib.mb_ref_free.argtypes = [ctypes.c_void_p]
        lib.mb_ref_last_error.restype = ctypes.c_char_p
        _lib = lib
    return _lib


@dataclass
class BlendResult:
    planes: np.ndarray      # (3, H, W) u8 or u16, channel order as fed
    mask: np.ndarray        # (H, W) u8, 0/255 (decoded full_mask)
    min_x: int
    min_y: int
    no_mask: bool
    timing: dict

    @property
    def width(self):
        return self.planes.shape[2]

    @property
    def height(self):
        return self.planes.shape[1]

    def interleaved(self, zero_outside_mask=True) -> np.ndarray:
        """(H, W, 3) as xpano returns it (blenders.cc:64-77, 151-154)."""
        out = np.ascontiguousarray(np.moveaxis(self.planes, 0, 2))
        if zero_outside_mask:
            out[self.mask == 0] = 0
        return out


def blend(images, positions, output_bpp=0, fixed_levels=0, add_levels=0,
          dither=True, reverse=False, n_threads=None,
          wideblend=False) -> BlendResult:
    """images: list of (H, W, 3|4) u8/u16 arrays; positions: list of (x, y)."""
    lib = _load()
    n = len(images)
    if n_threads is None:
        n_threads = max(2, len(os.sched_getaffinity(0)) - 1)
    arrs = [np.ascontiguousarray(im) for im in images]
    for a in arrs:
        assert a.ndim == 3 and a.shape[2] in (3, 4)
        assert a.dtype in (np.uint8, np.uint16)
    iarr = lambda v: (ctypes.c_int * n)(*[int(x) for x in v])
    w = iarr(a.shape[1] for a in arrs)
    h = iarr(a.shape[0] for a in arrs)
    bpp = iarr(8 * a.dtype.itemsize for a in arrs)
    spp = iarr(a.shape[2] for a in arrs)
    xs = iarr(p[0] for p in positions)
    ys = iarr(p[1] for p in positions)
    ptrs = (ctypes.c_void_p * n)(*[a.ctypes.data for a in arrs])
    lib.mb_ref_set_wideblend(int(bool(wideblend)))
    hnd = lib.mb_ref_blend(n, w, h, bpp, spp, xs, ys, ptrs, output_bpp,
                           fixed_levels, add_levels, int(dither),
                           int(reverse), n_threads)
    if not hnd:
        raise RuntimeError("mb_ref: " + lib.mb_ref_last_error().decode())
    try:
        info = (ctypes.c_int * 8)()
        lib.mb_ref_info(hnd, info)
        W, H, obpp, min_x, min_y, no_mask = list(info)[:6]
        dt = np.uint16 if obpp == 16 else np.uint8
        planes = np.empty((3, H, W), dt)
        mask = np.empty((H, W), np.uint8)
        timing = (ctypes.c_double * 10)()
        lib.mb_ref_copy(hnd, planes[0].ctypes.data, planes[1].ctypes.data,
                        planes[2].ctypes.data, mask.ctypes.data, timing)
    finally:
        lib.mb_ref_free(hnd)
    return BlendResult(planes, mask, min_x, min_y, bool(no_mask),
                       dict(zip(TIMING_KEYS, list(timing))))


def read_cereal_raw(path):
    """Decode the reference's InMemoryImage fixtures
    (mb/tests/data/img_*.png.raw; mb/tests/multiblend_test.cpp:20-31):
    cereal binary = <iiHHiiQ> header + bytes."""
    with open(path, "rb") as f:
        d = f.read()
    w, h, bpp, spp, x, y, nbytes = struct.unpack("<iiHHiiQ", d[:28])
    dt = np.uint16 if bpp == 16 else np.uint8
    img = np.frombuffer(d[28:28 + nbytes], dt).reshape(h, w, spp).copy()
    return img, (x, y)

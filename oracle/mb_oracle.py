"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/libmb_oracle.so, the
plain-C restatement of the reference blend algorithm (oracle/mb_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module. The product (xpano_b200) never does.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libmb_oracle.so")
_lib = None

MAX_LEVELS = 24


class _Image(ctypes.Structure):
    _fields_ = [("width", ctypes.c_int), ("height", ctypes.c_int),
                ("bpp", ctypes.c_int), ("spp", ctypes.c_int),
                ("xpos", ctypes.c_int), ("ypos", ctypes.c_int),
                ("data", ctypes.c_void_p)]


class _Options(ctypes.Structure):
    _fields_ = [("output_bpp", ctypes.c_int), ("fixed_levels", ctypes.c_int),
                ("add_levels", ctypes.c_int), ("dither", ctypes.c_int),
                ("reverse", ctypes.c_int), ("wideblend", ctypes.c_int)]


class _Trimmed(ctypes.Structure):
    _fields_ = [("left", ctypes.c_int), ("top", ctypes.c_int),
                ("width", ctypes.c_int), ("height", ctypes.c_int),
                ("xpos", ctypes.c_int), ("ypos", ctypes.c_int),
                ("opaque", ctypes.c_void_p),
                ("chan", ctypes.c_void_p * 3)]


class _Result(ctypes.Structure):
    _fields_ = [("width", ctypes.c_int), ("height", ctypes.c_int),
                ("bpp", ctypes.c_int), ("min_xpos", ctypes.c_int),
                ("min_ypos", ctypes.c_int), ("no_mask", ctypes.c_int),
                ("levels", ctypes.c_int),
                ("planes", ctypes.c_void_p * 3),
                ("full_mask", ctypes.c_void_p),
                ("xor_mask", ctypes.c_void_p),
                ("labels", ctypes.c_void_p),
                ("seam_bwd", ctypes.c_void_p),
                ("seam_fwd", ctypes.c_void_p),
                ("channel_totals", ctypes.c_uint64 * 3),
                ("total_pixels", ctypes.c_uint64),
                ("collapsed", ctypes.c_void_p * 3),
                ("out_pitch", ctypes.c_int),
                ("n_images", ctypes.c_int),
                ("images", ctypes.POINTER(_Trimmed))]


def build(force=False):
    """Compile oracle/mb_oracle.c (gcc, no GPU needed)."""
    src = os.path.join(_HERE, "mb_oracle.c")
    if (force or not os.path.exists(_LIB_PATH)
            or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-C", _HERE, "libmb_oracle.so"],
                              stdout=subprocess.DEVNULL)


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(_LIB_PATH)
        lib.mbo_blend.restype = ctypes.c_int
        lib.mbo_blend.argtypes = [ctypes.c_int, ctypes.POINTER(_Image),
                                  ctypes.POINTER(_Options), ctypes.c_int,
                                  ctypes.POINTER(_Result)]
        lib.mbo_free_result.argtypes = [ctypes.POINTER(_Result)]
        lib.mbo_last_error.restype = ctypes.c_char_p
        lib.mbo_trim.restype = ctypes.c_int
        lib.mbo_trim.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                 ctypes.c_int, ctypes.c_void_p]
        lib.mbo_inpaint.argtypes = [ctypes.c_void_p, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_void_p,
                                    ctypes.c_void_p]
        lib.mbo_seams.argtypes = [ctypes.c_int] + [ctypes.c_void_p] * 5 + [
            ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
            ctypes.c_void_p, ctypes.c_void_p]
        lib.mbo_pyramid_geometry.restype = ctypes.c_int
        lib.mbo_pyramid_geometry.argtypes = [ctypes.c_int] * 5 + [
            ctypes.c_void_p] * 7
        lib.mbo_blend_levels.restype = ctypes.c_int
        lib.mbo_blend_levels.argtypes = [ctypes.c_int, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_int]
        _lib = lib
    return _lib


def _np_from(ptr, shape, dtype):
    n = int(np.prod(shape))
    if not ptr or n == 0:
        return None
    buf = (ctypes.c_uint8 * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).reshape(shape).copy()


@dataclass
class Trimmed:
    left: int
    top: int
    width: int
    height: int
    xpos: int
    ypos: int
    opaque: np.ndarray | None = None
    chan: list = field(default_factory=list)


@dataclass
class OracleResult:
    planes: np.ndarray
    mask: np.ndarray
    min_x: int
    min_y: int
    no_mask: bool
    levels: int
    labels: np.ndarray | None = None
    xor_mask: np.ndarray | None = None
    seam_bwd: np.ndarray | None = None
    seam_fwd: np.ndarray | None = None
    channel_totals: tuple = ()
    total_pixels: int = 0
    collapsed: np.ndarray | None = None
    images: list = field(default_factory=list)

    @property
    def width(self):
        return self.planes.shape[2]

    @property
    def height(self):
        return self.planes.shape[1]

    def interleaved(self, zero_outside_mask=True):
        out = np.ascontiguousarray(np.moveaxis(self.planes, 0, 2))
        if zero_outside_mask:
            out[self.mask == 0] = 0
        return out


def blend(images, positions, output_bpp=0, fixed_levels=0, add_levels=0,
          dither=True, reverse=False, debug=False,
          wideblend=False) -> OracleResult:
    lib = _load()
    n = len(images)
    arrs = [np.ascontiguousarray(im) for im in images]
    ims = (_Image * n)()
    for i, a in enumerate(arrs):
        assert a.ndim == 3 and a.shape[2] in (3, 4)
        assert a.dtype in (np.uint8, np.uint16)
        ims[i] = _Image(a.shape[1], a.shape[0], 8 * a.dtype.itemsize,
                        a.shape[2], int(positions[i][0]),
                        int(positions[i][1]), a.ctypes.data)
    opts = _Options(output_bpp, fixed_levels, add_levels, int(dither),
                    int(reverse), int(wideblend))
    res = _Result()
    rc = lib.mbo_blend(n, ims, ctypes.byref(opts), int(debug),
                       ctypes.byref(res))
    if rc != 0:
        raise RuntimeError("mb_oracle: " + lib.mbo_last_error().decode())
    try:
        W, H = res.width, res.height
        dt = np.uint16 if res.bpp == 16 else np.uint8
        planes = np.stack([_np_from(res.planes[c], (H, W), dt)
                           for c in range(3)])
        out = OracleResult(
            planes=planes, mask=_np_from(res.full_mask, (H, W), np.uint8),
            min_x=res.min_xpos, min_y=res.min_ypos,
            no_mask=bool(res.no_mask), levels=res.levels,
            labels=_np_from(res.labels, (H, W), np.int32),
            xor_mask=_np_from(res.xor_mask, (H, W), np.uint8),
            channel_totals=tuple(res.channel_totals),
            total_pixels=int(res.total_pixels))
        if debug:
            out.seam_bwd = _np_from(res.seam_bwd, (H, W), np.uint64)
            out.seam_fwd = _np_from(res.seam_fwd, (H, W), np.uint64)
            out.collapsed = np.stack([
                _np_from(res.collapsed[c], (H, res.out_pitch), np.float32)
                for c in range(3)])[:, :, :W]
            for i in range(n):
                t = res.images[i]
                tr = Trimmed(t.left, t.top, t.width, t.height, t.xpos, t.ypos)
                tr.opaque = _np_from(t.opaque, (t.height, t.width), np.uint8)
                sdt = arrs[i].dtype
                tr.chan = [_np_from(t.chan[c], (t.height, t.width), sdt)
                           for c in range(3)]
                out.images.append(tr)
    finally:
        lib.mbo_free_result(ctypes.byref(res))
    return out


def trim(bgra: np.ndarray):
    """image.cpp:396-512 -> (left, top, width, height) or None."""
    lib = _load()
    a = np.ascontiguousarray(bgra)
    out = (ctypes.c_int * 4)()
    rc = lib.mbo_trim(a.ctypes.data, a.shape[1], a.shape[0],
                      8 * a.dtype.itemsize, out)
    return None if rc else tuple(out)


def inpaint(bgra: np.ndarray, bbox=None):
    """image.cpp:514-812. Returns (filled copy, opaque[h,w] of the bbox)."""
    lib = _load()
    a = np.ascontiguousarray(bgra).copy()
    if bbox is None:
        bbox = trim(a)
    bb = (ctypes.c_int * 4)(*bbox)
    opaque = np.zeros((bbox[3], bbox[2]), np.uint8)
    lib.mbo_inpaint(a.ctypes.data, a.shape[1], 8 * a.dtype.itemsize, bb,
                    opaque.ctypes.data)
    return a, opaque


def seams(opaques, positions, width, height, reverse=False, debug=False):
    """multiblend.cpp:254-823 on dense opacity maps (uint8 0/1 arrays)."""
    lib = _load()
    n = len(opaques)
    ops = [np.ascontiguousarray(o, dtype=np.uint8) for o in opaques]
    iarr = lambda v: (ctypes.c_int * n)(*[int(x) for x in v])
    xs = iarr(p[0] for p in positions)
    ys = iarr(p[1] for p in positions)
    ws = iarr(o.shape[1] for o in ops)
    hs = iarr(o.shape[0] for o in ops)
    ptrs = (ctypes.c_void_p * n)(*[o.ctypes.data for o in ops])
    labels = np.empty((height, width), np.int32)
    bwd = np.empty((height, width), np.uint64) if debug else None
    fwd = np.empty((height, width), np.uint64) if debug else None
    lib.mbo_seams(n, xs, ys, ws, hs, ptrs, width, height, int(reverse),
                  labels.ctypes.data,
                  bwd.ctypes.data if debug else None,
                  fwd.ctypes.data if debug else None)
    return (labels, bwd, fwd) if debug else labels


def pyramid_geometry(width, height, levels, x, y):
    """pyramid.cpp:21-98 -> list of dicts per level."""
    lib = _load()
    arr = [(ctypes.c_int * MAX_LEVELS)() for _ in range(7)]
    n = lib.mbo_pyramid_geometry(width, height, levels, x, y, *arr)
    keys = ("w", "h", "x", "y", "xs", "ys", "pitch")
    return [dict(zip(keys, [a[l] for a in arr])) for l in range(n)]


def blend_levels(widths, heights, fixed_levels=0, add_levels=0):
    lib = _load()
    n = len(widths)
    w = (ctypes.c_int * n)(*widths)
    h = (ctypes.c_int * n)(*heights)
    return lib.mbo_blend_levels(n, w, h, fixed_levels, add_levels)

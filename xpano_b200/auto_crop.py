"""Host-side mirror of xpano::algorithm::crop::FindLargestCrop
(xpano/algorithm/auto_crop.cc:118-143): the largest inscribed rectangle of the
panorama mask (approximate, seeded growth), executed by libxpano_b200.so."""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from ._lib import c_void_p, check


def _rect(out, found):
    return tuple(int(v) for v in out) if found.value else None


def FindLargestCrop(mask, device=0):
    """mask: (H, W) uint8 host array (255 = set) -> (x0, y0, x1, y1) or None,
    as utils::RectPPi {start, end}."""
    lib = _lib.load()
    out = (ctypes.c_int * 4)()
    found = ctypes.c_int()
    if mask is None or mask.size == 0:
        return None
    mask = np.asarray(mask)
    if mask.dtype != np.uint8 or mask.ndim != 2:
        raise ValueError("FindLargestCrop: CV_8U single-channel mask expected")
    if mask.strides[1] != 1 or mask.strides[0] < 0:
        mask = np.ascontiguousarray(mask)
    h, w = mask.shape
    check(lib.xp_find_largest_crop(mask.ctypes.data, w, h, mask.strides[0],
                                   _lib.MEM_HOST, int(device), out,
                                   ctypes.byref(found)))
    return _rect(out, found)


def largest_crop_from_lines(lines, rows):
    """The host half alone: lines = (W, 2) int32 [start, end) per column."""
    lib = _lib.load()
    lines = np.ascontiguousarray(lines, np.int32)
    out = (ctypes.c_int * 4)()
    found = ctypes.c_int()
    check(lib.xp_largest_crop_from_lines(lines.ctypes.data, len(lines),
                                         int(rows), out, ctypes.byref(found)))
    return _rect(out, found)


def blender_auto_crop(blender_handle):
    """FindLargestCrop on the mask a blender (or composer's blender) holds on
    the device after blend()."""
    lib = _lib.load()
    out = (ctypes.c_int * 4)()
    found = ctypes.c_int()
    check(lib.xp_blender_auto_crop(c_void_p(blender_handle), out,
                                   ctypes.byref(found)))
    return _rect(out, found)


def blender_inpaint_mask(blender_handle, size):
    """The inputs of the auto-fill step (stitcher_pipeline.cc:384-389) from
    the mask a blended blender holds on the device: -> (inpaint_mask (h, w)
    uint8 = ~pano_mask, pixels_filled). size = (w, h) of the panorama."""
    lib = _lib.load()
    w, h = size
    inv = np.empty((h, w), np.uint8)
    n = ctypes.c_longlong()
    check(lib.xp_blender_inpaint_mask(c_void_p(blender_handle),
                                      inv.ctypes.data, 0, _lib.MEM_HOST,
                                      ctypes.byref(n)))
    return inv, int(n.value)

"""Host-side mirror of xpano::algorithm::blenders::Multiblend
(xpano/algorithm/blenders.h:25-39, blenders.cc:116-158): prepare / feed /
blend, executed by libxpano_b200.so on the GPU.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from ._lib import c_void_p, check


class Multiblend:
    """cv::detail::Blender look-alike behind xpano's multiblend wrapper.

    Options follow multiblend::Options (mb/multiblend.h:29-45); xpano itself
    only sets output_bpp = 8 (blenders.cc:145-149)."""

    def __init__(self, output_bpp=0, fixed_levels=0, add_levels=0,
                 dither=True, reverse=False, device=0, keep_debug=False,
                 wideblend=False):
        self._lib = _lib.load()
        self._h = c_void_p()
        o = _lib.BlendOpts(int(output_bpp), int(fixed_levels),
                           int(add_levels), int(bool(dither)),
                           int(bool(reverse)), int(device),
                           int(bool(keep_debug)), int(bool(wideblend)))
        check(self._lib.xp_blender_create(ctypes.byref(o),
                                          ctypes.byref(self._h)))
        self.timing = None
        self.size = None
        self.min_pos = None
        self.no_mask = None
        self.bpp = None

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.xp_blender_destroy(self._h)
            self._h = c_void_p()

    def set_stream(self, cuda_stream):
        check(self._lib.xp_blender_set_stream(self._h, c_void_p(cuda_stream)))

    def prepare(self, dst_roi):
        x, y, w, h = dst_roi
        check(self._lib.xp_blender_prepare(self._h, int(x), int(y), int(w),
                                           int(h)))

    def feed(self, img, mask, tl):
        """img: (h, w, 3) u8/u16, mask: (h, w) u8, tl: (x, y). Rows may be
        strided (a cv::Mat ROI: step != cols * elemSize), as blenders.cc:86-95
        handles non-continuous inputs."""
        img, mask = np.asarray(img), np.asarray(mask)
        if img.ndim != 3 or img.shape[2] != 3 or \
                img.dtype not in (np.uint8, np.uint16):
            raise ValueError("feed: CV_Assert(img.type() == CV_8UC3)")
        if mask.dtype != np.uint8 or mask.ndim != 2:
            raise ValueError("feed: CV_Assert(mask.type() == CV_8U)")
        if mask.shape != img.shape[:2]:
            raise ValueError("feed: mask size differs from the image")
        it = img.dtype.itemsize
        if img.strides[1:] != (3 * it, it) or img.strides[0] < 0:
            img = np.ascontiguousarray(img)
        if mask.strides[1] != 1 or mask.strides[0] < 0:
            mask = np.ascontiguousarray(mask)
        h, w = mask.shape
        check(self._lib.xp_blender_feed(
            self._h, img.ctypes.data, img.strides[0], mask.ctypes.data,
            mask.strides[0], w, h, 8 * it, int(tl[0]), int(tl[1]),
            _lib.MEM_HOST))

    def feed_device(self, img_ptr, stride, mask_ptr, mask_stride, w, h,
                    depth, tl):
        """feed() of buffers already in device memory (XP_MEM_DEVICE): raw
        device pointers, strides in bytes."""
        check(self._lib.xp_blender_feed(
            self._h, c_void_p(img_ptr), int(stride), c_void_p(mask_ptr),
            int(mask_stride), int(w), int(h), int(depth), int(tl[0]),
            int(tl[1]), _lib.MEM_DEVICE))

    def feed_interleaved(self, data, tl):
        """multiblend::io::InMemoryImage with spp 3 or 4."""
        data = np.ascontiguousarray(data)
        h, w, spp = data.shape
        check(self._lib.xp_blender_feed_interleaved(
            self._h, data.ctypes.data, 0, w, h, 8 * data.dtype.itemsize, spp,
            int(tl[0]), int(tl[1]), _lib.MEM_HOST))

    def run(self):
        """Run the blend on the device; results stay resident."""
        wh = (ctypes.c_int * 2)()
        mn = (ctypes.c_int * 2)()
        bpp, nm = ctypes.c_int(), ctypes.c_int()
        t = _lib.Timing()
        check(self._lib.xp_blender_blend(self._h, wh, mn, ctypes.byref(bpp),
                                         ctypes.byref(nm), ctypes.byref(t)))
        self.size = (wh[0], wh[1])
        self.min_pos = (mn[0], mn[1])
        self.bpp, self.no_mask = bpp.value, bool(nm.value)
        if self.timing is None:
            self.timing = t.as_dict()
        return self.size

    def blend(self):
        """-> (dst (H, W, 3) zeroed outside the mask, dst_mask (H, W) 0/255)
        as Blender::blend returns them (blenders.cc:151-154)."""
        W, H = self.run()
        dt = np.uint16 if self.bpp == 16 else np.uint8
        dst = np.empty((H, W, 3), dt)
        mask = np.empty((H, W), np.uint8)
        check(self._lib.xp_blender_fetch(self._h, dst.ctypes.data, 0,
                                         mask.ctypes.data, 0, _lib.MEM_HOST))
        return dst, mask

    def blend_device(self, dst_ptr, dst_stride, mask_ptr, mask_stride):
        """blend() into caller-owned device buffers (XP_MEM_DEVICE); either
        pointer may be 0. -> (W, H)"""
        size = self.run()
        check(self._lib.xp_blender_fetch(
            self._h, c_void_p(dst_ptr or None), int(dst_stride),
            c_void_p(mask_ptr or None), int(mask_stride), _lib.MEM_DEVICE))
        return size

    def mask_rle(self):
        """Result.full_mask as the reference returns it (utils::Flex): ->
        (runs uint32 array of (white << 31) | length words, row byte offsets).
        `decode_mask_rle` is xpano's ToMat (blenders.cc:38-62)."""
        return fetch_mask_rle(self._lib, self._h, self.run()[1])

    def planes(self):
        """Result.output_channels: (3, H, W), not zeroed outside the mask."""
        W, H = self.run()
        dt = np.uint16 if self.bpp == 16 else np.uint8
        out = np.empty((3, H, W), dt)
        check(self._lib.xp_blender_fetch_planes(self._h, out.ctypes.data,
                                                _lib.MEM_HOST))
        return out

    def debug(self, what, image=0, channel=0, level=0, dtype=None):
        """Intermediate arrays for stage-level parity tests."""
        return debug_fetch(self._lib, self._h, what, image, channel, level,
                           dtype)


_DEBUG_DTYPES = {"labels": np.int32, "seam_bwd": np.uint64,
                 "seam_fwd": np.uint64, "xor_mask": np.uint8,
                 "full_mask": np.uint8, "trim": np.int32,
                 "totals": np.uint64, "opaque": np.uint8,
                 "collapsed": np.float32, "mask": np.float32,
                 "mask_rect": np.int32, "lap": np.float32, "out": np.float32}


def debug_fetch(lib, handle, what, image=0, channel=0, level=0, dtype=None):
    dims = (ctypes.c_int * 3)()
    nbytes = ctypes.c_size_t()
    key = what.encode()
    check(lib.xp_blender_debug_fetch(handle, key, image, channel, level, None,
                                     0, dims, ctypes.byref(nbytes)))
    dt = np.dtype(dtype or _DEBUG_DTYPES.get(what, np.uint8))
    buf = np.empty(max(nbytes.value, 1), np.uint8)
    check(lib.xp_blender_debug_fetch(handle, key, image, channel, level,
                                     buf.ctypes.data, buf.size, dims,
                                     ctypes.byref(nbytes)))
    arr = buf[:nbytes.value].view(dt)
    rows, cols, pitch = dims[0], dims[1], dims[2]
    if rows * pitch == arr.size and rows > 1:
        arr = arr.reshape(rows, pitch)[:, :cols]
    return arr


def fetch_mask_rle(lib, handle, height):
    n = ctypes.c_size_t()
    check(lib.xp_blender_fetch_mask_rle(handle, None, 0, ctypes.byref(n),
                                        None, _lib.MEM_HOST))
    runs = np.empty(max(n.value, 1), np.uint32)
    rows = np.empty(height, np.int32)
    check(lib.xp_blender_fetch_mask_rle(handle, runs.ctypes.data, runs.size,
                                        ctypes.byref(n), rows.ctypes.data,
                                        _lib.MEM_HOST))
    return runs[:n.value], rows


def decode_mask_rle(runs, width, height):
    """blenders.cc:38-62 ToMat: walk the words row by row; a word with the
    top bit set is a run of 255, otherwise a run of 0; zero-length runs and
    runs past the end of a row are errors ("invalid mask format" / "mask out
    of bounds")."""
    mask = np.empty((height, width), np.uint8)
    k = 0
    for y in range(height):
        x = 0
        while x < width:
            w = int(runs[k])
            k += 1
            n = w & 0x7FFFFFFF
            if n == 0:
                raise RuntimeError("Multiblend: invalid mask format")
            if x + n > width:
                raise RuntimeError("Multiblend: mask out of bounds")
            mask[y, x:x + n] = 255 if w & 0x80000000 else 0
            x += n
    return mask


class MultiBandOpenCV:
    """xpano::algorithm::blenders::MultiBandOpenCV look-alike
    (xpano/algorithm/blenders.h:41-50, blenders.cc:160-175): OpenCV's
    MultiBandBlender(false, num_bands = 5, CV_32F) fed with 16-bit signed
    copies of the 8-bit images, result converted back to 8 bits --
    BlendingMethod::kOpenCV (algorithm.cc:135-136)."""

    def __init__(self, num_bands=5, device=0):
        self._lib = _lib.load()
        self._h = c_void_p()
        check(self._lib.xp_multiband_create(int(num_bands), int(device),
                                            ctypes.byref(self._h)))
        self.num_bands = None

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.xp_multiband_destroy(self._h)
            self._h = c_void_p()

    def prepare(self, dst_roi):
        x, y, w, h = dst_roi
        check(self._lib.xp_multiband_prepare(self._h, int(x), int(y), int(w),
                                             int(h)))

    def feed(self, img, mask, tl):
        """img: (h, w, 3) uint8, mask: (h, w) uint8 (mask / 255 is the blend
        weight), tl: (x, y); rows may be strided."""
        img, mask = np.asarray(img), np.asarray(mask)
        if img.ndim != 3 or img.shape[2] != 3 or img.dtype != np.uint8:
            raise ValueError("feed: CV_8UC3 image expected")
        if mask.dtype != np.uint8 or mask.shape != img.shape[:2]:
            raise ValueError("feed: CV_8U mask of the image size expected")
        if img.strides[2] != 1 or img.strides[1] != 3 or img.strides[0] < 0:
            img = np.ascontiguousarray(img)
        if mask.strides[1] != 1 or mask.strides[0] < 0:
            mask = np.ascontiguousarray(mask)
        check(self._lib.xp_multiband_feed(
            self._h, img.ctypes.data, img.strides[0], mask.ctypes.data,
            mask.strides[0], img.shape[1], img.shape[0], int(tl[0]),
            int(tl[1]), _lib.MEM_HOST))

    def blend(self):
        """-> (dst (h, w, 3) uint8, dst_mask (h, w) uint8)."""
        wh = (ctypes.c_int * 2)()
        nb = ctypes.c_int()
        check(self._lib.xp_multiband_blend(self._h, wh, ctypes.byref(nb)))
        self.num_bands = nb.value
        dst = np.empty((wh[1], wh[0], 3), np.uint8)
        mask = np.empty((wh[1], wh[0]), np.uint8)
        check(self._lib.xp_multiband_fetch(self._h, dst.ctypes.data, 0,
                                           mask.ctypes.data, 0,
                                           _lib.MEM_HOST))
        return dst, mask

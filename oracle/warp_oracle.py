"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/libwarp_oracle.so
(oracle/warp_oracle.c), the CPU restatement of OpenCV's RotationWarper +
remap + resize(INTER_LINEAR_EXACT) + dilate as xpano calls them
(xpano/algorithm/stitcher.cc:144-169, 377-401).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module. The product (xpano_b200) never does.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libwarp_oracle.so")
_lib = None

# xpano ProjectionType order (xpano/algorithm/options.h:16-26)
PLANE, CYLINDRICAL, SPHERICAL, FISHEYE, STEREOGRAPHIC, \
    COMPRESSED_RECTILINEAR, PANINI, MERCATOR, TRANSVERSE_MERCATOR = range(9)

# name accepted by cv2.PyRotationWarper -> (projection, portrait, a, b)
CV2_NAMES = {
    "plane": (PLANE, 0, 0, 0),
    "cylindrical": (CYLINDRICAL, 0, 0, 0),
    "spherical": (SPHERICAL, 0, 0, 0),
    "fisheye": (FISHEYE, 0, 0, 0),
    "stereographic": (STEREOGRAPHIC, 0, 0, 0),
    "compressedPlaneA2B1": (COMPRESSED_RECTILINEAR, 0, 2, 1),
    "compressedPlaneA1.5B1": (COMPRESSED_RECTILINEAR, 0, 1.5, 1),
    "compressedPlanePortraitA2B1": (COMPRESSED_RECTILINEAR, 1, 2, 1),
    "compressedPlanePortraitA1.5B1": (COMPRESSED_RECTILINEAR, 1, 1.5, 1),
    "paniniA2B1": (PANINI, 0, 2, 1),
    "paniniA1.5B1": (PANINI, 0, 1.5, 1),
    "paniniPortraitA2B1": (PANINI, 1, 2, 1),
    "paniniPortraitA1.5B1": (PANINI, 1, 1.5, 1),
    "mercator": (MERCATOR, 0, 0, 0),
    "transverseMercator": (TRANSVERSE_MERCATOR, 0, 0, 0),
}


class _Projector(ctypes.Structure):
    _fields_ = [("type", ctypes.c_int), ("portrait", ctypes.c_int),
                ("scale", ctypes.c_float), ("a", ctypes.c_float),
                ("b", ctypes.c_float),
                ("k", ctypes.c_float * 9), ("rinv", ctypes.c_float * 9),
                ("r_kinv", ctypes.c_float * 9), ("k_rinv", ctypes.c_float * 9),
                ("t", ctypes.c_float * 3)]


def build(force=False):
    src = os.path.join(_HERE, "warp_oracle.c")
    if (force or not os.path.exists(_LIB_PATH)
            or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-C", _HERE, "libwarp_oracle.so"],
                              stdout=subprocess.DEVNULL)


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(_LIB_PATH)
        P = ctypes.POINTER(_Projector)
        vp = ctypes.c_void_p
        lib.wo_init.argtypes = [P, ctypes.c_int, ctypes.c_int, ctypes.c_float,
                                ctypes.c_float, ctypes.c_float, vp, vp]
        lib.wo_map_forward.argtypes = [P, ctypes.c_float, ctypes.c_float,
                                       vp, vp]
        lib.wo_map_backward.argtypes = [P, ctypes.c_float, ctypes.c_float,
                                        vp, vp]
        lib.wo_detect_roi.argtypes = [P, ctypes.c_int, ctypes.c_int, vp]
        lib.wo_build_maps.argtypes = [P, vp, vp, vp]
        lib.wo_remap_linear_reflect.argtypes = [vp] + [ctypes.c_int] * 4 + [
            vp, vp, ctypes.c_int, ctypes.c_int, vp]
        lib.wo_remap_nearest_mask.argtypes = [ctypes.c_int, ctypes.c_int, vp,
                                              vp, ctypes.c_int, ctypes.c_int,
                                              vp]
        lib.wo_resize_linear_exact_u8.argtypes = [vp, ctypes.c_int,
                                                  ctypes.c_int, vp,
                                                  ctypes.c_int, ctypes.c_int]
        lib.wo_dilate3x3_u8.argtypes = [vp, ctypes.c_int, ctypes.c_int, vp]
        lib.wo_resize_linear_f32.argtypes = [vp, ctypes.c_int, ctypes.c_int,
                                             vp, ctypes.c_int, ctypes.c_int]
        lib.wo_apply_gain_u8.argtypes = [vp, ctypes.c_int, ctypes.c_int, vp]
        _lib = lib
    return _lib


class Warper:
    """cv::detail::RotationWarper look-alike over the C restatement."""

    def __init__(self, projection, scale, portrait=False, a=2.0, b=1.0):
        self.projection = int(projection)
        self.portrait = int(bool(portrait))
        self.scale = float(scale)
        self.a, self.b = float(a), float(b)

    @classmethod
    def from_cv2_name(cls, name, scale):
        proj, portrait, a, b = CV2_NAMES[name]
        return cls(proj, scale, portrait, a, b)

    def _proj(self, K, R):
        lib = _load()
        p = _Projector()
        K = np.ascontiguousarray(K, np.float32).reshape(9)
        R = np.ascontiguousarray(R, np.float32).reshape(9)
        lib.wo_init(ctypes.byref(p), self.projection, self.portrait,
                    self.scale, self.a, self.b, K.ctypes.data, R.ctypes.data)
        return p

    def warp_point(self, pt, K, R):
        p = self._proj(K, R)
        u, v = ctypes.c_float(), ctypes.c_float()
        _load().wo_map_forward(ctypes.byref(p), pt[0], pt[1],
                               ctypes.byref(u), ctypes.byref(v))
        return u.value, v.value

    def warp_point_backward(self, pt, K, R):
        p = self._proj(K, R)
        x, y = ctypes.c_float(), ctypes.c_float()
        _load().wo_map_backward(ctypes.byref(p), pt[0], pt[1],
                                ctypes.byref(x), ctypes.byref(y))
        return x.value, y.value

    def _roi(self, p, size):
        out = (ctypes.c_int * 4)()
        _load().wo_detect_roi(ctypes.byref(p), size[0], size[1], out)
        return list(out)

    def warp_roi(self, size, K, R):
        """size = (w, h) -> (x, y, w, h) like RotationWarper::warpRoi."""
        tlx, tly, brx, bry = self._roi(self._proj(K, R), size)
        return tlx, tly, brx - tlx + 1, bry - tly + 1

    def build_maps(self, size, K, R):
        p = self._proj(K, R)
        roi = self._roi(p, size)
        W, H = roi[2] - roi[0] + 1, roi[3] - roi[1] + 1
        xmap = np.empty((H, W), np.float32)
        ymap = np.empty((H, W), np.float32)
        r = (ctypes.c_int * 4)(*roi)
        _load().wo_build_maps(ctypes.byref(p), r, xmap.ctypes.data,
                              ymap.ctypes.data)
        return (roi[0], roi[1], W, H), xmap, ymap

    def warp(self, src, K, R):
        """INTER_LINEAR + BORDER_REFLECT image warp (stitcher.cc:377-378).
        Returns (corner, warped)."""
        src = np.ascontiguousarray(src)
        h, w = src.shape[:2]
        roi, xmap, ymap = self.build_maps((w, h), K, R)
        return (roi[0], roi[1]), remap_linear_reflect(src, xmap, ymap)

    def warp_mask(self, size, K, R):
        """INTER_NEAREST + BORDER_CONSTANT warp of an all-255 mask
        (stitcher.cc:382-385)."""
        roi, xmap, ymap = self.build_maps(size, K, R)
        return (roi[0], roi[1]), remap_nearest_mask(size, xmap, ymap)


def remap_linear_reflect(src, xmap, ymap):
    src = np.ascontiguousarray(src)
    assert src.dtype in (np.uint8, np.uint16)
    h, w = src.shape[:2]
    cn = 1 if src.ndim == 2 else src.shape[2]
    dh, dw = xmap.shape
    dst = np.empty((dh, dw) + (() if src.ndim == 2 else (cn,)), src.dtype)
    xm = np.ascontiguousarray(xmap, np.float32)
    ym = np.ascontiguousarray(ymap, np.float32)
    _load().wo_remap_linear_reflect(src.ctypes.data, w, h, cn,
                                    8 * src.dtype.itemsize, xm.ctypes.data,
                                    ym.ctypes.data, dw, dh, dst.ctypes.data)
    return dst


def remap_nearest_mask(size, xmap, ymap):
    dh, dw = xmap.shape
    dst = np.empty((dh, dw), np.uint8)
    xm = np.ascontiguousarray(xmap, np.float32)
    ym = np.ascontiguousarray(ymap, np.float32)
    _load().wo_remap_nearest_mask(size[0], size[1], xm.ctypes.data,
                                  ym.ctypes.data, dw, dh, dst.ctypes.data)
    return dst


def resize_linear_exact_u8(src, dsize):
    """cv2.resize(src, dsize, interpolation=INTER_LINEAR_EXACT), 1 channel."""
    src = np.ascontiguousarray(src, np.uint8)
    dw, dh = dsize
    dst = np.empty((dh, dw), np.uint8)
    _load().wo_resize_linear_exact_u8(src.ctypes.data, src.shape[1],
                                      src.shape[0], dst.ctypes.data, dw, dh)
    return dst


def dilate3x3_u8(src):
    """cv2.dilate(src, None): 3x3 rectangle, border ignored."""
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty_like(src)
    _load().wo_dilate3x3_u8(src.ctypes.data, src.shape[1], src.shape[0],
                            dst.ctypes.data)
    return dst


def resize_linear_f32(src, dsize):
    """cv2.resize(src, dsize, interpolation=INTER_LINEAR) for float32."""
    src = np.ascontiguousarray(src, np.float32)
    dw, dh = dsize
    dst = np.empty((dh, dw), np.float32)
    _load().wo_resize_linear_f32(src.ctypes.data, src.shape[1], src.shape[0],
                                 dst.ctypes.data, dw, dh)
    return dst


def apply_gain_u8(img, gain):
    """BlocksCompensator::apply after the gain map was resized: per-pixel
    multiply(img, gain) with saturate-round. Returns a new array."""
    img = np.ascontiguousarray(img, np.uint8).copy()
    gain = np.ascontiguousarray(gain, np.float32)
    assert img.shape[:2] == gain.shape
    _load().wo_apply_gain_u8(img.ctypes.data, img.shape[1], img.shape[0],
                             gain.ctypes.data)
    return img

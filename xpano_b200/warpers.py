"""Host-side mirror of the warper surface the stitcher uses.

Same names and argument meaning as cv::detail::RotationWarper as called from
xpano/algorithm/stitcher.cc:160,377-385 and as created by
xpano/algorithm/algorithm.cc:43-105 (PickWarper / PickWarperPortrait); the
work is done by libxpano_b200.so on the GPU.
"""
from __future__ import annotations

import ctypes
import enum

import numpy as np

from . import _lib
from ._lib import c_void_p, check

INTER_NEAREST, INTER_LINEAR = 0, 1
BORDER_CONSTANT, BORDER_REFLECT = 0, 2


class ProjectionType(enum.IntEnum):
    """xpano::algorithm::ProjectionType (xpano/algorithm/options.h:16-26)."""
    kPerspective = 0
    kCylindrical = 1
    kSpherical = 2
    kFisheye = 3
    kStereographic = 4
    kCompressedRectilinear = 5
    kPanini = 6
    kMercator = 7
    kTransverseMercator = 8


def _f9(m):
    a = np.ascontiguousarray(m, np.float32).reshape(9)
    return a, a.ctypes.data


class RotationWarper:
    """cv::detail::RotationWarper look-alike."""

    def __init__(self, projection, scale, portrait=False, a=2.0, b=1.0,
                 device=0):
        self._h = c_void_p()
        self._lib = _lib.load()
        check(self._lib.xp_warper_create(int(projection), int(bool(portrait)),
                                         float(scale), float(a), float(b),
                                         int(device), ctypes.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.xp_warper_destroy(self._h)
            self._h = c_void_p()

    def getScale(self):
        return self._lib.xp_warper_get_scale(self._h)

    def setScale(self, scale):
        check(self._lib.xp_warper_set_scale(self._h, float(scale)))

    def set_stream(self, cuda_stream):
        check(self._lib.xp_warper_set_stream(self._h, c_void_p(cuda_stream)))

    def warpRoi(self, src_size, K, R):
        """src_size = (w, h) -> (x, y, w, h)."""
        k, kp = _f9(K)
        r, rp = _f9(R)
        out = (ctypes.c_int * 4)()
        check(self._lib.xp_warper_warp_roi(self._h, int(src_size[0]),
                                           int(src_size[1]), kp, rp, out))
        return tuple(out)

    def warpPoint(self, pt, K, R):
        k, kp = _f9(K)
        r, rp = _f9(R)
        xy = (ctypes.c_float * 2)(float(pt[0]), float(pt[1]))
        uv = (ctypes.c_float * 2)()
        check(self._lib.xp_warper_warp_point(self._h, xy, kp, rp, uv))
        return uv[0], uv[1]

    def warpPointBackward(self, pt, K, R):
        k, kp = _f9(K)
        r, rp = _f9(R)
        uv = (ctypes.c_float * 2)(float(pt[0]), float(pt[1]))
        xy = (ctypes.c_float * 2)()
        check(self._lib.xp_warper_warp_point_backward(self._h, uv, kp, rp, xy))
        return xy[0], xy[1]

    def buildMaps(self, src_size, K, R):
        """-> ((x, y, w, h), xmap, ymap) like RotationWarper::buildMaps."""
        k, kp = _f9(K)
        r, rp = _f9(R)
        roi = (ctypes.c_int * 4)()
        check(self._lib.xp_warper_build_maps(
            self._h, int(src_size[0]), int(src_size[1]), kp, rp, None, None, 0,
            _lib.MEM_HOST, roi))
        W, H = roi[2], roi[3]
        xmap = np.empty((H, W), np.float32)
        ymap = np.empty((H, W), np.float32)
        check(self._lib.xp_warper_build_maps(
            self._h, int(src_size[0]), int(src_size[1]), kp, rp,
            xmap.ctypes.data, ymap.ctypes.data, 0, _lib.MEM_HOST, roi))
        return tuple(roi), xmap, ymap

    def warp(self, src, K, R, interp_mode=INTER_LINEAR,
             border_mode=BORDER_REFLECT):
        """-> (corner, dst) like RotationWarper::warp (host arrays)."""
        src = np.ascontiguousarray(src)
        assert src.dtype in (np.uint8, np.uint16)
        h, w = src.shape[:2]
        cn = 1 if src.ndim == 2 else src.shape[2]
        x, y, W, H = self.warpRoi((w, h), K, R)
        dst = np.empty((H, W) + (() if src.ndim == 2 else (cn,)), src.dtype)
        k, kp = _f9(K)
        r, rp = _f9(R)
        corner = (ctypes.c_int * 2)()
        check(self._lib.xp_warper_warp(
            self._h, src.ctypes.data, w, h, 0, 8 * src.dtype.itemsize, cn, kp,
            rp, int(interp_mode), int(border_mode), dst.ctypes.data, 0,
            _lib.MEM_HOST, corner))
        return (corner[0], corner[1]), dst

    def warp_bgra(self, src, K, R, gain=None, seam_mask=None):
        """Fused stitcher.cc:377-405 + blenders.cc:99-112 for one image:
        -> ((x, y, w, h), BGRA array)."""
        src = np.ascontiguousarray(src)
        h, w = src.shape[:2]
        k, kp = _f9(K)
        r, rp = _f9(R)
        aux = _lib.ComposeAux()
        keep = []
        if gain is not None:
            g = np.ascontiguousarray(gain, np.float32)
            keep.append(g)
            aux.gain, aux.gain_w, aux.gain_h = (g.ctypes.data, g.shape[1],
                                                g.shape[0])
        if seam_mask is not None:
            m = np.ascontiguousarray(seam_mask, np.uint8)
            keep.append(m)
            aux.seam_mask, aux.seam_w, aux.seam_h = (m.ctypes.data, m.shape[1],
                                                     m.shape[0])
        roi = (ctypes.c_int * 4)()
        depth = 8 * src.dtype.itemsize
        check(self._lib.xp_warper_warp_bgra(
            self._h, src.ctypes.data, w, h, 0, depth, kp, rp,
            ctypes.byref(aux), _lib.MEM_HOST, None, 0, _lib.MEM_HOST, roi))
        dst = np.empty((roi[3], roi[2], 4), src.dtype)
        check(self._lib.xp_warper_warp_bgra(
            self._h, src.ctypes.data, w, h, 0, depth, kp, rp,
            ctypes.byref(aux), _lib.MEM_HOST, dst.ctypes.data, 0,
            _lib.MEM_HOST, roi))
        return tuple(roi), dst


class WarperCreator:
    """cv::WarperCreator look-alike: create(scale) -> RotationWarper."""

    def __init__(self, projection, portrait=False, a=2.0, b=1.0, device=0):
        self.projection, self.portrait = ProjectionType(projection), portrait
        self.a, self.b, self.device = a, b, device

    def create(self, scale):
        return RotationWarper(self.projection, scale, self.portrait, self.a,
                              self.b, self.device)


def PickWarper(projection, a=2.0, b=1.0, device=0):
    """algorithm.cc:43-77"""
    return WarperCreator(projection, False, a, b, device)


def PickWarperPortrait(projection, a=2.0, b=1.0, device=0):
    """algorithm.cc:79-105: None where no portrait variant exists."""
    if ProjectionType(projection) in (
            ProjectionType.kPerspective, ProjectionType.kCylindrical,
            ProjectionType.kSpherical, ProjectionType.kCompressedRectilinear,
            ProjectionType.kPanini):
        return WarperCreator(projection, True, a, b, device)
    return None


def Rotate(cameras_R, rotation_matrix):
    """algorithm::Rotate (algorithm.cc:362-369): R <- M * R for every camera
    (pitch / yaw / roll adjustment)."""
    M = np.asarray(rotation_matrix, np.float32)
    return [(M @ np.asarray(R, np.float32)).astype(np.float32)
            for R in cameras_R]

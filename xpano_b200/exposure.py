"""Host-side mirror of the exposure step of Stitcher::EstimateSeams
(xpano/algorithm/stitcher.cc:241-296): cv::detail::BlocksGainCompensator as
xpano constructs it (stitcher.cc:200-201) and the seam-scale warps that feed
it, executed by libxpano_b200.so. SURVEY.md 8(f1).

The graph-cut seam finder that follows (stitcher.cc:290,
cv::detail::GraphCutSeamFinder) is sequential max-flow and is not part of this
package: EstimateSeams takes it as a callable, exactly where the reference
calls seam_finder_->find.
"""
from __future__ import annotations

import ctypes
import math

import numpy as np

from . import _lib
from ._lib import c_void_p, check
from .warpers import (BORDER_CONSTANT, BORDER_REFLECT, INTER_LINEAR,
                      INTER_NEAREST, RotationWarper)


class BlocksGainCompensator:
    """cv::detail::BlocksGainCompensator look-alike (feed / apply /
    getMatGains / setMatGains); OpenCV's defaults."""

    def __init__(self, bl_width=32, bl_height=32, nr_gain_filtering_iterations=2,
                 device=0):
        self._lib = _lib.load()
        self._h = c_void_p()
        self._n = 0
        check(self._lib.xp_gain_compensator_create(
            int(bl_width), int(bl_height), int(nr_gain_filtering_iterations),
            int(device), ctypes.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.xp_gain_compensator_destroy(self._h)
            self._h = c_void_p()

    @staticmethod
    def _corners(corners):
        c = np.ascontiguousarray(corners, np.int32).reshape(-1, 2)
        return c

    def feed(self, corners, images, masks):
        """corners: [(x, y)], images: [(h, w, 3) uint8], masks: [(h, w) uint8,
        255 = inside] -- ExposureCompensator::feed(corners, images, masks)."""
        n = len(images)
        if not (len(corners) == n and len(masks) == n):
            raise ValueError("feed: corners, images and masks differ in length")
        ims, mks = [], []
        for im, m in zip(images, masks):
            im, m = np.asarray(im), np.asarray(m)
            if im.dtype != np.uint8 or im.ndim != 3 or im.shape[2] != 3:
                raise ValueError("feed: CV_8UC3 images expected")
            if m.dtype != np.uint8 or m.shape != im.shape[:2]:
                raise ValueError("feed: CV_8U masks of the image size expected")
            if im.strides[2] != 1 or im.strides[1] != 3 or im.strides[0] < 0:
                im = np.ascontiguousarray(im)
            if m.strides[1] != 1 or m.strides[0] < 0:
                m = np.ascontiguousarray(m)
            ims.append(im)
            mks.append(m)
        c = self._corners(corners)
        w = np.array([im.shape[1] for im in ims], np.int32)
        h = np.array([im.shape[0] for im in ims], np.int32)
        ip = (ctypes.c_void_p * max(1, n))(*[im.ctypes.data for im in ims])
        mp = (ctypes.c_void_p * max(1, n))(*[m.ctypes.data for m in mks])
        st = (ctypes.c_size_t * max(1, n))(*[im.strides[0] for im in ims])
        ms = (ctypes.c_size_t * max(1, n))(*[m.strides[0] for m in mks])
        check(self._lib.xp_gain_compensator_feed(
            self._h, n, c.ctypes.data, ip, w.ctypes.data, h.ctypes.data, st,
            mp, ms, _lib.MEM_HOST))
        self._n = n

    # -- the host-only halves (block pairs, system) ------------------------
    def plan(self, corners, sizes):
        """sizes: [(w, h)] -> (n_pairs, 6) int32 {block i, block j, x, y, w,
        h} of every overlapping block pair."""
        c = self._corners(corners)
        w = np.array([s[0] for s in sizes], np.int32)
        h = np.array([s[1] for s in sizes], np.int32)
        n_pairs = ctypes.c_int()
        check(self._lib.xp_gain_compensator_plan(
            self._h, len(sizes), c.ctypes.data, w.ctypes.data, h.ctypes.data,
            ctypes.byref(n_pairs)))
        out = np.zeros((n_pairs.value, 6), np.int32)
        if n_pairs.value:
            check(self._lib.xp_gain_compensator_pairs(self._h,
                                                      out.ctypes.data))
        self._n = len(sizes)
        return out

    def solve(self, counts, sums_first, sums_second):
        counts = np.ascontiguousarray(counts, np.int64)
        s1 = np.ascontiguousarray(sums_first, np.float64)
        s2 = np.ascontiguousarray(sums_second, np.float64)
        check(self._lib.xp_gain_compensator_solve(
            self._h, len(counts), counts.ctypes.data, s1.ctypes.data,
            s2.ctypes.data))

    def stats(self):
        """-> (block pairs, unknowns, solver iterations) of the last feed."""
        out = (ctypes.c_int * 3)()
        check(self._lib.xp_gain_compensator_stats(self._h, out))
        return tuple(out)

    def getMatGains(self):
        maps = []
        for i in range(self._n):
            wh = (ctypes.c_int * 2)()
            check(self._lib.xp_gain_compensator_map_size(self._h, i, wh))
            g = np.empty((wh[1], wh[0]), np.float32)
            check(self._lib.xp_gain_compensator_get_map(self._h, i,
                                                        g.ctypes.data))
            maps.append(g)
        return maps

    def setMatGains(self, maps):
        for i, g in enumerate(maps):
            g = np.ascontiguousarray(g, np.float32)
            check(self._lib.xp_gain_compensator_set_map(
                self._h, i, g.ctypes.data, g.shape[1], g.shape[0]))
        self._n = len(maps)

    def apply(self, index, corner, image, mask=None):
        """BlocksCompensator::apply: -> the compensated image (a copy; cv2's
        binding also returns the image). corner and mask are unused, as in
        the reference."""
        img = np.array(image, copy=True, order="C")
        if img.dtype not in (np.uint8, np.uint16) or img.ndim != 3 \
                or img.shape[2] != 3:
            raise ValueError("apply: CV_8UC3 / CV_16UC3 image expected")
        check(self._lib.xp_gain_compensator_apply(
            self._h, int(index), img.ctypes.data, img.shape[1], img.shape[0],
            img.strides[0], 8 * img.dtype.itemsize, _lib.MEM_HOST))
        return img


def ComputeSeamScale(img_size, seam_est_resol=0.1):
    """xpano/algorithm/stitcher.cc:124-126; img_size = (w, h)."""
    return min(1.0, math.sqrt(seam_est_resol * 1e6
                              / (img_size[0] * img_size[1])))


def scale_camera(K, s):
    """utils::opencv::Scale + ToFloat on one intrinsic matrix."""
    k = np.array(K, np.float64)
    k[0, 0] *= s
    k[1, 1] *= s
    k[0, 2] *= s
    k[1, 2] *= s
    return k.astype(np.float32)


def EstimateSeams(seam_est_imgs, Ks, Rs, projection, warped_image_scale,
                  seam_work_aspect, seam_finder=None, compensate=True,
                  portrait=False, a=2.0, b=1.0, device=0):
    """Stitcher::EstimateSeams (stitcher.cc:241-296).

    seam_est_imgs: the seam-scale images ((h, w, 3) uint8) the registration
    stage keeps (stitcher.cc:483-484); Ks: the work-scale intrinsics, scaled
    here by seam_work_aspect (stitcher.cc:260); Rs: rotations.
    seam_finder: callable(images_warped_f32, corners, masks_warped) -> masks,
    the place of seam_finder_->find (stitcher.cc:290); None leaves the warped
    masks as they are.
    -> dict(corners, images_warped (compensated), masks_warped (= the seams),
            gains (per-image float maps or None), compensator)."""
    warper = RotationWarper(projection,
                            float(np.float32(warped_image_scale
                                             * seam_work_aspect)),
                            portrait=portrait, a=a, b=b, device=device)
    corners, images_warped, masks_warped = [], [], []
    for img, K, R in zip(seam_est_imgs, Ks, Rs):
        img = np.ascontiguousarray(img)
        k = scale_camera(K, seam_work_aspect)
        c, iw = warper.warp(img, k, R, INTER_LINEAR, BORDER_REFLECT)
        mask = np.full(img.shape[:2], 255, np.uint8)
        _, mw = warper.warp(mask, k, R, INTER_NEAREST, BORDER_CONSTANT)
        corners.append(c)
        images_warped.append(iw)
        masks_warped.append(mw)
    comp, gains = None, None
    if compensate:
        comp = BlocksGainCompensator(device=device)
        comp.feed(corners, images_warped, masks_warped)
        images_warped = [comp.apply(i, corners[i], images_warped[i],
                                    masks_warped[i])
                         for i in range(len(images_warped))]
        gains = comp.getMatGains()
    if seam_finder is not None:
        masks_warped = seam_finder(
            [iw.astype(np.float32) for iw in images_warped], corners,
            masks_warped)
    return {"corners": corners, "images_warped": images_warped,
            "masks_warped": masks_warped, "gains": gains,
            "compensator": comp}

"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the compose loop
(Stitcher::ComposePanorama, xpano/algorithm/stitcher.cc:299-432, with
ComputeRoi :144-169 and the size cap :318-335) on top of the oracle's warp
and blend halves. Used as the checker for the device compose driver and as
the timed CPU port in bench.py's cpu_baseline leg."""
from __future__ import annotations

import math

import numpy as np

from oracle import mb_oracle
from oracle import warp_oracle as wo


from xpano_b200.synthetic import make_cameras, rot  # noqa: E402,F401


def compute_roi(warper, sizes, Ks, Rs):
    rois = [warper.warp_roi(s, K, R) for s, K, R in zip(sizes, Ks, Rs)]
    x0 = min(r[0] for r in rois)
    y0 = min(r[1] for r in rois)
    x1 = max(r[0] + r[2] for r in rois)
    y1 = max(r[1] + r[3] for r in rois)
    return rois, (x0, y0, x1 - x0, y1 - y0)


def capped_scale(warped_image_scale, compose_work_aspect, roi_w, roi_h,
                 max_pano_mpx):
    """stitcher.cc:312-329 + utils/opencv.cc:35-38, type for type: MPx is a
    float product over 1e6f, max_pano_mpx_ a float (stitcher.h:141-143),
    std::sqrt(float) the float overload; warped_image_scale_ is a double that
    is multiplied by the float ratio, and only the product with the (double)
    aspect is rounded to float. -> (float32 warp scale, capped)."""
    f32 = np.float32
    wis = float(warped_image_scale)
    pano_mpx = f32(f32(f32(roi_w) * f32(roi_h)) / f32(1e6))
    max_mpx = f32(max_pano_mpx)
    capped = bool(max_pano_mpx > 0 and pano_mpx > max_mpx)
    if capped:
        downscale_ratio = f32(np.sqrt(f32(max_mpx / pano_mpx)))
        wis = wis * float(downscale_ratio)
    return f32(wis * float(compose_work_aspect)), capped


def compose(images, Ks, Rs, projection=wo.SPHERICAL, warp_scale=1.0,
            portrait=False, a=2.0, b=1.0, max_pano_mpx=100.0, gains=None,
            seam_masks=None, keep_fed=False, warped_image_scale=None,
            compose_work_aspect=1.0, **blend_kw):
    """-> dict(pano (H, W, 3) zeroed outside mask, mask, planes, min_x, min_y,
    warp_scale, capped, fed=[(bgra, corner)])."""
    sizes = [(im.shape[1], im.shape[0]) for im in images]
    if warped_image_scale is None:
        warped_image_scale = float(np.float32(warp_scale))
        compose_work_aspect = 1.0
    scale = np.float32(warped_image_scale * compose_work_aspect)
    warper = wo.Warper(projection, float(scale), portrait, a, b)
    rois, rect = compute_roi(warper, sizes, Ks, Rs)
    scale, capped = capped_scale(warped_image_scale, compose_work_aspect,
                                 rect[2], rect[3], max_pano_mpx)
    if capped:
        warper = wo.Warper(projection, float(scale), portrait, a, b)
        rois, rect = compute_roi(warper, sizes, Ks, Rs)
    fed, posn = [], []
    for i, im in enumerate(images):
        if seam_masks is not None and seam_masks[i] is not None \
                and not seam_masks[i].any():
            continue
        corner, warped = warper.warp(im, Ks[i], Rs[i])
        _, mask = warper.warp_mask(sizes[i], Ks[i], Rs[i])
        H, W = mask.shape
        if gains is not None and gains[i] is not None:
            assert warped.dtype == np.uint8
            gm = wo.resize_linear_f32(gains[i], (W, H))
            warped = wo.apply_gain_u8(warped, gm)
        if seam_masks is not None and seam_masks[i] is not None:
            sm = wo.resize_linear_exact_u8(wo.dilate3x3_u8(seam_masks[i]),
                                           (W, H))
            mask = np.bitwise_and(sm, mask)
        maxv = 255 if warped.dtype == np.uint8 else 65535
        alpha = np.where(mask != 0, maxv, 0).astype(warped.dtype)
        fed.append(np.dstack([warped, alpha]))
        posn.append(corner)
    o = mb_oracle.blend(fed, posn, **blend_kw)
    out = dict(pano=o.interleaved(True), mask=o.mask, planes=o.planes,
               min_x=o.min_x, min_y=o.min_y, warp_scale=float(scale),
               capped=capped, rect=rect, levels=o.levels)
    if keep_fed:
        out["fed"] = list(zip(fed, posn))
    return out

"""Regenerates tests/golden/real_photos.npz -- the first real-photograph
fixture (SURVEY.md 8f1 / BASELINE.json configs[0]). Run in the build container
only (it reads /root/reference/tests/data):

    python tests/golden/make_real_fixture.py

Inputs: tests/data/image06-08.jpg, the group the reference's own pipeline test
stitches to a 976 x 1335 (+-2 %) preview panorama
(tests/stitcher_pipeline_test.cc:655-657, 718-719). The JPEG files are stored
as byte strings; the test decodes them and forms the 1024-px previews the
pipeline loads (pipeline/options.h preview_longer_side).

Cameras: a cv2.detail registration that follows Stitcher::MatchImages /
EstimateCameraParams (stitcher.cc:447-560): SIFT features at the work scale
(0.6 MPx), BestOf2NearestMatcher, HomographyBasedEstimator, BundleAdjusterRay
(conf 1), warped_image_scale = median focal, horizontal wave correction.
Registration is outside the hot path (out of scope); its outputs
are inputs of the path and are stored.

Reference outputs stored for the path's first step: the gain maps and the
graph-cut seam masks of Stitcher::EstimateSeams (stitcher.cc:241-296) computed
with cv2 4.13.0, and the SHA-1 of the decoded previews they were computed
from (the test compares against them only when its own decode agrees)."""
import hashlib
import math
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
D = "/root/reference/tests/data/"
NAMES = ["image06.jpg", "image07.jpg", "image08.jpg"]


def preview(im, longer=1024):
    h, w = im.shape[:2]
    s = longer / max(h, w)
    return cv2.resize(im, (int(round(w * s)), int(round(h * s))),
                      interpolation=cv2.INTER_AREA)


def scaled_K(K, s):
    k = np.array(K, np.float64)
    k[0, 0] *= s
    k[1, 1] *= s
    k[0, 2] *= s
    k[1, 2] *= s
    return k.astype(np.float32)


def estimate_seams_cv2(imgs, Ks, Rs, wis, work_scale, seam_scale):
    """Stitcher::EstimateSeams with cv2 (the reference's CPU path)."""
    swa = seam_scale / work_scale
    warper = cv2.PyRotationWarper("spherical", float(np.float32(wis * swa)))
    corners, warped, masks = [], [], []
    for im, K, R in zip(imgs, Ks, Rs):
        small = cv2.resize(im, None, fx=seam_scale, fy=seam_scale,
                           interpolation=cv2.INTER_LINEAR_EXACT)
        k = scaled_K(K, swa)
        c, iw = warper.warp(small, k, R, cv2.INTER_LINEAR, cv2.BORDER_REFLECT)
        _, mw = warper.warp(np.full(small.shape[:2], 255, np.uint8), k, R,
                            cv2.INTER_NEAREST, cv2.BORDER_CONSTANT)
        corners.append(c)
        warped.append(iw)
        masks.append(mw)
    comp = cv2.detail_BlocksGainCompensator()
    comp.feed(corners, warped, masks)
    warped = [comp.apply(i, corners[i], warped[i], masks[i])
              for i in range(len(imgs))]
    gains = [np.ascontiguousarray(g) for g in comp.getMatGains()]
    finder = cv2.detail_GraphCutSeamFinder("COST_COLOR")
    out = finder.find([w.astype(np.float32) for w in warped], corners, masks)
    seams = [m.get() if hasattr(m, "get") else np.asarray(m) for m in out]
    return corners, gains, seams


def main():
    raw = [open(D + n, "rb").read() for n in NAMES]
    imgs = [preview(cv2.imdecode(np.frombuffer(b, np.uint8),
                                 cv2.IMREAD_COLOR)) for b in raw]
    w, h = imgs[0].shape[1], imgs[0].shape[0]
    work_scale = min(1.0, math.sqrt(0.6e6 / (w * h)))   # stitcher.cc:120-122
    seam_scale = min(1.0, math.sqrt(0.1e6 / (w * h)))   # stitcher.cc:124-126
    finder = cv2.SIFT_create()
    feats = []
    for i, im in enumerate(imgs):
        small = cv2.resize(im, None, fx=work_scale, fy=work_scale,
                           interpolation=cv2.INTER_LINEAR_EXACT)
        f = cv2.detail.computeImageFeatures2(finder, small)
        f.img_idx = i
        feats.append(f)
    matcher = cv2.detail_BestOf2NearestMatcher(False, 0.3)
    pm = matcher.apply2(feats)
    matcher.collectGarbage()
    idx = np.array(cv2.detail.leaveBiggestComponent(feats, pm, 1.0)).ravel()
    assert list(idx) == [0, 1, 2]
    ok, cams = cv2.detail_HomographyBasedEstimator().apply(feats, pm, None)
    assert ok
    for c in cams:
        c.R = c.R.astype(np.float32)
    ba = cv2.detail_BundleAdjusterRay()
    ba.setConfThresh(1.0)
    ok, cams = ba.apply(feats, pm, cams)
    assert ok
    focals = sorted(c.focal for c in cams)
    wis = focals[1]                                      # median of three
    rm = cv2.detail.waveCorrect([np.copy(c.R) for c in cams],
                                cv2.detail.WAVE_CORRECT_HORIZ)
    Ks = [c.K() for c in cams]
    Rs = [np.asarray(r, np.float32) for r in rm]
    corners, gains, seams = estimate_seams_cv2(imgs, Ks, Rs, wis, work_scale,
                                               seam_scale)
    g = {"names": np.array(NAMES), "K": np.array(Ks, np.float64),
         "R": np.array(Rs, np.float32), "warped_image_scale": wis,
         "work_scale": work_scale, "seam_scale": seam_scale,
         "expected_rows_cols": np.array([976, 1335]),
         "preview_sha1": np.array([hashlib.sha1(i.tobytes()).hexdigest()
                                   for i in imgs]),
         "seam_corners": np.array(corners, np.int32)}
    for i in range(3):
        g[f"jpeg_{i}"] = np.frombuffer(raw[i], np.uint8)
        g[f"gain_{i}"] = gains[i]
        g[f"seam_{i}"] = np.packbits(seams[i] != 0)
        g[f"seam_shape_{i}"] = np.array(seams[i].shape)
    path = os.path.join(HERE, "real_photos.npz")
    np.savez_compressed(path, **g)
    print(path, os.path.getsize(path) // 1024, "KiB", "wis", wis)


if __name__ == "__main__":
    main()

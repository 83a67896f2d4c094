"""blenders::MultiBandOpenCV (xpano/algorithm/blenders.h:41-50, blenders.cc:
160-175; BlendingMethod::kOpenCV) -- SURVEY.md 8(f4).

CPU: the oracle (oracle/mbcv_oracle.py) against vectors recorded from cv2
4.13.0's cv::detail::MultiBandBlender (tests/golden/mbcv_cv2_goldens.npz),
live cv2 where importable, and its pyramid primitives against cv2.pyrDown /
cv2.pyrUp. GPU: xp_multiband_* against the oracle (same arithmetic: bit-exact)
and against the cv2 goldens.

Bars: CV_16S pyrDown / pyrUp, feed geometry, masks: bit-exact. Blended 8-bit
panorama against cv2: >= 99.9 % of the pixels within 1 LSB (the weight
pyramid's float sums are ordered differently inside cv2's SIMD code, see the
oracle's header); on every recorded case it is in fact identical."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from golden.make_golden import MBCV_CASES, mbcv_scene  # noqa: E402
from oracle import mbcv_oracle as mo  # noqa: E402

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def oracle_blend(roi, feeds, nb=5):
    b = mo.MultiBandBlender(nb)
    b.prepare(roi)
    for f in feeds:
        b.feed(*f)
    return b.blend() + (b.nb,)


def close(dst, want):
    d = np.abs(dst.astype(int) - want.astype(int))
    return (d <= 1).mean() >= 0.999 and d.max() <= 2, d


@pytest.mark.parametrize("case", range(len(MBCV_CASES)))
def test_oracle_matches_cv2_golden(case):
    g = np.load(os.path.join(GOLD, "mbcv_cv2_goldens.npz"))
    seed, kw = MBCV_CASES[case]
    roi, feeds = mbcv_scene(seed, **kw)
    dst, mask, _ = oracle_blend(roi, feeds)
    assert np.array_equal(mask, g[f"mask_{seed}"])
    ok, d = close(dst, g[f"dst_{seed}"])
    assert ok, (int(d.max()), float((d > 0).mean()))


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_pyramid_primitives_match_cv2_live():
    rng = np.random.default_rng(0)
    for (h, w) in [(32, 64), (33, 47), (2, 2), (6, 3), (64, 96), (1, 8)]:
        s = rng.integers(-3000, 3000, (h, w, 3)).astype(np.int16)
        assert np.array_equal(mo.pyr_down(s), cv2.pyrDown(s))
        assert np.array_equal(mo.pyr_up(s), cv2.pyrUp(s))
        f = rng.random((h, w)).astype(np.float32)
        np.testing.assert_allclose(mo.pyr_down(f), cv2.pyrDown(f), rtol=3e-7)


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
@pytest.mark.parametrize("nb", [1, 3, 5, 7])
def test_oracle_matches_cv2_live_band_counts(nb):
    roi, feeds = mbcv_scene(11 + nb, n=3, w=150, h=97, step=100)
    b = cv2.detail_MultiBandBlender(0, nb, cv2.CV_32F)
    b.prepare(roi)
    for img, mask, tl in feeds:
        b.feed(img.astype(np.int16), mask, tl)
    ref, rmask = b.blend(None, None)
    ref = np.clip(ref, 0, 255).astype(np.uint8)
    dst, mask, _ = oracle_blend(roi, feeds, nb)
    assert np.array_equal(mask, rmask)
    ok, d = close(dst, ref)
    assert ok, (int(d.max()), float((d > 0).mean()))


@pytest.mark.gpu
@pytest.mark.parametrize("case", range(len(MBCV_CASES)))
def test_device_matches_oracle_and_cv2(case):
    import xpano_b200 as xb
    g = np.load(os.path.join(GOLD, "mbcv_cv2_goldens.npz"))
    seed, kw = MBCV_CASES[case]
    roi, feeds = mbcv_scene(seed, **kw)
    want, wmask, nb = oracle_blend(roi, feeds)
    b = xb.MultiBandOpenCV()
    b.prepare(roi)
    for img, mask, tl in feeds:
        b.feed(img, mask, tl)
    dst, mask = b.blend()
    assert b.num_bands == nb
    assert np.array_equal(mask, wmask) and np.array_equal(dst, want)
    assert np.array_equal(mask, g[f"mask_{seed}"])
    ok, d = close(dst, g[f"dst_{seed}"])
    assert ok, (int(d.max()), float((d > 0).mean()))


@pytest.mark.gpu
def test_device_strided_inputs_band_counts_and_errors():
    import xpano_b200 as xb
    roi, feeds = mbcv_scene(21, n=3, w=150, h=97, step=100)
    for nb in (0, 2, 7):
        want, wmask, got_nb = oracle_blend(roi, feeds, nb)
        b = xb.MultiBandOpenCV(num_bands=nb)
        b.prepare(roi)
        for img, mask, tl in feeds:
            big = np.zeros((img.shape[0] + 2, img.shape[1] + 5, 3), np.uint8)
            big[1:-1, 2:-3] = img
            bm = np.zeros((mask.shape[0] + 3, mask.shape[1] + 4), np.uint8)
            bm[2:-1, 1:-3] = mask
            b.feed(big[1:-1, 2:-3], bm[2:-1, 1:-3], tl)
        dst, mask = b.blend()
        assert b.num_bands == got_nb
        assert np.array_equal(mask, wmask) and np.array_equal(dst, want)
    b = xb.MultiBandOpenCV()
    with pytest.raises(xb._lib.XpError):
        b.feed(feeds[0][0], feeds[0][1], feeds[0][2])    # before prepare
    b.prepare(roi)
    with pytest.raises(xb._lib.XpError):
        b.feed(feeds[0][0], feeds[0][1], (10 ** 6, 0))   # outside the ROI
    b.feed(*feeds[0])
    b.blend()
    with pytest.raises(xb._lib.XpError):
        b.blend()                                        # single use
    with pytest.raises(xb._lib.XpError):
        b.feed(*feeds[1])


@pytest.mark.gpu
def test_device_on_a_composed_panorama():
    """The stitcher's loop (stitcher.cc:357-419) with this blender installed:
    warp on the device, feed, blend; against the oracle fed with the same
    warped images."""
    import math
    import compose_oracle as co
    import synth
    import xpano_b200 as xb
    n_cols, w_, h_, f = 4, 400, 300, 300.0
    hfov = 2 * math.atan(w_ / 2 / f)
    Ks, Rs = co.make_cameras(n_cols, 1, w_, h_, f, 0.62 * hfov, [0.0], 17,
                             None)
    ims = [synth.texture(h_, w_, 900 + i, 3, 8) for i in range(n_cols)]
    warper = xb.RotationWarper(xb.ProjectionType.kSpherical, f)
    feeds, rois = [], []
    for im, K, R in zip(ims, Ks, Rs):
        corner, wi = warper.warp(im, K, R)
        _, wm = warper.warp(np.full((h_, w_), 255, np.uint8), K, R,
                            xb.INTER_NEAREST, xb.BORDER_CONSTANT)
        feeds.append((wi, wm, corner))
        rois.append((corner[0], corner[1], wi.shape[1], wi.shape[0]))
    x0 = min(r[0] for r in rois)
    y0 = min(r[1] for r in rois)
    x1 = max(r[0] + r[2] for r in rois)
    y1 = max(r[1] + r[3] for r in rois)
    roi = (x0, y0, x1 - x0, y1 - y0)
    want, wmask, _ = oracle_blend(roi, feeds)
    b = xb.MultiBandOpenCV()
    b.prepare(roi)
    for f_ in feeds:
        b.feed(*f_)
    dst, mask = b.blend()
    assert np.array_equal(mask, wmask) and np.array_equal(dst, want)
    assert mask.mean() > 0.5 * 255


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(12))
def test_device_random_cases(seed):
    """Seeded random layouts, band counts and fractional masks: the device
    result is the oracle's bit for bit."""
    import xpano_b200 as xb
    rng = np.random.default_rng(400 + seed)
    n = int(rng.integers(1, 6))
    nb = int(rng.integers(0, 8))
    feeds = []
    for i in range(n):
        w, h = int(rng.integers(17, 260)), int(rng.integers(9, 200))
        img = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
        m = rng.choice(np.array([0, 64, 128, 255], np.uint8), (h, w),
                       p=[0.1, 0.1, 0.1, 0.7])
        feeds.append((img, m, (int(rng.integers(-150, 300)),
                               int(rng.integers(-80, 120)))))
    x0 = min(f[2][0] for f in feeds)
    y0 = min(f[2][1] for f in feeds)
    x1 = max(f[2][0] + f[0].shape[1] for f in feeds)
    y1 = max(f[2][1] + f[0].shape[0] for f in feeds)
    roi = (x0, y0, x1 - x0, y1 - y0)
    want, wmask, got_nb = oracle_blend(roi, feeds, nb)
    b = xb.MultiBandOpenCV(num_bands=nb)
    b.prepare(roi)
    for f in feeds:
        b.feed(*f)
    dst, mask = b.blend()
    assert b.num_bands == got_nb
    assert np.array_equal(mask, wmask) and np.array_equal(dst, want)

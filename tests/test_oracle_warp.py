"""Pins the warp oracle (oracle/warp_oracle.c, a CPU restatement of OpenCV's
RotationWarper + remap + resize(INTER_LINEAR_EXACT) + dilate) against vectors
recorded from cv2 4.13.0 (tests/golden/warp_cv2_goldens.npz) and, when cv2 is
importable, against cv2 live.

OpenCV is a third-party dependency of the reference, not vendored under
/root/reference (CI pins 4.11.0, .github/workflows/test.yml:16); the call
sites being matched are xpano/algorithm/stitcher.cc:160, 377-401.

Bars: ROIs, maps, remap (u8/u16/nearest), exact resize, dilate: bit-exact.
Float gain-map resize: 1e-5 relative (cv2 itself is ~15 ulp off the exact
value there).
"""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from golden.make_golden import (RESIZE_CASES, WARP_NAMES, WARP_SCALE,  # noqa
                                WARP_SIZE, WARP_STEP, rot)
from oracle import warp_oracle as wo  # noqa: E402

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLD, "warp_cv2_goldens.npz"))


@pytest.mark.parametrize("name", WARP_NAMES)
def test_roi_maps_points_golden(g, name):
    K, R = g["K"], g["R"]
    w = wo.Warper.from_cv2_name(name, WARP_SCALE)
    assert tuple(g[f"roi_{name}"]) == w.warp_roi(WARP_SIZE, K, R)
    roi, xm, ym = w.build_maps(WARP_SIZE, K, R)
    assert np.array_equal(xm[::WARP_STEP, ::WARP_STEP], g[f"xmap_{name}"])
    assert np.array_equal(ym[::WARP_STEP, ::WARP_STEP], g[f"ymap_{name}"])
    pt = np.array(w.warp_point((123.5, 77.25), K, R), np.float32)
    assert np.array_equal(pt, g[f"pt_{name}"])


@pytest.mark.parametrize("name", ["spherical", "cylindrical", "mercator"])
def test_warp_golden(g, name):
    K, R = g["K"], g["R"]
    w = wo.Warper.from_cv2_name(name, WARP_SCALE)
    corner, d8 = w.warp(g["src8"], K, R)
    assert corner == tuple(g[f"corner_{name}"])
    assert np.array_equal(d8, g[f"warp8_{name}"])
    _, m = w.warp_mask(WARP_SIZE, K, R)
    assert np.array_equal(m, g[f"warpmask_{name}"])
    if name == "spherical":
        _, d16 = w.warp(g["src16"], K, R)
        assert np.array_equal(d16, g["warp16_spherical"])


def test_remap_wild_maps_golden(g):
    xm, ym = g["wild_x"], g["wild_y"]
    assert np.array_equal(wo.remap_linear_reflect(g["src8"], xm, ym),
                          g["wild8"])
    assert np.array_equal(wo.remap_linear_reflect(g["src16"], xm, ym),
                          g["wild16"])
    assert np.array_equal(wo.remap_nearest_mask(WARP_SIZE, xm, ym),
                          g["wildmask"])


@pytest.mark.parametrize("k", range(len(RESIZE_CASES)))
def test_resize_dilate_golden(g, k):
    sw, sh, dw, dh = RESIZE_CASES[k]
    src = g[f"rs_src_{k}"]
    assert np.array_equal(wo.resize_linear_exact_u8(src, (dw, dh)),
                          g[f"rs_dst_{k}"])
    assert np.array_equal(wo.dilate3x3_u8(src), g[f"dil_{k}"])
    got = wo.resize_linear_f32(g[f"gain_src_{k}"], (dw, dh))
    np.testing.assert_allclose(got, g[f"gain_dst_{k}"], rtol=1e-5)


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
@pytest.mark.parametrize("name", sorted(wo.CV2_NAMES))
def test_live_cv2_maps(name):
    w, h = 333, 211
    K = np.array([[290, 0, w / 2 + 3], [0, 285, h / 2 - 2], [0, 0, 1]],
                 np.float32)
    for ypr in [(0, 0, 0), (-0.7, 0.3, -0.2), (1.9, -0.6, 0.4)]:
        R = rot(*ypr)
        cw = cv2.PyRotationWarper(name, 240.0)
        ow = wo.Warper.from_cv2_name(name, 240.0)
        roi = tuple(cw.warpRoi((w, h), K, R))
        assert roi == ow.warp_roi((w, h), K, R)
        if roi[2] * roi[3] > 2e7:
            continue
        _, xm, ym = cw.buildMaps((w, h), K, R)
        _, xo, yo = ow.build_maps((w, h), K, R)
        assert np.array_equal(xm, xo, equal_nan=True)
        assert np.array_equal(ym, yo, equal_nan=True)


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_live_cv2_gain_apply():
    rng = np.random.default_rng(1)
    h, w = 171, 211
    img = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
    gsmall = rng.uniform(0.7, 1.4, (13, 17)).astype(np.float32)
    comp = cv2.detail_BlocksGainCompensator()
    comp.setMatGains([gsmall])
    ref = comp.apply(0, (0, 0), img.copy(), np.full((h, w), 255, np.uint8))
    gm = cv2.resize(gsmall, (w, h), interpolation=cv2.INTER_LINEAR)
    assert np.array_equal(wo.apply_gain_u8(img, gm), ref)
    mine = wo.apply_gain_u8(img, wo.resize_linear_f32(gsmall, (w, h)))
    d = np.abs(mine.astype(int) - ref.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3


PORTRAIT_PROJ = {"planePortrait": wo.PLANE, "cylindricalPortrait":
                 wo.CYLINDRICAL, "sphericalPortrait": wo.SPHERICAL}


def _portrait_keys():
    gp = np.load(os.path.join(GOLD, "warp_portrait_goldens.npz"))
    return sorted(k[4:] for k in gp.files if k.startswith("roi_"))


@pytest.mark.parametrize("key", _portrait_keys())
def test_portrait_warpers_golden(key):
    """The plain portrait warpers of xpano/algorithm/warpers.h:9-31 are not
    constructible through cv2; their goldens come from cv2's landscape
    warpers through the row-swap symmetry of OpenCV's portrait projectors
    (tests/golden/make_golden.py portrait_goldens). Bit-exact."""
    gp = np.load(os.path.join(GOLD, "warp_portrait_goldens.npz"))
    g8 = np.load(os.path.join(GOLD, "warp_cv2_goldens.npz"))["src8"]
    name, k = key.rsplit("_", 1)
    K, R = gp["K"], gp[f"R_{k}"]
    w = wo.Warper(PORTRAIT_PROJ[name], WARP_SCALE, True)
    assert w.warp_roi(WARP_SIZE, K, R) == tuple(gp[f"roi_{key}"])
    _, xm, ym = w.build_maps(WARP_SIZE, K, R)
    assert np.array_equal(xm[::WARP_STEP, ::WARP_STEP], gp[f"xmap_{key}"],
                          equal_nan=True)
    assert np.array_equal(ym[::WARP_STEP, ::WARP_STEP], gp[f"ymap_{key}"],
                          equal_nan=True)
    pt = np.array(w.warp_point((123.5, 77.25), K, R), np.float32)
    assert np.array_equal(pt, gp[f"pt_{key}"])
    corner, d8 = w.warp(g8, K, R)
    assert corner == tuple(gp[f"roi_{key}"][:2])
    assert np.array_equal(d8, gp[f"warp8_{key}"])
    _, m = w.warp_mask(WARP_SIZE, K, R)
    assert np.array_equal(m, gp[f"warpmask_{key}"])


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
@pytest.mark.parametrize("name", sorted(PORTRAIT_PROJ))
def test_live_cv2_portrait_symmetry(name):
    from golden.make_golden import P_SWAP, PORTRAIT_OF
    w, h = 333, 211
    K = np.array([[290, 0, w / 2 + 3], [0, 285, h / 2 - 2], [0, 0, 1]],
                 np.float32)
    for ypr in [(0, 0, 0), (-0.7, 0.3, -0.2), (0.3, 1.1, 0.2)]:
        R = rot(*ypr)
        cw = cv2.PyRotationWarper(PORTRAIT_OF[name], 240.0)
        ow = wo.Warper(PORTRAIT_PROJ[name], 240.0, True)
        x, y, rw, rh = cw.warpRoi((w, h), K, (P_SWAP @ R).astype(np.float32))
        assert ow.warp_roi((w, h), K, R) == (-(x + rw - 1), y, rw, rh)
        if rw * rh > 2e7:
            continue
        _, xm, ym = cw.buildMaps((w, h), K, (P_SWAP @ R).astype(np.float32))
        _, xo, yo = ow.build_maps((w, h), K, R)
        assert np.array_equal(xm[:, ::-1], xo, equal_nan=True)
        assert np.array_equal(ym[:, ::-1], yo, equal_nan=True)

"""CPU tests of the C-ABI library: it loads, exports every symbol that
include/xpano_b200.h declares, its host-side logic (scalar point maps, border
ROI detection, pyramid geometry, level count) matches the oracle, and compute
entry points fail loudly without a CUDA device (no CPU fallback)."""
import os
import re
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from golden.make_golden import WARP_NAMES, WARP_SCALE, WARP_SIZE, rot  # noqa
from oracle import mb_oracle  # noqa: E402
from oracle import warp_oracle as wo  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def xb():
    import __graft_entry__ as ge
    if not os.path.exists(os.path.join(ROOT, "xpano_b200",
                                       "libxpano_b200.so")):
        ge.build()
    import xpano_b200
    return xpano_b200


def test_exports_every_declared_symbol(xb):
    hdr = open(os.path.join(ROOT, "include", "xpano_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(xp_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 35
    lib = xb._lib.load()
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert declared == set(xb._lib.SYMBOLS), \
        declared ^ set(xb._lib.SYMBOLS)
    assert b"sm_100a" in lib.xp_version()


def _has_gpu(xb):
    return xb._lib.load().xp_device_count() > 0


@pytest.mark.parametrize("name", WARP_NAMES)
def test_host_point_maps_match_cv2_goldens(xb, name):
    g = np.load(os.path.join(ROOT, "tests", "golden", "warp_cv2_goldens.npz"))
    proj, portrait, a, b = wo.CV2_NAMES[name]
    w = xb.RotationWarper(proj, WARP_SCALE, bool(portrait), a or 2.0, b or 1.0)
    pt = np.array(w.warpPoint((123.5, 77.25), g["K"], g["R"]), np.float32)
    assert np.array_equal(pt, g[f"pt_{name}"])
    if f"ptb_{name}" in g:
        pb = np.array(w.warpPointBackward((10.0, 20.0), g["K"], g["R"]),
                      np.float32)
        assert np.array_equal(pb, g[f"ptb_{name}"])
    ow = wo.Warper.from_cv2_name(name, WARP_SCALE)
    for p in [(0.0, 0.0), (55.5, 10.25), (199.0, 149.0)]:
        assert w.warpPoint(p, g["K"], g["R"]) == ow.warp_point(p, g["K"],
                                                               g["R"])
        assert np.array_equal(
            np.array(w.warpPointBackward(p, g["K"], g["R"])),
            np.array(ow.warp_point_backward(p, g["K"], g["R"])),
            equal_nan=True)


@pytest.mark.parametrize("name", ["plane", "cylindrical", "spherical"])
def test_host_roi_detection(xb, name):
    """Border / corner ROI detection is host arithmetic (as in OpenCV)."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "warp_cv2_goldens.npz"))
    proj, portrait, a, b = wo.CV2_NAMES[name]
    w = xb.RotationWarper(proj, WARP_SCALE)
    assert w.warpRoi(WARP_SIZE, g["K"], g["R"]) == tuple(g[f"roi_{name}"])
    K = np.array([[290, 0, 170], [0, 285, 100], [0, 0, 1]], np.float32)
    for ypr in [(0, 0, 0), (-0.7, 0.3, -0.2), (1.9, -0.6, 0.4),
                (0.1, 1.4, 0.0)]:
        R = rot(*ypr)
        for portrait in (False, True):
            if portrait and name == "plane" and not _has_gpu(xb):
                continue  # PlanePortrait maps every pixel: device work
            ow = wo.Warper(proj, 240.0, portrait)
            gw = xb.RotationWarper(proj, 240.0, portrait)
            assert gw.warpRoi((333, 211), K, R) == ow.warp_roi((333, 211), K,
                                                               R)
    w.setScale(12.5)
    assert w.getScale() == 12.5


def test_geometry_and_levels_match_oracle(xb):
    import ctypes
    lib = xb._lib.load()
    rng = np.random.default_rng(0)
    for _ in range(50):
        w, h = int(rng.integers(1, 5000)), int(rng.integers(1, 5000))
        x, y = int(rng.integers(0, 3000)), int(rng.integers(0, 3000))
        lv = int(rng.integers(1, 12))
        arrs = [(ctypes.c_int * 24)() for _ in range(7)]
        n = lib.xp_pyramid_geometry(w, h, lv, x, y, *arrs)
        want = mb_oracle.pyramid_geometry(w, h, lv, x, y)
        keys = ("w", "h", "x", "y", "xs", "ys", "pitch")
        got = [dict(zip(keys, [a[l] for a in arrs])) for l in range(n)]
        assert got == want
        k = int(rng.integers(1, 9))
        ws = [int(v) for v in rng.integers(1, 9000, k)]
        hs = [int(v) for v in rng.integers(1, 9000, k)]
        fl, al = int(rng.integers(0, 3)) * 3, int(rng.integers(0, 2))
        iw = (ctypes.c_int * k)(*ws)
        ih = (ctypes.c_int * k)(*hs)
        assert lib.xp_blend_levels(k, iw, ih, fl, al) == \
            mb_oracle.blend_levels(ws, hs, fl, al)


def test_no_cpu_fallback(xb):
    if _has_gpu(xb):
        pytest.skip("a CUDA device is present")
    b = xb.Multiblend()
    with pytest.raises(xb._lib.XpError) as e:
        b.feed_interleaved(np.full((8, 8, 4), 255, np.uint8), (0, 0))
    assert e.value.code == 2 and "no CPU fallback" in str(e.value)
    w = xb.RotationWarper(xb.ProjectionType.kSpherical, 100.0)
    K = np.array([[100, 0, 8], [0, 100, 8], [0, 0, 1]], np.float32)
    with pytest.raises(xb._lib.XpError):
        w.warp(np.zeros((16, 16, 3), np.uint8), K, np.eye(3, dtype=np.float32))
    with pytest.raises(xb._lib.XpError):
        # generic projections forward-map every pixel on the device
        xb.RotationWarper(xb.ProjectionType.kFisheye, 100.0).warpRoi(
            (16, 16), K, np.eye(3, dtype=np.float32))


def test_argument_errors(xb):
    lib = xb._lib.load()
    import ctypes
    h = ctypes.c_void_p()
    assert lib.xp_warper_create(99, 0, 1.0, 1.0, 1.0, 0, ctypes.byref(h)) == 1
    assert b"projection" in lib.xp_last_error()
    b = xb.Multiblend()
    with pytest.raises(ValueError):
        b.feed(np.zeros((4, 4), np.uint8), np.zeros((4, 4), np.uint8), (0, 0))


def test_size_cap_arithmetic_matches_the_reference(xb):
    """stitcher.cc:318-335: the capped warp scale is formed in float/double
    exactly as the reference mixes them; an all-double evaluation differs in
    the last bit for part of the inputs, which this test must see."""
    import ctypes
    import math
    import compose_oracle as co
    lib = xb._lib.load()
    rng = np.random.default_rng(11)
    differs_from_double = 0
    for _ in range(4000):
        wis = float(rng.uniform(200.0, 9000.0))
        cwa = float(rng.uniform(1.0, 9.0))
        w, h = int(rng.integers(500, 40000)), int(rng.integers(300, 12000))
        cap = float(rng.choice([0.35, 20, 50, 100, 1000]))
        got = ctypes.c_float()
        capped = lib.xp_compose_capped_scale(wis, cwa, w, h, cap,
                                             ctypes.byref(got))
        want, want_capped = co.capped_scale(wis, cwa, w, h, cap)
        assert bool(capped) == want_capped
        assert np.float32(got.value) == want, (wis, cwa, w, h, cap)
        if want_capped:
            mpx = w * h / 1e6
            old = np.float32(float(np.float32(wis * cwa)) *
                             float(np.float32(math.sqrt(cap / mpx))))
            differs_from_double += int(old != want)
    assert differs_from_double > 0
    # not capped: the scale is just float(wis * cwa); 0 switches the cap off
    got = ctypes.c_float()
    assert lib.xp_compose_capped_scale(1234.5678, 3.3, 20000, 5000, 0.0,
                                       ctypes.byref(got)) == 0
    assert np.float32(got.value) == np.float32(1234.5678 * 3.3)
    # exactly at the limit is not capped (pano_mpx > max_pano_mpx_)
    assert lib.xp_compose_capped_scale(1000.0, 1.0, 10000, 10000, 100.0,
                                       ctypes.byref(got)) == 0

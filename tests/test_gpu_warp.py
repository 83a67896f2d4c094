"""GPU parity tests of the warp half and of the compose driver, through the C
ABI, against the warp oracle (oracle/warp_oracle.c, pinned bit-exact against
cv2 4.13.0 vectors in tests/test_oracle_warp.py) and the cv2 goldens.

Bars (BASELINE.json north_star), written as the thresholds below. The
separable projections (spherical, cylindrical, Mercator -- the sines and
cosines of their maps come from the host's libm, like OpenCV's): maps, warped
masks and warped pixels bit-exact. The other projections (device libm): ROIs
and warped masks bit-exact except where a map value differs in the last bit
exactly at a rounding boundary (< 1e-4 of pixels); maps within 2e-5 relative;
warped pixels within 1 LSB (8-bit) / 1e-3 relative (16-bit) at every pixel;
composed panorama within 1 LSB (8-bit) / 1e-3 relative (16-bit) at >= 99.9 %
of pixels and PSNR >= 50 dB.
"""
import json
import math
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import compose_oracle as co  # noqa: E402
import synth  # noqa: E402
from golden.make_golden import (WARP_NAMES, WARP_SCALE, WARP_SIZE,  # noqa
                                WARP_STEP, rot)
from oracle import warp_oracle as wo  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REPORT = {}


def _note(key, **kw):
    REPORT[key] = kw
    out = os.path.join(os.path.dirname(GOLD), "..", "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, "parity_warp.json"), "w") as f:
            json.dump(REPORT, f, indent=1, default=str)
    except OSError:
        pass


def _warper(name, scale):
    import xpano_b200
    proj, portrait, a, b = wo.CV2_NAMES[name]
    return xpano_b200.RotationWarper(proj, scale, bool(portrait), a or 2.0,
                                     b or 1.0)


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLD, "warp_cv2_goldens.npz"))


@pytest.mark.parametrize("name", WARP_NAMES)
def test_roi_maps_points_vs_cv2_goldens(g, name):
    K, R = g["K"], g["R"]
    w = _warper(name, WARP_SCALE)
    roi = w.warpRoi(WARP_SIZE, K, R)
    assert roi == tuple(g[f"roi_{name}"])
    roi2, xm, ym = w.buildMaps(WARP_SIZE, K, R)
    assert roi2 == roi
    gx, gy = g[f"xmap_{name}"], g[f"ymap_{name}"]
    dx = np.abs(xm[::WARP_STEP, ::WARP_STEP] - gx)
    dy = np.abs(ym[::WARP_STEP, ::WARP_STEP] - gy)
    ok = np.isfinite(gx) & np.isfinite(gy)
    exact = float(((dx == 0) & (dy == 0))[ok].mean())
    _note(f"maps_{name}", max_dx=float(dx[ok].max()), max_dy=float(dy[ok].max()),
          bit_exact_fraction=exact)
    # relative to the magnitude of the coordinate (float32 ulps grow with it)
    mag = np.maximum(1.0, np.maximum(np.abs(gx), np.abs(gy)))
    assert (np.maximum(dx, dy)[ok] / mag[ok]).max() < 2e-5
    if name in ("spherical", "cylindrical", "mercator", "plane"):
        assert exact == 1.0  # trig from the host's libm, the rest is IEEE
    pt = np.array(w.warpPoint((123.5, 77.25), K, R), np.float32)
    assert np.array_equal(pt, g[f"pt_{name}"])  # host libm: bit-exact


@pytest.mark.parametrize("name", ["spherical", "cylindrical", "mercator"])
def test_warp_vs_cv2_goldens(g, name):
    import xpano_b200 as xb
    K, R = g["K"], g["R"]
    w = _warper(name, WARP_SCALE)
    corner, d8 = w.warp(g["src8"], K, R)
    assert corner == tuple(g[f"corner_{name}"])
    d = np.abs(d8.astype(int) - g[f"warp8_{name}"].astype(int))
    _, m = w.warp(np.full(WARP_SIZE[::-1], 255, np.uint8), K, R,
                  xb.INTER_NEAREST, xb.BORDER_CONSTANT)
    mdiff = float((m != g[f"warpmask_{name}"]).mean())
    _note(f"warp8_{name}", max_diff=int(d.max()), frac=float((d > 0).mean()),
          mask_diff_fraction=mdiff)
    assert d.max() == 0 and mdiff == 0.0   # bit-exact against cv2
    if name == "spherical":
        _, d16 = w.warp(g["src16"], K, R)
        d = np.abs(d16.astype(int) - g["warp16_spherical"].astype(int))
        _note("warp16_spherical", max_diff=int(d.max()),
              frac=float((d > 0).mean()))
        assert d.max() == 0


PORTRAIT_PROJ = {"planePortrait": wo.PLANE, "cylindricalPortrait":
                 wo.CYLINDRICAL, "sphericalPortrait": wo.SPHERICAL}


def _portrait_keys():
    gp = np.load(os.path.join(GOLD, "warp_portrait_goldens.npz"))
    return sorted(k[4:] for k in gp.files if k.startswith("roi_"))


@pytest.mark.parametrize("key", _portrait_keys())
def test_portrait_warpers_vs_goldens(g, key):
    """The three warper classes xpano itself defines (warpers.h:9-31):
    ROI, maps, point map, image warp and mask warp against goldens derived
    from cv2 (tests/golden/make_golden.py portrait_goldens)."""
    import xpano_b200 as xb
    gp = np.load(os.path.join(GOLD, "warp_portrait_goldens.npz"))
    name, k = key.rsplit("_", 1)
    K, R = gp["K"], gp[f"R_{k}"]
    w = xb.PickWarperPortrait(PORTRAIT_PROJ[name]).create(WARP_SCALE)
    roi = w.warpRoi(WARP_SIZE, K, R)
    assert roi == tuple(gp[f"roi_{key}"])
    _, xm, ym = w.buildMaps(WARP_SIZE, K, R)
    gx, gy = gp[f"xmap_{key}"], gp[f"ymap_{key}"]
    ok = np.isfinite(gx) & np.isfinite(gy)
    dx = np.abs(xm[::WARP_STEP, ::WARP_STEP] - gx)
    dy = np.abs(ym[::WARP_STEP, ::WARP_STEP] - gy)
    mag = np.maximum(1.0, np.maximum(np.abs(gx), np.abs(gy)))
    assert (np.maximum(dx, dy)[ok] / mag[ok]).max() < 2e-5
    if name in ("spherical", "cylindrical", "mercator", "plane"):
        assert exact == 1.0  # trig from the host's libm, the rest is IEEE
    pt = np.array(w.warpPoint((123.5, 77.25), K, R), np.float32)
    assert np.array_equal(pt, gp[f"pt_{key}"])
    corner, d8 = w.warp(g["src8"], K, R)
    assert corner == roi[:2]
    d = np.abs(d8.astype(int) - gp[f"warp8_{key}"].astype(int))
    _, m = w.warp(np.full(WARP_SIZE[::-1], 255, np.uint8), K, R,
                  xb.INTER_NEAREST, xb.BORDER_CONSTANT)
    mdiff = float((m != gp[f"warpmask_{key}"]).mean())
    _note(f"portrait_{key}", max_diff=int(d.max()),
          frac=float((d > 0).mean()), mask_diff_fraction=mdiff)
    assert d.max() <= 1 and (d > 0).mean() < 1e-3
    assert mdiff < 1e-4


@pytest.mark.parametrize("name", sorted(wo.CV2_NAMES))
def test_all_projections_vs_oracle(name):
    """Every projection (and portrait variant) with a rotated camera
    (pitch/yaw/roll reach the path only as R, algorithm.cc:362-369)."""
    w_, h_ = 333, 211
    K = np.array([[290, 0, w_ / 2 + 3], [0, 285, h_ / 2 - 2], [0, 0, 1]],
                 np.float32)
    src = synth.texture(h_, w_, 5, 3, 8)
    for ypr in [(0, 0, 0), (-0.7, 0.3, -0.2)]:
        R = rot(*ypr)
        ow = wo.Warper.from_cv2_name(name, 240.0)
        gw = _warper(name, 240.0)
        roi = ow.warp_roi((w_, h_), K, R)
        groi = gw.warpRoi((w_, h_), K, R)
        assert max(abs(a - b) for a, b in zip(roi, groi)) <= 1, (roi, groi)
        if roi != groi or roi[2] * roi[3] > 4e6:
            continue
        _, xo, yo = ow.build_maps((w_, h_), K, R)
        _, xg, yg = gw.buildMaps((w_, h_), K, R)
        ok = np.isfinite(xo) & np.isfinite(yo) & (np.abs(xo) < 1e5) \
            & (np.abs(yo) < 1e5)
        mag = np.maximum(1.0, np.maximum(np.abs(xo), np.abs(yo)))
        rel = np.maximum(np.abs(xg - xo), np.abs(yg - yo)) / mag
        assert rel[ok].max() < 5e-5, (name, ypr, rel[ok].max())
        _, want = ow.warp(src, K, R)
        _, got = gw.warp(src, K, R)
        d = np.abs(got.astype(int) - want.astype(int))
        # pixels whose map is sane: within 1 LSB, every one of them
        assert d[ok].max() <= 1, (name, ypr, int(d[ok].max()))
        assert (d[ok] > 0).mean() < 1e-3, (name, ypr, (d[ok] > 0).mean())
        _note(f"oracle_{name}_{ypr[0]}", max_map_rel=float(rel[ok].max()),
              px_diff_frac=float((d[ok] > 0).mean()),
              px_gt1_frac=float((d[ok] > 1).mean()))


def test_fused_bgra_vs_oracle():
    """warp + mask + gain + seam-mask upsample + pack in one kernel
    (stitcher.cc:377-405, blenders.cc:99-112)."""
    rng = np.random.default_rng(3)
    w_, h_ = 400, 300
    src = synth.texture(h_, w_, 9, 3, 8)
    K = np.array([[350, 0, w_ / 2], [0, 350, h_ / 2], [0, 0, 1]], np.float32)
    R = rot(0.3, -0.05, 0.02)
    gain = rng.uniform(0.8, 1.25, (9, 12)).astype(np.float32)
    seam = (rng.random((24, 31)) < 0.6).astype(np.uint8) * 255
    seam[:, :6] = 0
    res = co.compose([src], [K], [R], warp_scale=330.0, gains=[gain],
                     seam_masks=[seam], keep_fed=True, fixed_levels=1)
    want, corner = res["fed"][0]
    gw = _warper("spherical", 330.0)
    roi, got = gw.warp_bgra(src, K, R, gain=gain, seam_mask=seam)
    assert roi[:2] == corner and got.shape == want.shape
    adiff = float((got[..., 3] != want[..., 3]).mean())
    d = np.abs(got[..., :3].astype(int) - want[..., :3].astype(int))
    _note("fused_bgra", alpha_diff_fraction=adiff, max_diff=int(d.max()),
          frac=float((d > 0).mean()))
    assert adiff < 1e-4
    assert d.max() <= 1 and (d > 0).mean() < 1e-3


def _psnr(a, b, maxv):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10 * math.log10(maxv * maxv / mse)


@pytest.mark.parametrize("cfg", ["spherical_row", "cyl_rotated_u16",
                                 "mercator_grid_capped",
                                 "spherical_capped_doubles"])
def test_compose_vs_oracle(cfg):
    """Device compose driver against the CPU compose oracle (small versions
    of BASELINE.json configs 2-4)."""
    import xpano_b200 as xb
    if cfg == "spherical_row":
        n_cols, n_rows, w_, h_, f, bpp = 4, 1, 400, 300, 300.0, 8
        proj, kw, pitches = wo.SPHERICAL, dict(fixed_levels=5), [0.0]
        extra, cap = None, 100.0
    elif cfg == "cyl_rotated_u16":
        n_cols, n_rows, w_, h_, f, bpp = 3, 2, 360, 240, 320.0, 16
        proj, kw = wo.CYLINDRICAL, dict()
        pitches = [-0.25, 0.25]
        extra, cap = co.rot(math.radians(10), math.radians(5),
                            math.radians(3)), 100.0
    elif cfg == "spherical_capped_doubles":
        # the cap branch with warped_image_scale_ and compose_work_aspect
        # given as the two doubles of stitcher.cc:306-313; chosen so that an
        # all-double evaluation of the cap rounds to a different float scale
        n_cols, n_rows, w_, h_, f, bpp = 4, 1, 400, 300, 300.0, 8
        proj, kw, pitches = wo.SPHERICAL, dict(fixed_levels=3), [0.0]
        extra, cap = None, 0.23
        scales = dict(warped_image_scale=f * 0.6 / 1.6180339887,
                      compose_work_aspect=1.6180339887 / 0.6)
    else:
        n_cols, n_rows, w_, h_, f, bpp = 3, 2, 320, 240, 300.0, 8
        proj, kw = wo.MERCATOR, dict(fixed_levels=4)
        pitches = [-0.3, 0.3]
        extra, cap = None, 0.35  # exercises the size-cap branch
    if cfg != "spherical_capped_doubles":
        scales = dict()
    hfov = 2 * math.atan(w_ / 2 / f)
    Ks, Rs = co.make_cameras(n_cols, n_rows, w_, h_, f, 0.62 * hfov, pitches,
                             17, extra)
    ims = [synth.texture(h_, w_, 100 + i, 3, bpp) for i in range(len(Ks))]
    want = co.compose(ims, Ks, Rs, projection=proj, warp_scale=f,
                      max_pano_mpx=cap, **scales, **kw)
    c = xb.Composer(projection=proj, warp_scale=f, max_pano_mpx=cap,
                    depth=bpp, **scales, **kw)
    c.set_images(ims, Ks, Rs)
    info = c.run()
    pano, mask = c.fetch()
    assert info["resolution_capped"] == want["capped"]
    assert np.float32(info["warp_scale"]) == np.float32(want["warp_scale"])
    assert (info["width"], info["height"]) == want["pano"].shape[1::-1]
    assert (info["min_x"], info["min_y"]) == (want["min_x"], want["min_y"])
    mdiff = float((mask != want["mask"]).mean())
    maxv = 255 if bpp == 8 else 65535
    both = (mask > 0) & (want["mask"] > 0)
    d = np.abs(pano.astype(int) - want["pano"].astype(int))[both]
    # 1 LSB for 8-bit samples, 1e-3 relative for 16-bit ones
    tol = 1 if bpp == 8 else np.maximum(1, 1e-3 * want["pano"].astype(int)
                                        [both])
    within = float((d <= tol).mean())
    assert d.max() <= (1 if bpp == 8 else 66), d.max()  # 1e-3 of full scale
    psnr = _psnr(pano[both], want["pano"][both], maxv)
    _note(f"compose_{cfg}", size=(info["width"], info["height"]),
          mask_diff_fraction=mdiff, within_1lsb=within, psnr=psnr,
          max_diff=int(d.max()), levels=info["levels"])
    assert mdiff < 1e-4
    assert within >= 0.999 and psnr >= 50.0, (within, psnr)

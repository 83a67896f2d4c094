"""Regenerates tests/golden/*.npz. Run in the build container only (it reads
/root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

blend_reference_goldens.npz
    The reference's own golden vectors for the blend path
    (external/multiblend/tests/data, used by tests/multiblend_test.cpp:39-44
    and tests/test_multiblend.py:47-185), decoded to arrays: inputs, positions
    and expected outputs for expected.png, expected_alpha.png,
    expected_seams.png, expected.tif and expected_16bit.tif.

warp_cv2_goldens.npz
    Vectors recorded from Python cv2 for the OpenCV half of the path
    (RotationWarper maps/ROIs/points, remap, resize(INTER_LINEAR_EXACT),
    dilate, float resize).

blend_ref_outputs.npz
    Outputs of the UNMODIFIED reference library (oracle/_ref/libmb_ref.so) on
    the seeded synthetic cases of tests/synth.py (inputs are regenerated from
    the seed by the tests, only outputs are stored).
"""
import os
import sys

import cv2
import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import mb_ref  # noqa: E402
import synth  # noqa: E402

D = "/root/reference/external/multiblend/tests/data/"

# (seed, kwargs for synth.blend_case, kwargs for blend)
REF_CASES = [
    (1, dict(n=2, mode="ellipse", bpp=8), dict()),
    (2, dict(n=3, mode="full", bpp=8), dict(fixed_levels=3)),
    (3, dict(n=5, mode="blobs", bpp=8), dict(fixed_levels=5)),
    (4, dict(n=4, mode="ellipse", bpp=16), dict()),
    (5, dict(n=3, mode="blobs", bpp=16), dict(fixed_levels=4, dither=False)),
    (6, dict(n=6, mode="ellipse", bpp=8, w=320, h=240), dict(fixed_levels=7)),
    (7, dict(n=3, mode="full", bpp=8, spp=3), dict()),
    (8, dict(n=4, mode="ellipse", bpp=8), dict(reverse=True, fixed_levels=2)),
    (9, dict(n=1, mode="ellipse", bpp=8), dict()),
    (10, dict(n=3, mode="ellipse", bpp=8), dict(fixed_levels=1)),
    (11, dict(n=3, mode="ellipse", bpp=8), dict(wideblend=True)),
]


def rot(yaw, pitch, roll):
    cy, sy = np.cos(yaw), np.sin(yaw)
    cp, sp = np.cos(pitch), np.sin(pitch)
    cr, sr = np.cos(roll), np.sin(roll)
    ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    rz = np.array([[cr, -sr, 0], [sr, cr, 0], [0, 0, 1]])
    return (ry @ rx @ rz).astype(np.float32)


WARP_NAMES = ["plane", "cylindrical", "spherical", "fisheye", "stereographic",
              "compressedPlaneA2B1", "compressedPlaneA1.5B1",
              "compressedPlanePortraitA2B1", "paniniA2B1", "paniniA1.5B1",
              "paniniPortraitA2B1", "mercator", "transverseMercator"]
WARP_SIZE = (200, 150)
WARP_SCALE = 160.0
WARP_ROT = (0.4, -0.1, 0.05)
WARP_STEP = 5          # maps are stored subsampled
RESIZE_CASES = [(57, 43, 311, 271), (33, 21, 330, 210), (40, 30, 41, 31),
                (1, 5, 7, 9), (64, 48, 32, 24)]


def warp_goldens():
    """Vectors recorded from Python cv2 (the OpenCV half of the path is a
    third-party dependency)."""
    w, h = WARP_SIZE
    K = np.array([[150, 0, w / 2], [0, 150, h / 2], [0, 0, 1]], np.float32)
    R = rot(*WARP_ROT)
    g = {"K": K, "R": R, "cv2_version": np.array(cv2.__version__)}
    rng = np.random.default_rng(5)
    src8 = synth.texture(h, w, 77, 3, 8)
    src16 = synth.texture(h, w, 78, 3, 16)
    g["src8"], g["src16"] = src8, src16
    for name in WARP_NAMES:
        cw = cv2.PyRotationWarper(name, WARP_SCALE)
        g[f"roi_{name}"] = np.array(cw.warpRoi((w, h), K, R))
        _, xm, ym = cw.buildMaps((w, h), K, R)
        g[f"xmap_{name}"] = xm[::WARP_STEP, ::WARP_STEP].copy()
        g[f"ymap_{name}"] = ym[::WARP_STEP, ::WARP_STEP].copy()
        g[f"pt_{name}"] = np.array(cw.warpPoint((123.5, 77.25), K, R),
                                   np.float32)
        try:  # not implemented for every warper in the Python binding
            g[f"ptb_{name}"] = np.array(
                cw.warpPointBackward((10.0, 20.0), K, R), np.float32)
        except cv2.error:
            pass
    for name in ("spherical", "cylindrical", "mercator"):
        cw = cv2.PyRotationWarper(name, WARP_SCALE)
        c, d8 = cw.warp(src8, K, R, cv2.INTER_LINEAR, cv2.BORDER_REFLECT)
        if name == "spherical":
            _, d16 = cw.warp(src16, K, R, cv2.INTER_LINEAR,
                             cv2.BORDER_REFLECT)
            g[f"warp16_{name}"] = d16
        _, m = cw.warp(np.full((h, w), 255, np.uint8), K, R,
                       cv2.INTER_NEAREST, cv2.BORDER_CONSTANT)
        g[f"warp8_{name}"] = d8
        g[f"warpmask_{name}"] = m
        g[f"corner_{name}"] = np.array(c)
    # remap on wild maps (far outside, integers, -1)
    xm = rng.uniform(-3 * w, 4 * w, (60, 90)).astype(np.float32)
    ym = rng.uniform(-3 * h, 4 * h, (60, 90)).astype(np.float32)
    xm[:15] = np.round(xm[:15])
    ym[:15] = np.round(ym[:15])
    xm[15:20] = ym[15:20] = -1
    g["wild_x"], g["wild_y"] = xm, ym
    g["wild8"] = cv2.remap(src8, xm, ym, cv2.INTER_LINEAR,
                           borderMode=cv2.BORDER_REFLECT)
    g["wild16"] = cv2.remap(src16, xm, ym, cv2.INTER_LINEAR,
                            borderMode=cv2.BORDER_REFLECT)
    g["wildmask"] = cv2.remap(np.full((h, w), 255, np.uint8), xm, ym,
                              cv2.INTER_NEAREST,
                              borderMode=cv2.BORDER_CONSTANT)
    for k, (sw, sh, dw, dh) in enumerate(RESIZE_CASES):
        m = (rng.random((sh, sw)) < 0.5).astype(np.uint8) * 255
        g[f"rs_src_{k}"] = m
        g[f"rs_dst_{k}"] = cv2.resize(m, (dw, dh),
                                      interpolation=cv2.INTER_LINEAR_EXACT)
        g[f"dil_{k}"] = cv2.dilate(m, None)
        gm = rng.uniform(0.7, 1.4, (sh, sw)).astype(np.float32)
        g[f"gain_src_{k}"] = gm
        g[f"gain_dst_{k}"] = cv2.resize(gm, (dw, dh),
                                        interpolation=cv2.INTER_LINEAR)
    np.savez_compressed(os.path.join(HERE, "warp_cv2_goldens.npz"), **g)


# The three portrait warpers xpano defines itself (xpano/algorithm/warpers.h:
# 9-31: PlanePortraitWarper, CylindricalPortraitWarper, SphericalPortraitWarper
# creators) exist only in OpenCV's C++ detail namespace; cv2.PyRotationWarper
# cannot construct them. OpenCV's portrait projectors (warpers_inl.hpp) swap
# the first two rows of r_kinv, evaluate the landscape formula and negate u.
# Swapping those rows is R' = P R with P = [[0,1,0],[1,0,0],[0,0,1]] (then
# k_rinv' = K R'^T has its first two columns swapped, which is exactly the
# portrait mapBackward, and rinv' has the pole column the portrait ROI test
# uses), so cv2's LANDSCAPE warper called with R' yields the portrait result
# mirrored in u: roi.x = -(x' + w' - 1), maps / warped images flipped
# left-right, point.u = -u'.
PORTRAIT_OF = {"planePortrait": "plane", "cylindricalPortrait": "cylindrical",
               "sphericalPortrait": "spherical"}
PORTRAIT_ROTS = [(0.4, -0.1, 0.05), (-0.2, 0.9, 0.1)]  # 2nd: pole in view
P_SWAP = np.array([[0, 1, 0], [1, 0, 0], [0, 0, 1]], np.float32)


def portrait_goldens():
    w, h = WARP_SIZE
    K = np.array([[150, 0, w / 2], [0, 150, h / 2], [0, 0, 1]], np.float32)
    g = {"K": K, "cv2_version": np.array(cv2.__version__)}
    src8 = synth.texture(h, w, 77, 3, 8)  # = warp_cv2_goldens.npz "src8"
    for k, ypr in enumerate(PORTRAIT_ROTS):
        R = rot(*ypr)
        g[f"R_{k}"] = R
        Rp = (P_SWAP @ R).astype(np.float32)
        for name, land in PORTRAIT_OF.items():
            cw = cv2.PyRotationWarper(land, WARP_SCALE)
            x, y, rw, rh = cw.warpRoi((w, h), K, Rp)
            if rw * rh > 4e6:
                continue
            key = f"{name}_{k}"
            g[f"roi_{key}"] = np.array([-(x + rw - 1), y, rw, rh])
            _, xm, ym = cw.buildMaps((w, h), K, Rp)
            g[f"xmap_{key}"] = xm[:, ::-1][::WARP_STEP, ::WARP_STEP].copy()
            g[f"ymap_{key}"] = ym[:, ::-1][::WARP_STEP, ::WARP_STEP].copy()
            u, v = cw.warpPoint((123.5, 77.25), K, Rp)
            g[f"pt_{key}"] = np.array([-u, v], np.float32)
            c, d8 = cw.warp(src8, K, Rp, cv2.INTER_LINEAR, cv2.BORDER_REFLECT)
            _, m = cw.warp(np.full((h, w), 255, np.uint8), K, Rp,
                           cv2.INTER_NEAREST, cv2.BORDER_CONSTANT)
            g[f"warp8_{key}"] = np.ascontiguousarray(d8[:, ::-1])
            g[f"warpmask_{key}"] = np.ascontiguousarray(m[:, ::-1])
    np.savez_compressed(os.path.join(HERE, "warp_portrait_goldens.npz"), **g)


def crop_goldens():
    """The reference's "Real life example" of tests/auto_crop_test.cc:154-160
    (tests/data/mask.png, expected crop (67, 659)-(5985, 2950)) as a packed
    bit array."""
    m = cv2.imread("/root/reference/tests/data/mask.png",
                   cv2.IMREAD_UNCHANGED)
    np.savez_compressed(os.path.join(HERE, "crop_real_life_mask.npz"),
                        shape=np.array(m.shape), bits=np.packbits(m == 255),
                        expected=np.array([67, 659, 5985, 2950]))


def gain_scenes():
    """Two small exposure-compensation scenes (corners, images, masks):
    A: four overlapping exposures of one smooth texture with warped-looking
       masks (cut corners, empty margins) plus one isolated image;
    B: block-grid edge cases (33x20, 31x65, 100x31; one half-empty mask)."""
    rng = np.random.default_rng(11)
    noise = rng.integers(0, 256, (260, 520, 3)).astype(np.float32)
    base = noise.copy()
    for _ in range(6):  # smooth without cv2: repeated 5-point averaging
        base = (base + np.roll(base, 1, 0) + np.roll(base, -1, 0)
                + np.roll(base, 1, 1) + np.roll(base, -1, 1)) / 5
    base = (base - base.min()) / (base.max() - base.min()) * 255
    cs, ims, ms = [], [], []
    for k, (x, y, w, h) in enumerate([(0, 10, 170, 120), (120, 0, 180, 131),
                                      (250, 30, 160, 110),
                                      (60, 90, 130, 70)]):
        im = np.clip(base[y:y + h, x:x + w] * (0.75 + 0.17 * k)
                     + rng.normal(0, 2, (h, w, 3)), 0, 255).astype(np.uint8)
        m = np.full((h, w), 255, np.uint8)
        m[:5] = 0
        m[:, :4] = 0
        yy, xx = np.mgrid[0:h, 0:w]
        m[(xx + yy) < 40] = 0
        m[(w - xx) + (h - yy) < 25] = 0
        cs.append((x - 50, y + 3))
        ims.append(im)
        ms.append(m)
    ims.append(np.full((70, 80, 3), 99, np.uint8))
    ms.append(np.full((70, 80), 255, np.uint8))
    cs.append((2000, 2000))
    a = (cs, ims, ms)
    ims = [rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
           for (w, h) in [(33, 20), (31, 65), (100, 31)]]
    ms = [np.full(i.shape[:2], 255, np.uint8) for i in ims]
    ms[1][:30] = 0
    b = ([(0, 0), (20, -5), (10, 10)], ims, ms)
    return {"A": a, "B": b}


def gain_goldens():
    """cv::detail::BlocksGainCompensator (cv2 4.13.0) on gain_scenes():
    getMatGains with 0 and 2 filtering iterations, and apply on every image
    (8-bit: BlocksCompensator::apply asserts CV_8UC3)."""
    g = {}
    for name, (cs, ims, ms) in gain_scenes().items():
        g[f"{name}_corners"] = np.array(cs, np.int32)
        for i, (im, m) in enumerate(zip(ims, ms)):
            g[f"{name}_img_{i}"] = im
            g[f"{name}_mask_{i}"] = m
        for it in (0, 2):
            comp = cv2.detail_BlocksGainCompensator()
            comp.setNrGainsFilteringIterations(it)
            comp.feed(cs, ims, ms)
            for i, gm in enumerate(comp.getMatGains()):
                g[f"{name}_gain{it}_{i}"] = np.ascontiguousarray(gm)
        for i, (im, m) in enumerate(zip(ims, ms)):
            g[f"{name}_applied_{i}"] = comp.apply(i, cs[i], im.copy(), m)
    np.savez_compressed(os.path.join(HERE, "gain_cv2_goldens.npz"), **g)


SEAM_SCENE = dict(n_cols=4, n_rows=1, w=400, h=300, f=300.0, step=0.62,
                  pitches=[0.0], seed=23)


def seam_scene():
    """A registered 4-image row at the seam scale (the images double as
    seam_est_imgs_, seam_work_aspect = 1) with per-image exposure errors."""
    import math
    import compose_oracle as co
    import synth
    c = SEAM_SCENE
    hfov = 2 * math.atan(c["w"] / 2 / c["f"])
    Ks, Rs = co.make_cameras(c["n_cols"], c["n_rows"], c["w"], c["h"], c["f"],
                             c["step"] * hfov, c["pitches"], c["seed"], None)
    ims = []
    for i in range(len(Ks)):
        t = synth.texture(c["h"], c["w"], 300 + i, 3, 8).astype(np.float32)
        ims.append(np.clip(t * (0.8 + 0.12 * i), 0, 255).astype(np.uint8))
    return ims, Ks, Rs


def seam_goldens():
    """Stitcher::EstimateSeams (stitcher.cc:241-296) with cv2 4.13.0 on
    seam_scene(): corners, gain maps, the compensated warped images and the
    GraphCutSeamFinder(COST_COLOR) masks."""
    ims, Ks, Rs = seam_scene()
    warper = cv2.PyRotationWarper("spherical", SEAM_SCENE["f"])
    corners, warped, masks = [], [], []
    for im, K, R in zip(ims, Ks, Rs):
        k = np.asarray(K, np.float32)
        c, iw = warper.warp(im, k, R, cv2.INTER_LINEAR, cv2.BORDER_REFLECT)
        _, mw = warper.warp(np.full(im.shape[:2], 255, np.uint8), k, R,
                            cv2.INTER_NEAREST, cv2.BORDER_CONSTANT)
        corners.append(c)
        warped.append(iw)
        masks.append(mw)
    comp = cv2.detail_BlocksGainCompensator()
    comp.feed(corners, warped, masks)
    warped = [comp.apply(i, corners[i], warped[i], masks[i])
              for i in range(len(ims))]
    g = {"corners": np.array(corners, np.int32)}
    for i, gm in enumerate(comp.getMatGains()):
        g[f"gain_{i}"] = np.ascontiguousarray(gm)
        g[f"warped_{i}"] = warped[i]
    finder = cv2.detail_GraphCutSeamFinder("COST_COLOR")
    out = finder.find([w.astype(np.float32) for w in warped], corners, masks)
    for i, m in enumerate(out):
        g[f"seam_{i}"] = m.get() if hasattr(m, "get") else np.asarray(m)
    np.savez_compressed(os.path.join(HERE, "seam_cv2_goldens.npz"), **g)


def mbcv_scene(seed, n=3, w=200, h=150, step=130):
    """Images, masks (with fractional values on one image) and corners for
    the MultiBandOpenCV tests; regenerated from the seed by the tests."""
    feeds = []
    for i in range(n):
        img = synth.texture(h, w, seed * 10 + i, 3, 8)
        m = np.full((h, w), 255, np.uint8)
        yy, xx = np.mgrid[0:h, 0:w]
        m[(xx + yy) < 30] = 0
        m[:, : 3 + i] = 0
        if i == 1:
            m[h // 2:, w // 2:] = 128        # a half-weight region
        feeds.append((img, m, (i * step - 40, 7 * i - 5)))
    x0 = min(f[2][0] for f in feeds)
    y0 = min(f[2][1] for f in feeds)
    x1 = max(f[2][0] + f[0].shape[1] for f in feeds)
    y1 = max(f[2][1] + f[0].shape[0] for f in feeds)
    return (x0, y0, x1 - x0, y1 - y0), feeds


MBCV_CASES = [(1, dict()), (2, dict(n=4, w=331, h=207, step=250)),
              (3, dict(n=2, w=64, h=40, step=30)),
              (4, dict(n=1, w=20, h=9, step=0))]


def mbcv_goldens():
    """blenders::MultiBandOpenCV as xpano runs it (blenders.cc:160-175) with
    cv2 4.13.0: MultiBandBlender(0, 5, CV_32F), feed(int16 image),
    blend, convert to 8 bits."""
    g = {}
    for seed, kw in MBCV_CASES:
        roi, feeds = mbcv_scene(seed, **kw)
        b = cv2.detail_MultiBandBlender(0, 5, cv2.CV_32F)
        b.prepare(roi)
        for img, mask, tl in feeds:
            b.feed(img.astype(np.int16), mask, tl)
        dst, m = b.blend(None, None)
        g[f"dst_{seed}"] = np.clip(dst, 0, 255).astype(np.uint8)
        g[f"mask_{seed}"] = m
    np.savez_compressed(os.path.join(HERE, "mbcv_cv2_goldens.npz"), **g)


def main():
    warp_goldens()
    portrait_goldens()
    crop_goldens()
    gain_goldens()
    seam_goldens()
    mbcv_goldens()
    g = {}
    pos_png = [(-91, 204), (-167, 202), (-41, 193)]
    pos_tif = [(76, 11), (0, 9), (126, 0)]
    for i in range(3):
        raw, p = mb_ref.read_cereal_raw(D + f"img_{i}.png.raw")
        assert p == pos_png[i]
        g[f"raw_{i}"] = raw                                   # RGB order
        g[f"alpha_{i}"] = cv2.imread(D + f"img_{i}_alpha.png",
                                     cv2.IMREAD_UNCHANGED)    # BGRA
        g[f"tif16_{i}"] = cv2.imread(D + f"img_{i}_16bit.tif",
                                     cv2.IMREAD_UNCHANGED)    # BGR u16
    g["pos_png"] = np.array(pos_png)
    g["pos_tif"] = np.array(pos_tif)
    g["expected_png"] = cv2.imread(D + "expected.png", cv2.IMREAD_UNCHANGED)
    g["expected_alpha_png"] = cv2.imread(D + "expected_alpha.png",
                                         cv2.IMREAD_UNCHANGED)
    g["expected_16bit_tif"] = cv2.imread(D + "expected_16bit.tif",
                                         cv2.IMREAD_UNCHANGED)
    g["expected_seams_png"] = np.array(Image.open(D + "expected_seams.png"))
    # multiblend_test.cpp:39-44
    g["kat_size_minpos"] = np.array([331, 244, -167, 193])
    np.savez_compressed(os.path.join(HERE, "blend_reference_goldens.npz"), **g)

    r = {}
    for seed, ckw, bkw in REF_CASES:
        ims, pos = synth.blend_case(seed, **ckw)
        res = mb_ref.blend(ims, pos, n_threads=4, **bkw)
        r[f"planes_{seed}"] = res.planes
        r[f"mask_{seed}"] = res.mask
        r[f"meta_{seed}"] = np.array([res.min_x, res.min_y, int(res.no_mask)])
    np.savez_compressed(os.path.join(HERE, "blend_ref_outputs.npz"), **r)
    for f in ("blend_reference_goldens.npz", "blend_ref_outputs.npz",
              "warp_cv2_goldens.npz", "warp_portrait_goldens.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)) // 1024, "KiB")


if __name__ == "__main__":
    main()



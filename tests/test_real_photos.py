"""The first real-photograph fixture (BASELINE.json configs[0]):
tests/data/image06-08.jpg of the reference, which its own pipeline test
stitches to a 976 x 1335 (+-2 %) preview panorama
(tests/stitcher_pipeline_test.cc:655-657, 718-719), with cameras from the
committed cv2.detail registration script (tests/golden/make_real_fixture.py).

CPU: the oracle chain (warp oracle + exposure oracle + multiblend oracle)
reproduces the reference's expected panorama size and cv2's gain maps.
GPU: EstimateSeams on the device (+ cv2's graph cut, the caller-side CPU step)
gives cv2's gain maps and seam masks; the composed panorama meets the
north-star bars against the oracle chain on the same inputs (mask identical,
>= 99.9 % of the pixels within 1 LSB, PSNR >= 50 dB; the spherical maps are
bit-exact, so every fed image is the oracle's apart from the roundings of the
gain multiply)."""
import hashlib
import math
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import compose_oracle as co  # noqa: E402
from oracle import gain_oracle as go  # noqa: E402
from oracle import warp_oracle as wo  # noqa: E402

cv2 = pytest.importorskip("cv2")
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EPS = 0.02  # stitcher_pipeline_test.cc:690


def scaled_K(K, s):
    k = np.array(K, np.float64)
    k[0, 0] *= s
    k[1, 1] *= s
    k[0, 2] *= s
    k[1, 2] *= s
    return k.astype(np.float32)


@pytest.fixture(scope="module")
def fx():
    g = np.load(os.path.join(GOLD, "real_photos.npz"))
    imgs = []
    for i in range(3):
        im = cv2.imdecode(g[f"jpeg_{i}"], cv2.IMREAD_COLOR)
        h, w = im.shape[:2]
        s = 1024 / max(h, w)
        imgs.append(cv2.resize(im, (int(round(w * s)), int(round(h * s))),
                               interpolation=cv2.INTER_AREA))
    same = all(hashlib.sha1(im.tobytes()).hexdigest() == str(g["preview_sha1"][i])
               for i, im in enumerate(imgs))
    work, seam = float(g["work_scale"]), float(g["seam_scale"])
    out = dict(g=g, imgs=imgs, same_decode=same, work_scale=work,
               seam_scale=seam, wis=float(g["warped_image_scale"]),
               K=[g["K"][i] for i in range(3)], R=[g["R"][i] for i in range(3)])
    out["seam_imgs"] = [cv2.resize(im, None, fx=seam, fy=seam,
                                   interpolation=cv2.INTER_LINEAR_EXACT)
                        for im in imgs]
    out["gains"] = [g[f"gain_{i}"] for i in range(3)]
    out["seams"] = [
        (np.unpackbits(g[f"seam_{i}"])[:int(np.prod(g[f"seam_shape_{i}"]))]
         .reshape(g[f"seam_shape_{i}"]) * 255).astype(np.uint8)
        for i in range(3)]
    return out


def compose_cpu(fx, gains, seams):
    cwa = 1.0 / fx["work_scale"]
    Ks = [scaled_K(k, cwa) for k in fx["K"]]
    return co.compose(fx["imgs"], Ks, fx["R"], projection=wo.SPHERICAL,
                      warped_image_scale=fx["wis"], compose_work_aspect=cwa,
                      gains=gains, seam_masks=seams)


def test_oracle_chain_on_the_reference_photos(fx):
    """Panorama size as the reference's test expects; gain maps of the
    exposure oracle on the warp oracle's seam-scale images as cv2's."""
    swa = fx["seam_scale"] / fx["work_scale"]
    warper = wo.Warper(wo.SPHERICAL, float(np.float32(fx["wis"] * swa)))
    corners, warped, masks = [], [], []
    for im, K, R in zip(fx["seam_imgs"], fx["K"], fx["R"]):
        k = scaled_K(K, swa)
        c, iw = warper.warp(im, k, R)
        _, mw = warper.warp_mask((im.shape[1], im.shape[0]), k, R)
        corners.append(c)
        warped.append(iw)
        masks.append(mw)
    assert np.array_equal(np.array(corners), fx["g"]["seam_corners"])
    maps, _ = go.blocks_gain_feed(corners, warped, masks)
    if fx["same_decode"]:
        for a, b in zip(maps, fx["gains"]):
            np.testing.assert_allclose(a, b, rtol=2e-6)
    res = compose_cpu(fx, fx["gains"], fx["seams"])
    rows, cols = res["pano"].shape[:2]
    want_rows, want_cols = fx["g"]["expected_rows_cols"]
    assert abs(rows - want_rows) <= EPS * want_rows
    assert abs(cols - want_cols) <= EPS * want_cols
    assert res["mask"].mean() > 0.8 * 255  # a well-filled panorama


@pytest.mark.gpu
def test_device_path_on_the_reference_photos(fx):
    import xpano_b200 as xb
    finder = cv2.detail_GraphCutSeamFinder("COST_COLOR")

    def find(images_f, corners, masks):
        out = finder.find(images_f, corners, masks)
        return [m.get() if hasattr(m, "get") else np.asarray(m) for m in out]

    swa = fx["seam_scale"] / fx["work_scale"]
    est = xb.EstimateSeams(fx["seam_imgs"], fx["K"], fx["R"],
                           xb.ProjectionType.kSpherical, fx["wis"], swa,
                           seam_finder=find)
    assert np.array_equal(np.array(est["corners"]), fx["g"]["seam_corners"])
    if fx["same_decode"]:
        for a, b in zip(est["gains"], fx["gains"]):
            np.testing.assert_allclose(a, b, rtol=2e-6)
        for a, b in zip(est["masks_warped"], fx["seams"]):
            assert ((a != 0) != (b != 0)).mean() < 1e-3

    cwa = 1.0 / fx["work_scale"]
    Ks = [scaled_K(k, cwa) for k in fx["K"]]
    c = xb.Composer(projection=wo.SPHERICAL, warped_image_scale=fx["wis"],
                    compose_work_aspect=cwa, depth=8)
    c.set_images(fx["imgs"], Ks, fx["R"], est["gains"], est["masks_warped"])
    info = c.run()
    pano, mask = c.fetch()
    want_rows, want_cols = fx["g"]["expected_rows_cols"]
    assert abs(info["height"] - want_rows) <= EPS * want_rows
    assert abs(info["width"] - want_cols) <= EPS * want_cols

    want = compose_cpu(fx, est["gains"], est["masks_warped"])
    assert pano.shape == want["pano"].shape
    assert (info["min_x"], info["min_y"]) == (want["min_x"], want["min_y"])
    assert np.array_equal(mask, want["mask"])
    both = mask > 0
    d = np.abs(pano.astype(int) - want["pano"].astype(int))[both]
    mse = float((d.astype(np.float64) ** 2).mean())
    psnr = 99.0 if mse == 0 else 10 * math.log10(255.0 ** 2 / mse)
    assert (d <= 1).mean() >= 0.999, ((d <= 1).mean(), int(d.max()))
    assert psnr >= 50.0, psnr
    # the auto-crop the export applies next (stitcher_pipeline.cc:374-403)
    crop = c.auto_crop()
    assert crop is not None
    x0, y0, x1, y1 = crop
    assert mask[y0:y1, x0:x1].all()

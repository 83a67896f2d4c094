"""Edge cases of the compose path through the C ABI against the compose
oracle (tests/compose_oracle.py): tiny images, an image the seam masks hide
completely (skipped like stitcher.cc:360-364), images that do not overlap
(a hole in the canvas), two images on top of each other, a single column of
overlap, 16-bit samples with the levels left to the library.

Bars as everywhere: masks bit-exact; colours within 1 LSB (8-bit) / 1e-3
relative (16-bit) at >= 99.9 % of the pixels."""
import math
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import compose_oracle as co  # noqa: E402
import synth  # noqa: E402
from oracle import warp_oracle as wo  # noqa: E402

pytestmark = pytest.mark.gpu


def run_both(ims, Ks, Rs, f, bpp=8, seams=None, gains=None, **kw):
    import xpano_b200 as xb
    want = co.compose(ims, Ks, Rs, projection=wo.SPHERICAL, warp_scale=f,
                      gains=gains, seam_masks=seams, **kw)
    c = xb.Composer(projection=wo.SPHERICAL, warp_scale=f, depth=bpp, **kw)
    c.set_images(ims, Ks, Rs, gains, seams)
    info = c.run()
    pano, mask = c.fetch()
    assert pano.shape == want["pano"].shape
    assert (info["min_x"], info["min_y"]) == (want["min_x"], want["min_y"])
    assert info["levels"] == want["levels"]
    assert np.array_equal(mask, want["mask"])
    both = mask > 0
    d = np.abs(pano.astype(int) - want["pano"].astype(int))[both]
    tol = 1 if bpp == 8 else np.maximum(
        1, 1e-3 * want["pano"].astype(int)[both])
    assert d.size == 0 or (d <= tol).mean() >= 0.999, \
        (int(d.max()), float((d <= tol).mean()))
    return pano, mask, info


def cams(n, w, h, f, step, seed=3):
    hfov = 2 * math.atan(w / 2 / f)
    return co.make_cameras(n, 1, w, h, f, step * hfov, [0.0], seed, None)


def test_tiny_images():
    w, h, f = 24, 17, 30.0
    Ks, Rs = cams(2, w, h, f, 0.5)
    ims = [synth.texture(h, w, 40 + i, 3, 8) for i in range(2)]
    run_both(ims, Ks, Rs, f)


def test_image_hidden_by_its_seam_mask_is_skipped():
    w, h, f = 200, 150, 160.0
    Ks, Rs = cams(3, w, h, f, 0.45)
    ims = [synth.texture(h, w, 50 + i, 3, 8) for i in range(3)]
    seams = [np.full((20, 26), 255, np.uint8) for _ in range(3)]
    seams[1][:] = 0           # fully obscured: stitcher.cc:360-364
    pano, mask, _ = run_both(ims, Ks, Rs, f, seams=seams, fixed_levels=3)
    assert mask.any()


def test_disjoint_images_leave_a_hole():
    w, h, f = 160, 120, 200.0
    Ks, Rs = cams(2, w, h, f, 1.6)     # 60 % of a field of view apart
    ims = [synth.texture(h, w, 60 + i, 3, 8) for i in range(2)]
    pano, mask, _ = run_both(ims, Ks, Rs, f, fixed_levels=4)
    cols = mask.any(axis=0)
    assert not cols.all()               # a gap of empty columns in between


def test_two_images_on_top_of_each_other():
    w, h, f = 180, 130, 170.0
    Ks, Rs = cams(1, w, h, f, 0.0)
    Ks, Rs = [Ks[0], Ks[0]], [Rs[0], Rs[0]]
    ims = [synth.texture(h, w, 70 + i, 3, 8) for i in range(2)]
    run_both(ims, Ks, Rs, f)


def test_thin_overlap_and_gain_maps():
    w, h, f = 220, 160, 200.0
    Ks, Rs = cams(3, w, h, f, 0.97)    # a few columns shared
    ims = [synth.texture(h, w, 80 + i, 3, 8) for i in range(3)]
    rng = np.random.default_rng(5)
    gains = [rng.uniform(0.8, 1.25, (6, 8)).astype(np.float32)
             for _ in range(3)]
    run_both(ims, Ks, Rs, f, gains=gains)


def test_sixteen_bit_levels_left_to_the_library():
    w, h, f = 210, 140, 190.0
    Ks, Rs = cams(3, w, h, f, 0.6)
    ims = [synth.texture(h, w, 90 + i, 3, 16) for i in range(3)]
    pano, _, info = run_both(ims, Ks, Rs, f, bpp=16)
    assert pano.dtype == np.uint16 and info["levels"] >= 5


@pytest.mark.parametrize("n_cols,reverse", [(3, False), (5, True)])
def test_cover_map_from_opacity_bits_equals_the_pixel_path(n_cols, reverse,
                                                           monkeypatch):
    """The composer's warps leave an opacity bit plane; the cover map built
    from it (cover_bits_kernel) is the one built from the pixels
    (cover_kernel), word for word, and so is everything after it."""
    import xpano_b200 as xb
    w, h, f = 333, 190, 210.0
    Ks, Rs = cams(n_cols, w, h, f, 0.55, seed=8)
    ims = [synth.texture(h, w, 120 + i, 3, 8) for i in range(n_cols)]
    rng = np.random.default_rng(2)
    seams = []
    for i in range(n_cols):   # ragged seam masks: holes and run starts
        m = (rng.random((24, 40)) > 0.25).astype(np.uint8) * 255
        m[:, 10:30] = 255
        seams.append(m)
    got = {}
    for mode in ("bits", "pixels"):
        if mode == "pixels":
            monkeypatch.setenv("XP_NO_COVER_BITS", "1")
        c = xb.Composer(projection=wo.SPHERICAL, warp_scale=f, depth=8,
                        fixed_levels=3, reverse=reverse, keep_debug=True)
        c.set_images(ims, Ks, Rs, None, seams)
        c.run()
        pano, mask = c.fetch()
        got[mode] = (c.debug("cover", dtype=np.uint32),
                     c.debug("labels", dtype=np.int32), pano, mask)
    for a, b in zip(got["bits"], got["pixels"]):
        assert a.shape == b.shape and np.array_equal(a, b)
    assert got["bits"][0].any()


def test_auto_fill_inputs_from_the_device_mask():
    """RunInpainting's inpaint_mask and pixels_filled
    (stitcher_pipeline.cc:384-389) from the mask on the device."""
    import xpano_b200 as xb
    w, h, f = 160, 120, 200.0
    Ks, Rs = cams(2, w, h, f, 1.6)
    ims = [synth.texture(h, w, 60 + i, 3, 8) for i in range(2)]
    c = xb.Composer(projection=wo.SPHERICAL, warp_scale=f, depth=8,
                    fixed_levels=4)
    c.set_images(ims, Ks, Rs)
    c.run()
    _, mask = c.fetch()
    inv, n = c.inpaint_mask()
    assert np.array_equal(inv, np.bitwise_not(mask))
    assert n == int(np.count_nonzero(inv)) and n > 0

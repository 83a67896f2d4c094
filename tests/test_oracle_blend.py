"""Pins the blend oracle (oracle/mb_oracle.c, a CPU restatement of
external/multiblend) against

 * the reference's own golden vectors (mb/tests/data/expected*.{png,tif},
   expected_seams.png, mb/tests/multiblend_test.cpp:39-44), decoded into
   tests/golden/blend_reference_goldens.npz by tests/golden/make_golden.py;
 * outputs of the unmodified reference library recorded on seeded synthetic
   cases (tests/golden/blend_ref_outputs.npz);
 * the reference library itself, live, when oracle/_ref/libmb_ref.so exists.

Integer/byte results: the bar is bit-exact.
"""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import synth  # noqa: E402
from golden.make_golden import REF_CASES  # noqa: E402
from oracle import mb_oracle, mb_ref  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "blend_reference_goldens.npz"))


@pytest.fixture(scope="module")
def ref_out():
    return np.load(os.path.join(GOLD, "blend_ref_outputs.npz"))


def _inside_alpha_equal(out_bgr, expected_bgra, maxv):
    a = expected_bgra[..., 3] == maxv
    return np.array_equal(out_bgr[a], expected_bgra[..., :3][a]), a


def test_expected_png_and_kat(gold):
    # raw fixtures are RGB, expected.png decoded by cv2 is BGRA
    ims = [gold[f"raw_{i}"] for i in range(3)]
    pos = [tuple(p) for p in gold["pos_png"]]
    o = mb_oracle.blend(ims, pos, output_bpp=8)
    w, h, mx, my = gold["kat_size_minpos"]
    assert (o.width, o.height, o.min_x, o.min_y) == (w, h, mx, my)
    assert o.no_mask is False
    same, a = _inside_alpha_equal(o.interleaved(False)[..., ::-1],
                                  gold["expected_png"], 255)
    assert same
    assert np.array_equal(o.mask == 255, a)


def test_expected_seams_png(gold):
    ims = [gold[f"raw_{i}"] for i in range(3)]
    pos = [tuple(p) for p in gold["pos_png"]]
    o = mb_oracle.blend(ims, pos, output_bpp=8)
    assert np.array_equal(o.labels, gold["expected_seams_png"])


def test_expected_alpha_png(gold):
    ims = [gold[f"alpha_{i}"] for i in range(3)]
    pos = [tuple(p) for p in gold["pos_png"]]
    o = mb_oracle.blend(ims, pos)
    exp = gold["expected_alpha_png"]
    assert (o.height, o.width) == exp.shape[:2]
    same, a = _inside_alpha_equal(o.interleaved(False), exp, 255)
    assert same
    assert np.array_equal(o.mask == 255, a)


def test_expected_16bit_tif(gold):
    ims = [gold[f"tif16_{i}"] for i in range(3)]
    pos = [tuple(p) for p in gold["pos_tif"]]
    o = mb_oracle.blend(ims, pos)
    exp = gold["expected_16bit_tif"]
    assert o.planes.dtype == np.uint16
    same, a = _inside_alpha_equal(o.interleaved(False), exp, 65535)
    assert same
    assert np.array_equal(o.mask == 255, a)


@pytest.mark.parametrize("case", REF_CASES, ids=[str(c[0]) for c in REF_CASES])
def test_recorded_reference_outputs(ref_out, case):
    seed, ckw, bkw = case
    ims, pos = synth.blend_case(seed, **ckw)
    o = mb_oracle.blend(ims, pos, **bkw)
    assert np.array_equal(o.planes, ref_out[f"planes_{seed}"])
    assert np.array_equal(o.mask, ref_out[f"mask_{seed}"])
    mx, my, nm = ref_out[f"meta_{seed}"]
    assert (o.min_x, o.min_y, int(o.no_mask)) == (mx, my, nm)


@pytest.mark.skipif(not mb_ref.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed", range(20, 32))
def test_live_reference(seed):
    mode = ("ellipse", "full", "blobs")[seed % 3]
    bpp = (8, 16)[seed % 2]
    ims, pos = synth.blend_case(seed, n=2 + seed % 4, mode=mode, bpp=bpp,
                                w=150 + 7 * (seed % 9), h=120 + 5 * (seed % 7))
    kw = dict(fixed_levels=(0, 3, 5, 6)[seed % 4], dither=bool(seed % 3))
    try:
        r = mb_ref.blend(ims, pos, n_threads=3, **kw)
    except RuntimeError as e:  # reference limitation ("bad p") on some masks
        pytest.skip(str(e))
    o = mb_oracle.blend(ims, pos, **kw)
    assert np.array_equal(o.planes, r.planes)
    assert np.array_equal(o.mask, r.mask)
    assert (o.min_x, o.min_y, o.no_mask) == (r.min_x, r.min_y, r.no_mask)


@pytest.mark.skipif(not mb_ref.available(), reason="oracle/_ref not built")
def test_reference_thread_count_invariance():
    ims, pos = synth.blend_case(3, n=4, mode="ellipse", w=260, h=200)
    a = mb_ref.blend(ims, pos, n_threads=2)
    b = mb_ref.blend(ims, pos, n_threads=8)
    assert np.array_equal(a.planes, b.planes)


def test_stage_entry_points():
    ims, pos = synth.blend_case(11, n=3, mode="ellipse")
    bb = mb_oracle.trim(ims[0])
    a = ims[0][..., 3] == 255
    ys, xs = np.nonzero(a)
    assert bb == (xs.min(), ys.min(), xs.max() - xs.min() + 1,
                  ys.max() - ys.min() + 1)
    filled, opaque = mb_oracle.inpaint(ims[0], bb)
    # opaque pixels untouched, every pixel of the bbox got an opaque colour
    assert np.array_equal(filled[a], ims[0][a])
    l, t, w, h = bb
    assert (filled[t:t + h, l:l + w, 3] == 255).all()
    assert np.array_equal(opaque.astype(bool), a[t:t + h, l:l + w])
    geo = mb_oracle.pyramid_geometry(200, 150, 4, 13, 7)
    assert [g["w"] for g in geo][0] == 200 and len(geo) == 4
    assert mb_oracle.blend_levels([198, 198, 205], [233, 231, 234]) == 6

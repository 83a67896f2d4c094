"""GPU parity tests of the blend half (trim, inpaint, seams, mask and image
pyramids, composite, collapse, out) through the C ABI, stage by stage against
the oracle (oracle/mb_oracle.c, pinned bit-exact against the reference
library and its golden images in tests/test_oracle_blend.py).

Bars: integer / index stages (trim, inpaint, seams, labels, masks, channel
sums) bit-exact; float pyramids bit-exact given identical inputs (the library
is built without FMA contraction); mask weights within 1 ulp of the oracle's
run-length evaluation; final 8/16-bit planes identical except where a 1-ulp
weight difference crosses a rounding boundary (<= 1 LSB, < 1e-4 of pixels).
"""
import json
import os
import sys
import tempfile

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import synth  # noqa: E402
from golden.make_golden import REF_CASES  # noqa: E402
from oracle import mb_oracle  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REPORT = {}


def _note(key, **kw):
    REPORT[key] = kw
    out = os.path.join(os.path.dirname(GOLD), "..", "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, "parity_blend.json"), "w") as f:
            json.dump(REPORT, f, indent=1, default=str)
    except OSError:
        pass


def _diff(name, got, want, exact=True, tol=0):
    got, want = np.asarray(got), np.asarray(want)
    assert got.shape == want.shape, f"{name}: shape {got.shape} vs {want.shape}"
    if got.dtype.kind == "f":
        bad = ~((got == want) | (np.isnan(got) & np.isnan(want)))
    else:
        bad = got != want
    nbad = int(bad.sum())
    msg = ""
    if nbad:
        idx = tuple(int(v[0]) for v in np.nonzero(bad))
        d = np.abs(got.astype(np.float64) - want.astype(np.float64))
        msg = (f"{name}: {nbad}/{bad.size} differ, first at {idx}: got "
               f"{got[idx]!r} want {want[idx]!r}, max abs diff {d.max():.6g}")
        print(msg)
    if exact:
        assert nbad == 0, msg
    return nbad


def _oracle_dump(ims, pos, **kw):
    """Run the oracle with MBO_DUMP and load its per-level dumps."""
    d = tempfile.mkdtemp(prefix="mbo_")
    os.environ["MBO_DUMP"] = d
    try:
        o = mb_oracle.blend(ims, pos, debug=True, **kw)
    finally:
        del os.environ["MBO_DUMP"]
    dumps = {}
    for fn in os.listdir(d):
        raw = np.fromfile(os.path.join(d, fn), np.uint8)
        w, h, pitch = raw[:12].view(np.int32)
        dumps[fn[3:-4]] = raw[12:].view(np.float32).reshape(h, pitch)[:, :w]
        os.remove(os.path.join(d, fn))
    os.rmdir(d)
    return o, dumps


def _gpu_blend(ims, pos, debug=True, **kw):
    import xpano_b200
    b = xpano_b200.Multiblend(keep_debug=debug, **kw)
    for im, p in zip(ims, pos):
        b.feed_interleaved(im, p)
    b.run()
    return b


CASES = [
    (1, dict(n=2, mode="ellipse", bpp=8), dict()),
    (2, dict(n=3, mode="full", bpp=8), dict(fixed_levels=3)),
    (3, dict(n=5, mode="blobs", bpp=8), dict(fixed_levels=5)),
    (4, dict(n=4, mode="ellipse", bpp=16), dict()),
    (6, dict(n=6, mode="ellipse", bpp=8, w=320, h=240), dict(fixed_levels=7)),
    (8, dict(n=4, mode="ellipse", bpp=8), dict(reverse=True, fixed_levels=2)),
    (10, dict(n=3, mode="ellipse", bpp=8), dict(fixed_levels=1)),
    (21, dict(n=3, mode="ellipse", bpp=8, w=700, h=500), dict(fixed_levels=6)),
]


@pytest.mark.parametrize("case", CASES, ids=[str(c[0]) for c in CASES])
def test_stages_against_oracle(case):
    seed, ckw, bkw = case
    ims, pos = synth.blend_case(seed, **ckw)
    o, dumps = _oracle_dump(ims, pos, **bkw)
    b = _gpu_blend(ims, pos, **bkw)
    n = len(ims)
    sdt = ims[0].dtype
    assert b.size == (o.width, o.height)
    assert b.min_pos == (o.min_x, o.min_y)
    assert b.no_mask == o.no_mask
    # trim + inpaint (image.cpp:396-868)
    for i in range(n):
        t = o.images[i]
        _diff(f"trim[{i}]", b.debug("trim", i),
              np.array([t.left, t.top, t.width, t.height, t.xpos, t.ypos]))
        _diff(f"opaque[{i}]", b.debug("opaque", i), t.opaque)
        for c in range(3):
            _diff(f"chan[{i}][{c}]", b.debug("chan", i, c, dtype=sdt),
                  t.chan[c])
    # seams (multiblend.cpp:254-823)
    _diff("seam_bwd", b.debug("seam_bwd"), o.seam_bwd)
    _diff("seam_fwd", b.debug("seam_fwd"), o.seam_fwd)
    _diff("labels", b.debug("labels"), o.labels)
    _diff("xor_mask", b.debug("xor_mask"), o.xor_mask)
    _diff("full_mask", b.debug("full_mask"), o.mask)
    _diff("totals", b.debug("totals"),
          np.array(list(o.channel_totals) + [o.total_pixels], np.uint64))
    # mask pyramids (functions.cpp:24-314): dense vs run-length evaluation
    levels = o.levels
    mask_bad = 0
    for i in range(n):
        for l in range(1, levels):
            want = dumps[f"mask_i{i}_l{l}"]
            x0, y0, w, h, cw, ch = b.debug("mask_rect", i, 0, l)
            assert (ch, cw) == want.shape
            full = np.zeros_like(want)
            if w and h:
                full[y0:y0 + h, x0:x0 + w] = b.debug("mask", i, 0, l)
            nb = _diff(f"mask[{i}][{l}]", full, want, exact=False)
            if nb:
                ulp = np.abs(full.view(np.int32).astype(np.int64)
                             - want.view(np.int32).astype(np.int64))
                assert ulp.max() <= 1, f"mask[{i}][{l}] off by {ulp.max()} ulp"
            mask_bad += nb
    # Laplacian pyramids of the inputs (pyramid.cpp:233-1038)
    for i in range(n):
        for c in range(3):
            for l in range(levels):
                _diff(f"lap[{i}][{c}][{l}]", b.debug("lap", i, c, l),
                      dumps[f"lap_c{c}_i{i}_l{l}"])
    # composited output pyramid, collapsed level 0, planes
    out_bad = 0
    for c in range(3):
        for l in range(levels):
            out_bad += _diff(f"out[{c}][{l}]", b.debug("out", 0, c, l),
                             dumps[f"out_c{c}_i0_l{l}"], exact=mask_bad == 0)
    planes = b.planes()
    d = np.abs(planes.astype(np.int64) - o.planes.astype(np.int64))
    frac = float((d > 0).mean())
    _note(f"case{seed}", mask_values_off_by_1ulp=mask_bad,
          out_pyramid_values_differing=out_bad, plane_max_diff=int(d.max()),
          plane_diff_fraction=frac)
    if mask_bad == 0:
        for c in range(3):
            _diff(f"collapsed[{c}]", b.debug("collapsed", 0, c), o.collapsed[c],
                  exact=False)
    assert d.max() <= 1 and frac < 1e-4, (d.max(), frac)
    dst, mask = b.blend()
    _diff("mask", mask, o.mask)
    _diff("dst", dst, np.where(o.mask[..., None] > 0,
                               np.moveaxis(planes, 0, 2), 0))


def _sparse_case(bpp=8):
    """Images far apart on a wide canvas, one of them with a mostly
    transparent bounding box: most 32x32 skewed tiles of the canvas are
    covered by no image (the wavefront's uncovered step variant, first band
    included) and run-start bits fall into such tiles."""
    rng = np.random.default_rng(5)
    dt = np.uint8 if bpp == 8 else np.uint16
    amax = 255 if bpp == 8 else 65535

    def img(h, w, opaque):
        a = np.zeros((h, w, 4), dt)
        a[..., :3] = rng.integers(0, amax + 1, (h, w, 3), dtype=np.uint32)
        a[..., 3] = np.where(opaque, amax, 0)
        return a

    W, H = 900, 400
    frame = np.zeros((H - 10, W - 20), bool)
    frame[:14, :17] = True
    frame[-11:, -15:] = True
    frame[150:170, 420:450] = True
    blob = rng.random((160, 220)) < 0.9
    blob[0, :] = blob[-1, :] = blob[:, 0] = blob[:, -1] = True
    ims = [img(*frame.shape, frame), img(160, 220, blob),
           img(100, 150, np.ones((100, 150), bool)),
           img(150, 200, np.ones((150, 200), bool))]
    pos = [(0, 0), (W - 220, H - 160), (330, 120), (5, 200)]
    return ims, pos


@pytest.mark.parametrize("bpp,reverse", [(8, False), (16, True)])
def test_sparse_canvas_uncovered_tiles(bpp, reverse):
    ims, pos = _sparse_case(bpp)
    kw = dict(fixed_levels=4, reverse=reverse)
    o, _ = _oracle_dump(ims, pos, **kw)
    b = _gpu_blend(ims, pos, **kw)
    assert b.size == (o.width, o.height)
    _diff("seam_bwd", b.debug("seam_bwd"), o.seam_bwd)
    _diff("seam_fwd", b.debug("seam_fwd"), o.seam_fwd)
    _diff("labels", b.debug("labels"), o.labels)
    _diff("full_mask", b.debug("full_mask"), o.mask)
    d = np.abs(b.planes().astype(np.int64) - o.planes.astype(np.int64))
    assert d.max() <= 1 and float((d > 0).mean()) < 1e-4
    # the streaming path (what xpano's export runs) gives the same planes
    s2 = _gpu_blend(ims, pos, debug=False, **kw)
    _diff("planes", s2.planes(), b.planes())
    _note(f"sparse_{bpp}", plane_max_diff=int(d.max()),
          uncovered_fraction=float((o.mask == 0).mean()))


def test_single_image_and_spp3():
    ims, pos = synth.blend_case(9, n=1, mode="ellipse", bpp=8)
    o = mb_oracle.blend(ims, pos)
    b = _gpu_blend(ims, pos, debug=False)
    _diff("planes n=1", b.planes(), o.planes)
    ims, pos = synth.blend_case(7, n=3, mode="full", bpp=8, spp=3)
    o = mb_oracle.blend(ims, pos)
    b = _gpu_blend(ims, pos, debug=False)
    d = np.abs(b.planes().astype(int) - o.planes.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 1e-4
    assert b.no_mask == o.no_mask


def test_reference_golden_images():
    """The reference's own goldens (mb/tests/data/expected*.png/.tif,
    multiblend_test.cpp:39-44) through the GPU path."""
    g = np.load(os.path.join(GOLD, "blend_reference_goldens.npz"))
    ims = [g[f"raw_{i}"] for i in range(3)]
    pos = [tuple(p) for p in g["pos_png"]]
    b = _gpu_blend(ims, pos, debug=True, output_bpp=8)
    w, h, mx, my = g["kat_size_minpos"]
    assert b.size == (w, h) and b.min_pos == (mx, my) and b.no_mask is False
    _diff("expected_seams.png", b.debug("labels"), g["expected_seams_png"])
    exp = g["expected_png"]
    a = exp[..., 3] == 255
    dst, mask = b.blend()
    _diff("mask vs expected.png alpha", mask == 255, a)
    d = np.abs(dst[..., ::-1][a].astype(int) - exp[..., :3][a].astype(int))
    _note("expected.png", max_diff=int(d.max()), frac=float((d > 0).mean()))
    assert d.max() <= 1 and (d > 0).mean() < 1e-4
    # alpha (trim + inpaint) and 16-bit goldens
    for key, posk, maxv in (("alpha", "pos_png", 255), ("tif16", "pos_tif",
                                                        65535)):
        ims = [g[f"{key}_{i}"] for i in range(3)]
        pos = [tuple(p) for p in g[posk]]
        b = _gpu_blend(ims, pos, debug=False)
        exp = g["expected_alpha_png" if key == "alpha" else
                "expected_16bit_tif"]
        dst, mask = b.blend()
        a = exp[..., 3] == maxv
        _diff(f"{key} mask", mask == 255, a)
        d = np.abs(dst[a].astype(int) - exp[..., :3][a].astype(int))
        _note(f"expected_{key}", max_diff=int(d.max()),
              frac=float((d > 0).mean()))
        assert d.max() <= 1 and (d > 0).mean() < 1e-4


@pytest.mark.parametrize("case", REF_CASES, ids=[str(c[0]) for c in REF_CASES])
def test_recorded_reference_outputs(case):
    """Outputs of the unmodified reference library recorded on seeded inputs
    (tests/golden/blend_ref_outputs.npz)."""
    seed, ckw, bkw = case
    ref = np.load(os.path.join(GOLD, "blend_ref_outputs.npz"))
    ims, pos = synth.blend_case(seed, **ckw)
    b = _gpu_blend(ims, pos, debug=False, **bkw)
    planes = b.planes()
    _, mask = b.blend()
    _diff("mask", mask, ref[f"mask_{seed}"])
    mx, my, nm = ref[f"meta_{seed}"]
    assert b.min_pos == (mx, my) and int(b.no_mask) == nm
    d = np.abs(planes.astype(int) - ref[f"planes_{seed}"].astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 1e-4, (d.max(), (d > 0).mean())


@pytest.mark.parametrize("case", CASES, ids=[str(c[0]) for c in CASES])
def test_streaming_path_equals_staged_and_oracle(case):
    """Without keep_debug the blender never materialises level 0 or the
    Laplacian (pyramid_fused.cu); the result must be bit-identical to the
    staged kernels and to the oracle's planes."""
    seed, ckw, bkw = case
    ims, pos = synth.blend_case(seed, **ckw)
    staged = _gpu_blend(ims, pos, debug=True, **bkw)
    stream = _gpu_blend(ims, pos, debug=False, **bkw)
    ps, pf = staged.planes(), stream.planes()
    assert np.array_equal(ps, pf), int(np.abs(ps.astype(int) - pf).max())
    o = mb_oracle.blend(ims, pos, **bkw)
    assert np.array_equal(pf, o.planes)
    _, ms = staged.blend()
    _, mf = stream.blend()
    assert np.array_equal(ms, mf)


@pytest.mark.parametrize("n", [34, 28])
def test_many_images_wide_cover(n):
    """More than 31 images switches the cover map to 64-bit words; 25..31 use
    32-bit bit words instead of the digest form (scan_problems.h Cov<>)."""
    ims, pos = synth.blend_case(33, n=n, w=60, h=50, mode="ellipse")
    o = mb_oracle.blend(ims, pos, fixed_levels=3, debug=True)
    b = _gpu_blend(ims, pos, debug=True, fixed_levels=3)
    _diff("labels", b.debug("labels"), o.labels)
    d = np.abs(b.planes().astype(int) - o.planes.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 1e-4


def test_error_convention():
    import xpano_b200
    b = xpano_b200.Multiblend()
    with pytest.raises(xpano_b200._lib.XpError):
        b.run()  # nothing fed
    b = xpano_b200.Multiblend()
    b.feed_interleaved(np.zeros((8, 8, 4), np.uint8), (0, 0))
    with pytest.raises(xpano_b200._lib.XpError) as e:
        b.run()
    assert "no opaque pixel" in str(e.value)

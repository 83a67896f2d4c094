"""GPU parity tests of the exposure step of Stitcher::EstimateSeams
(stitcher.cc:241-296) through the C ABI
(xp_gain_compensator_*), against vectors recorded from cv2 4.13.0
(tests/golden/gain_cv2_goldens.npz, seam_cv2_goldens.npz) and against the
exposure oracle (oracle/gain_oracle.py).

Bars: block-pair counts exact and sums 1e-12 relative (double sums in a
different order); gain maps 2e-6 relative against cv2 (they are float32 from
a double solve); apply: <= 1 LSB on fewer than 1e-3 of the samples (the float
gain resize is 1e-5 from cv2's, test_oracle_warp); seam-scale warps and
corners bit-exact."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from golden import make_golden as mg  # noqa: E402
from oracle import gain_oracle as go  # noqa: E402
from oracle import warp_oracle as wo  # noqa: E402
from test_oracle_gain import SCENES, load_scene  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", sorted(SCENES))
@pytest.mark.parametrize("iters", [0, 2])
def test_feed_matches_cv2(name, iters):
    from xpano_b200.exposure import BlocksGainCompensator
    g, cs, ims, ms = load_scene(name)
    comp = BlocksGainCompensator(32, 32, iters)
    comp.feed(cs, ims, ms)
    maps = comp.getMatGains()
    assert len(maps) == len(ims)
    for i, m in enumerate(maps):
        ref = g[f"{name}_gain{iters}_{i}"]
        assert m.shape == ref.shape
        np.testing.assert_allclose(m, ref, rtol=2e-6)
    want, _ = go.blocks_gain_feed(cs, ims, ms, iters=iters)
    for a, b in zip(maps, want):
        np.testing.assert_allclose(a, b, rtol=2e-6)


def test_feed_strided_inputs_and_repeat():
    """Row strides (views into larger arrays) and a second feed on the same
    object give the same maps."""
    from xpano_b200.exposure import BlocksGainCompensator
    g, cs, ims, ms = load_scene("A")
    comp = BlocksGainCompensator()
    comp.feed(cs, ims, ms)
    first = comp.getMatGains()
    big_i = [np.zeros((im.shape[0] + 3, im.shape[1] + 7, 3), np.uint8)
             for im in ims]
    big_m = [np.zeros((m.shape[0] + 2, m.shape[1] + 5), np.uint8) for m in ms]
    vi, vm = [], []
    for im, m, bi, bm in zip(ims, ms, big_i, big_m):
        bi[1:1 + im.shape[0], 2:2 + im.shape[1]] = im
        bm[1:1 + m.shape[0], 3:3 + m.shape[1]] = m
        vi.append(bi[1:1 + im.shape[0], 2:2 + im.shape[1]])
        vm.append(bm[1:1 + m.shape[0], 3:3 + m.shape[1]])
    comp.feed(cs, vi, vm)
    for a, b in zip(first, comp.getMatGains()):
        assert np.array_equal(a, b)


def test_feed_device_pointers():
    import ctypes
    import torch
    from xpano_b200 import _lib
    g, cs, ims, ms = load_scene("A")
    lib = _lib.load()
    h = ctypes.c_void_p()
    _lib.check(lib.xp_gain_compensator_create(32, 32, 2, 0, ctypes.byref(h)))
    try:
        ti = [torch.from_numpy(im).cuda() for im in ims]
        tm = [torch.from_numpy(m).cuda() for m in ms]
        n = len(ims)
        c = np.ascontiguousarray(cs, np.int32)
        w = np.array([im.shape[1] for im in ims], np.int32)
        hh = np.array([im.shape[0] for im in ims], np.int32)
        ip = (ctypes.c_void_p * n)(*[t.data_ptr() for t in ti])
        mp = (ctypes.c_void_p * n)(*[t.data_ptr() for t in tm])
        torch.cuda.synchronize()
        _lib.check(lib.xp_gain_compensator_feed(
            h, n, c.ctypes.data, ip, w.ctypes.data, hh.ctypes.data, None, mp,
            None, _lib.MEM_DEVICE))
        for i in range(n):
            wh = (ctypes.c_int * 2)()
            _lib.check(lib.xp_gain_compensator_map_size(h, i, wh))
            m = np.empty((wh[1], wh[0]), np.float32)
            _lib.check(lib.xp_gain_compensator_get_map(h, i, m.ctypes.data))
            np.testing.assert_allclose(m, g[f"A_gain2_{i}"], rtol=2e-6)
        # apply in place on a device image
        t = ti[0].clone()
        _lib.check(lib.xp_gain_compensator_apply(
            h, 0, t.data_ptr(), int(w[0]), int(hh[0]), 0, 8, _lib.MEM_DEVICE))
        torch.cuda.synchronize()
        d = np.abs(t.cpu().numpy().astype(int)
                   - g["A_applied_0"].astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3
    finally:
        lib.xp_gain_compensator_destroy(h)


@pytest.mark.parametrize("name", sorted(SCENES))
def test_apply_matches_cv2(name):
    from xpano_b200.exposure import BlocksGainCompensator
    g, cs, ims, ms = load_scene(name)
    comp = BlocksGainCompensator()
    comp.setMatGains([g[f"{name}_gain2_{i}"] for i in range(len(ims))])
    for i, im in enumerate(ims):
        got = comp.apply(i, cs[i], im, ms[i])
        ref = g[f"{name}_applied_{i}"]
        d = np.abs(got.astype(int) - ref.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3
        # and the same bits as the oracle chain (same float arithmetic)
        gm = g[f"{name}_gain2_{i}"]
        h, w = im.shape[:2]
        full = gm if gm.shape == (h, w) else wo.resize_linear_f32(gm, (w, h))
        assert np.array_equal(got, wo.apply_gain_u8(im, full))


def test_apply_16bit_and_errors():
    from xpano_b200 import _lib
    from xpano_b200.exposure import BlocksGainCompensator
    comp = BlocksGainCompensator()
    with pytest.raises(_lib.XpError):
        comp.apply(0, (0, 0), np.zeros((4, 4, 3), np.uint8))  # before feed
    gm = np.array([[0.5, 1.0], [1.5, 300.0]], np.float32)
    comp.setMatGains([gm])
    im = np.full((2, 2, 3), 1000, np.uint16)
    out = comp.apply(0, (0, 0), im)
    assert out[0, 0, 0] == 500 and out[0, 1, 1] == 1000
    assert out[1, 0, 2] == 1500 and out[1, 1, 0] == 65535  # saturates
    with pytest.raises(_lib.XpError):
        comp.apply(3, (0, 0), im)


def test_estimate_seams_matches_cv2():
    """EstimateSeams without the (CPU, caller-side) graph cut: corners and
    gain maps as cv2's; compensated seam-scale images within 1 LSB. With cv2
    importable, its GraphCutSeamFinder run on this package's compensated
    images gives the masks it gives on cv2's own."""
    import xpano_b200 as xb
    from xpano_b200.exposure import EstimateSeams
    g = np.load(os.path.join(GOLD, "seam_cv2_goldens.npz"))
    ims, Ks, Rs = mg.seam_scene()
    res = EstimateSeams(ims, Ks, Rs, xb.ProjectionType.kSpherical,
                        mg.SEAM_SCENE["f"], 1.0)
    assert np.array_equal(np.array(res["corners"], np.int32), g["corners"])
    changed = 0.0
    for i in range(len(ims)):
        np.testing.assert_allclose(res["gains"][i], g[f"gain_{i}"], rtol=2e-6)
        d = np.abs(res["images_warped"][i].astype(int)
                   - g[f"warped_{i}"].astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3
        changed = max(changed, float((d > 0).mean()))
    try:
        import cv2
    except Exception:  # pragma: no cover
        return
    finder = cv2.detail_GraphCutSeamFinder("COST_COLOR")

    def find(images_f, corners, masks):
        out = finder.find(images_f, corners, masks)
        return [m.get() if hasattr(m, "get") else np.asarray(m) for m in out]

    res = EstimateSeams(ims, Ks, Rs, xb.ProjectionType.kSpherical,
                        mg.SEAM_SCENE["f"], 1.0, seam_finder=find)
    for i in range(len(ims)):
        diff = float((res["masks_warped"][i] != g[f"seam_{i}"]).mean())
        # a rounding moved by the gain resize can move a cut by a pixel
        assert diff < (1e-3 if changed > 0 else 1e-12), (i, diff, changed)


@pytest.mark.parametrize("seed", range(8))
def test_feed_random_cases(seed):
    """Seeded random layouts, sizes, block sizes and ragged masks against the
    exposure oracle (pinned bit for bit against cv2)."""
    from xpano_b200.exposure import BlocksGainCompensator
    rng = np.random.default_rng(900 + seed)
    n = int(rng.integers(2, 7))
    bw, bh = int(rng.integers(8, 48)), int(rng.integers(8, 48))
    iters = int(rng.integers(0, 4))
    cs, ims, ms = [], [], []
    for _ in range(n):
        w, h = int(rng.integers(20, 180)), int(rng.integers(20, 140))
        ims.append(rng.integers(0, 256, (h, w, 3)).astype(np.uint8))
        m = np.full((h, w), 255, np.uint8)
        m[rng.random((h, w)) < 0.15] = 0
        if rng.random() < 0.3:
            m[:, : w // 2] = 0
        ms.append(m)
        cs.append((int(rng.integers(-90, 90)), int(rng.integers(-60, 60))))
    comp = BlocksGainCompensator(bw, bh, iters)
    comp.feed(cs, ims, ms)
    want, _ = go.blocks_gain_feed(cs, ims, ms, bw, bh, iters)
    for a, b in zip(comp.getMatGains(), want):
        assert a.shape == b.shape
        np.testing.assert_allclose(a, b, rtol=2e-6)

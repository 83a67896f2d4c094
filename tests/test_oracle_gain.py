"""Pins the exposure oracle (oracle/gain_oracle.py, a numpy restatement of
cv::detail::BlocksGainCompensator::feed as xpano runs it in
Stitcher::EstimateSeams, stitcher.cc:273-278) against vectors recorded from
cv2 4.13.0 (tests/golden/gain_cv2_goldens.npz) and, when cv2 is importable,
against cv2 live; then checks the host half of the product (block pairs,
sparse system, smoothing -- xp_gain_compensator_plan / _solve, no GPU) against
the oracle.

Bars: gain maps vs cv2: 2e-6 relative (float32 maps from a double solve; on
every recorded case the oracle's maps are the same bits as cv2's). Block
pairs and counts: exact."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle import gain_oracle as go  # noqa: E402
from oracle import warp_oracle as wo  # noqa: E402

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCENES = {"A": 5, "B": 3}


def load_scene(name):
    g = np.load(os.path.join(GOLD, "gain_cv2_goldens.npz"))
    n = SCENES[name]
    cs = [tuple(int(v) for v in c) for c in g[f"{name}_corners"]]
    ims = [g[f"{name}_img_{i}"] for i in range(n)]
    ms = [g[f"{name}_mask_{i}"] for i in range(n)]
    return g, cs, ims, ms


@pytest.mark.parametrize("name", sorted(SCENES))
@pytest.mark.parametrize("iters", [0, 2])
def test_oracle_matches_cv2_golden(name, iters):
    g, cs, ims, ms = load_scene(name)
    maps, _ = go.blocks_gain_feed(cs, ims, ms, iters=iters)
    for i, m in enumerate(maps):
        ref = g[f"{name}_gain{iters}_{i}"]
        assert m.shape == ref.shape
        np.testing.assert_allclose(m, ref, rtol=2e-6)


@pytest.mark.parametrize("name", sorted(SCENES))
def test_oracle_apply_matches_cv2_golden(name):
    """BlocksCompensator::apply = resize(gain, INTER_LINEAR) * image; the
    float resize is 1e-5 from cv2's (test_oracle_warp), which moves a rounding
    now and then: <= 1 LSB, fewer than 1e-3 of the samples."""
    g, cs, ims, ms = load_scene(name)
    for i, im in enumerate(ims):
        gm = g[f"{name}_gain2_{i}"]
        h, w = im.shape[:2]
        full = gm if gm.shape == (h, w) else wo.resize_linear_f32(gm, (w, h))
        got = wo.apply_gain_u8(im, full)
        ref = g[f"{name}_applied_{i}"]
        d = np.abs(got.astype(int) - ref.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_oracle_matches_cv2_live(seed):
    rng = np.random.default_rng(seed)
    n = 3 + seed
    cs, ims, ms = [], [], []
    for k in range(n):
        w, h = int(rng.integers(40, 150)), int(rng.integers(30, 120))
        im = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
        m = np.full((h, w), 255, np.uint8)
        m[rng.random((h, w)) < 0.2] = 0
        if k == 1:
            m[:, : w // 2] = 0
        cs.append((int(rng.integers(-60, 60)), int(rng.integers(-40, 40))))
        ims.append(im)
        ms.append(m)
    comp = cv2.detail_BlocksGainCompensator()
    comp.feed(cs, ims, ms)
    ref = comp.getMatGains()
    maps, _ = go.blocks_gain_feed(cs, ims, ms)
    for a, b in zip(maps, ref):
        np.testing.assert_allclose(a, b, rtol=2e-6)


def host_maps(cs, ims, ms, iters):
    """The product's host half fed with the oracle's per-pair statistics."""
    from xpano_b200.exposure import BlocksGainCompensator
    comp = BlocksGainCompensator(32, 32, iters)
    sizes = [(im.shape[1], im.shape[0]) for im in ims]
    pairs = comp.plan(cs, sizes)
    blocks, _ = go.make_blocks(cs, sizes)
    stats = {(i, j): (c, s1, s2)
             for i, j, c, s1, s2 in go.pair_stats(cs, ims, ms, blocks)}
    return comp, pairs, stats, blocks


@pytest.mark.parametrize("name", sorted(SCENES))
@pytest.mark.parametrize("iters", [0, 2])
def test_host_plan_and_solve(name, iters):
    g, cs, ims, ms = load_scene(name)
    comp, pairs, stats, blocks = host_maps(cs, ims, ms, iters)
    # the same pairs, each once, with the intersection rectangle
    assert sorted((int(p[0]), int(p[1])) for p in pairs) == sorted(stats)
    for i, j, x, y, w, h in pairs:
        _, ax0, ay0, ax1, ay1 = blocks[i]
        _, bx0, by0, bx1, by1 = blocks[j]
        assert (x, y) == (max(ax0, bx0), max(ay0, by0))
        assert (x + w, y + h) == (min(ax1, bx1), min(ay1, by1))
    st = [stats[(int(p[0]), int(p[1]))] for p in pairs]
    comp.solve([s[0] for s in st], [s[1] for s in st], [s[2] for s in st])
    n_pairs, unknowns, iterations = comp.stats()
    assert n_pairs == len(pairs) and 0 < iterations < 500
    for i, m in enumerate(comp.getMatGains()):
        ref = g[f"{name}_gain{iters}_{i}"]
        assert m.shape == ref.shape
        np.testing.assert_allclose(m, ref, rtol=2e-6)


def test_host_no_overlap_keeps_unit_gains():
    from xpano_b200.exposure import BlocksGainCompensator
    comp = BlocksGainCompensator()
    pairs = comp.plan([(0, 0), (500, 0)], [(64, 40), (40, 64)])
    assert all(p[0] == p[1] for p in pairs)
    comp.solve([p[4] * p[5] for p in pairs], [1.0] * len(pairs),
               [1.0] * len(pairs))
    assert comp.stats()[1] == 0
    for m in comp.getMatGains():
        assert np.all(m == 1.0)


def test_host_errors():
    from xpano_b200 import _lib
    from xpano_b200.exposure import BlocksGainCompensator
    comp = BlocksGainCompensator()
    with pytest.raises(_lib.XpError):
        comp.plan([(0, 0)], [(0, 10)])
    comp.plan([(0, 0)], [(10, 10)])
    with pytest.raises(_lib.XpError):
        comp.solve([1, 2], [0.0, 0.0], [0.0, 0.0])  # one pair planned
    with pytest.raises(_lib.XpError):
        BlocksGainCompensator(0, 32)


def test_host_solver_at_scale():
    """The sparse conjugate-gradient solve against a dense solve of the same
    system (GainCompensator::singleFeed's A, b) for a 24-image panorama's
    worth of blocks with made-up but plausible statistics."""
    from xpano_b200.exposure import BlocksGainCompensator
    rng = np.random.default_rng(7)
    n_img, w, h = 24, 420, 330
    corners = [(int(i % 8) * 300 + int(rng.integers(-20, 20)),
                int(i // 8) * 240 + int(rng.integers(-20, 20)))
               for i in range(n_img)]
    comp = BlocksGainCompensator(32, 32, 0)
    pairs = comp.plan(corners, [(w, h)] * n_img)
    n_pairs = len(pairs)
    area = pairs[:, 4].astype(np.int64) * pairs[:, 5]
    counts = (area * rng.uniform(0.0, 1.0, n_pairs)).astype(np.int64)
    counts[rng.random(n_pairs) < 0.1] = 0          # empty intersections
    exposure = rng.uniform(0.7, 1.4, n_img)
    blocks, per = go.make_blocks(corners, [(w, h)] * n_img)
    img_of = np.array([b[0] for b in blocks])
    mean = rng.uniform(60, 200, n_pairs)
    s1 = counts * mean * exposure[img_of[pairs[:, 0]]]
    s2 = counts * mean * exposure[img_of[pairs[:, 1]]]
    comp.solve(counts, s1, s2)
    _, unknowns, iterations = comp.stats()
    assert unknowns > 2500 and iterations < 400
    got = np.concatenate([m.ravel() for m in comp.getMatGains()])
    stats = [(int(p[0]), int(p[1]), int(c), float(a), float(b))
             for p, c, a, b in zip(pairs, counts, s1, s2)]
    want = go.solve_gains(len(blocks), stats).astype(np.float32)
    np.testing.assert_allclose(got, want, rtol=2e-6)

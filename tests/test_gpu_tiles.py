"""Canvas tiling of one panorama: the strips computed by the
ranks of a tiled run, put side by side, are BIT-IDENTICAL to the single-GPU
panorama. The ranks are emulated here one after the other on one device (the
partial sums are added on the host exactly as the all-reduce adds them); the
multi-process form of the same code is tests/test_sharding.py (gloo) and
bench.py --tiled (NCCL)."""
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


def _workload(n_cols, n_rows, w, h, f, bpp, pitches, seed, gains_seams):
    hfov = 2 * math.atan(w / 2 / f)
    Ks, Rs = co.make_cameras(n_cols, n_rows, w, h, f, 0.6 * hfov, pitches,
                             seed, None)
    ims = [synth.texture(h, w, 700 + i, 3, bpp) for i in range(len(Ks))]
    gains = seams = None
    if gains_seams:
        rng = np.random.default_rng(seed)
        gains = [rng.uniform(0.9, 1.1, (7, 9)).astype(np.float32)
                 for _ in ims] if bpp == 8 else None
        seams = []
        for _ in ims:
            m = np.full((40, 52), 255, np.uint8)
            m[:, :int(rng.integers(0, 6))] = 0   # cut a few columns away
            m[:int(rng.integers(0, 5))] = 0
            seams.append(m)
    return ims, Ks, Rs, gains, seams


def _tiled(xb, kw, wl, count):
    ims, Ks, Rs, gains, seams = wl
    comps, parts = [], []
    for r in range(count):
        c = xb.Composer(**kw)
        c.set_tile(r, count)
        c.set_images(ims, Ks, Rs, gains, seams)
        parts.append(c.run_tile_begin())
        comps.append(c)
    reduced = [sum(p[k] for p in parts) for k in range(8)]
    strips = []
    for c in comps:
        info = c.run_tile_end(reduced)
        pano, mask = c.fetch()
        strips.append((info, pano, mask, c.tile_needs(len(ims))))
    return strips


CASES = [
    # name, (n_cols, n_rows, w, h, f, bpp, pitches), levels, ranks, aux
    ("row8_L3", (8, 1, 400, 300, 330.0, 8, [0.0]), 3, (2, 3, 4), True),
    ("grid_L4", (6, 2, 360, 260, 300.0, 8, [-0.3, 0.3]), 4, (2, 5), True),
    ("row6_u16_L3", (6, 1, 420, 300, 340.0, 16, [0.0]), 3, (2, 4), False),
    ("row5_L2", (5, 1, 300, 220, 260.0, 8, [0.0]), 2, (8,), False),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_strips_equal_single_gpu(case):
    import xpano_b200 as xb
    name, shape, levels, ranks, aux = case
    n_cols, n_rows, w, h, f, bpp, pitches = shape
    wl = _workload(n_cols, n_rows, w, h, f, bpp, pitches, 31, aux)
    kw = dict(projection=wo.SPHERICAL, warp_scale=f, depth=bpp,
              fixed_levels=levels)
    one = xb.Composer(**kw)
    one.set_images(*wl)
    info1 = one.run()
    pano1, mask1 = one.fetch()
    W, H = info1["width"], info1["height"]
    halo = xb.sharding.tile_halo(levels)
    assert halo == 8 << levels
    for count in ranks:
        strips = _tiled(xb, kw, wl, count)
        x = 0
        uploads = 0
        for info, pano, mask, needs in strips:
            assert info["canvas"] == (W, H) and info["height"] == H
            assert info["levels"] == levels
            assert info["tile_x0"] == x, (name, count, info["tile_x0"], x)
            assert info["min_x"] == info1["min_x"] + x
            sw = info["width"]
            assert pano.shape == (H, sw, 3) and mask.shape == (H, sw)
            assert np.array_equal(mask, mask1[:, x:x + sw]), (name, count, x)
            d = pano.astype(int) - pano1[:, x:x + sw].astype(int)
            assert not d.any(), (name, count, x, int(np.abs(d).max()),
                                 np.argwhere(np.abs(d).max(axis=(0, 2)) > 0)
                                 [[0, -1]].ravel() if d.any() else None)
            x += sw
            uploads += sum(needs)
        assert x == W
        # a rank uploads only the sources that reach its strip
        if count >= 4 and n_cols >= 6:
            assert uploads < count * len(wl[0])


def test_tile_errors():
    import xpano_b200 as xb
    wl = _workload(3, 1, 200, 150, 180.0, 8, [0.0], 5, False)
    c = xb.Composer(projection=wo.SPHERICAL, warp_scale=180.0,
                    fixed_levels=3)
    c.set_tile(0, 2)
    c.set_images(*wl)
    with pytest.raises(xb._lib.XpError):
        c.run()                      # a tile runs through run_tile_begin/_end
    with pytest.raises(xb._lib.XpError):
        c.run_tile_end([0.0] * 8)    # before begin
    with pytest.raises(xb._lib.XpError):
        xb.Composer().set_tile(3, 2)
    one = xb.Composer(projection=wo.SPHERICAL, warp_scale=180.0,
                      fixed_levels=1)
    one.set_tile(1, 2)
    with pytest.raises(xb._lib.XpError):   # one level: nothing to tile
        one.set_images(*wl)

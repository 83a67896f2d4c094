"""Seeded random sweep of the compose path through the C ABI against the
compose oracle: projection, image count and layout, sizes, sample depth,
pyramid depth, seam direction, gain maps, seam masks and the size cap are
drawn per case. Bars as everywhere: canvas geometry, levels and mask
bit-exact (the separable projections' maps are the oracle's bit for bit);
colours within 1 LSB (8-bit) / 1e-3 relative (16-bit) at >= 99.9 % of the
pixels."""
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
PROJ = [wo.SPHERICAL, wo.CYLINDRICAL, wo.MERCATOR]


def draw(seed):
    rng = np.random.default_rng(1000 + seed)
    n_cols = int(rng.integers(2, 6))
    n_rows = int(rng.integers(1, 3))
    w = int(rng.integers(90, 330))
    h = int(rng.integers(70, 260))
    f = float(rng.uniform(0.7, 1.6) * w)
    bpp = 16 if seed % 4 == 3 else 8
    proj = PROJ[seed % 3]
    step = float(rng.uniform(0.45, 0.9))
    pitches = [0.0] if n_rows == 1 else [-0.22, 0.22]
    extra = None
    if seed % 5 == 0:
        extra = co.rot(*(math.radians(v) for v in rng.uniform(-12, 12, 3)))
    hfov = 2 * math.atan(w / 2 / f)
    Ks, Rs = co.make_cameras(n_cols, n_rows, w, h, f, step * hfov, pitches,
                             seed, extra)
    n = len(Ks)
    ims = [synth.texture(h, w, 5000 + 31 * seed + i, 3, bpp) for i in range(n)]
    kw = dict(fixed_levels=int(rng.integers(0, 6)),
              reverse=bool(rng.integers(0, 2)))
    gains = seams = None
    if bpp == 8 and seed % 2 == 0:
        gains = [rng.uniform(0.8, 1.25, (int(rng.integers(2, 9)),
                                         int(rng.integers(2, 9))))
                 .astype(np.float32) for _ in range(n)]
    if seed % 3 != 1:
        seams = []
        for _ in range(n):
            m = np.full((int(rng.integers(8, 30)), int(rng.integers(8, 30))),
                        255, np.uint8)
            m[rng.random(m.shape) < 0.12] = 0
            seams.append(m)
    cap = 100.0 if seed % 7 else float(rng.uniform(0.02, 0.06))
    return ims, Ks, Rs, proj, f, bpp, gains, seams, cap, kw


@pytest.mark.parametrize("seed", range(48))
def test_random_compose_case(seed):
    import xpano_b200 as xb
    ims, Ks, Rs, proj, f, bpp, gains, seams, cap, kw = draw(seed)
    want = co.compose(ims, Ks, Rs, projection=proj, warp_scale=f,
                      max_pano_mpx=cap, gains=gains, seam_masks=seams, **kw)
    c = xb.Composer(projection=proj, warp_scale=f, max_pano_mpx=cap,
                    depth=bpp, **kw)
    c.set_images(ims, Ks, Rs, gains, seams)
    info = c.run()
    pano, mask = c.fetch()
    assert info["resolution_capped"] == want["capped"]
    assert np.float32(info["warp_scale"]) == np.float32(want["warp_scale"])
    assert pano.shape == want["pano"].shape
    assert (info["min_x"], info["min_y"]) == (want["min_x"], want["min_y"])
    assert info["levels"] == want["levels"]
    assert np.array_equal(mask, want["mask"])
    both = mask > 0
    d = np.abs(pano.astype(int) - want["pano"].astype(int))[both]
    tol = 1 if bpp == 8 else np.maximum(
        1, 1e-3 * want["pano"].astype(int)[both])
    assert (d <= tol).mean() >= 0.999, (int(d.max()), float((d <= tol).mean()))

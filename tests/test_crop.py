"""Auto-crop (crop::FindLargestCrop, xpano/algorithm/auto_crop.cc:118-143):
the oracle (oracle/crop_oracle.py) against every known-answer test of the
reference (tests/auto_crop_test.cc), the library's host half against the
oracle on CPU, and the device path against both on the GPU."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle import crop_oracle as co  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ON = 255


def _full(h, w, v=ON):
    return np.full((h, w), v, np.uint8)


def _kats():
    """(name, mask, expected (x0, y0, x1, y1) | None): auto_crop_test.cc."""
    out = [("empty mask", _full(10, 20, 0), None),
           ("full even", _full(10, 20), (0, 0, 20, 10)),
           ("full odd", _full(10, 21), (0, 0, 21, 10)),
           ("single column", _full(10, 1), (0, 0, 1, 10)),
           ("two columns", _full(10, 2), (0, 0, 2, 10)),
           ("single row", _full(1, 20), (0, 0, 20, 1)),
           ("two rows", _full(2, 20), (0, 0, 20, 2))]
    m = _full(10, 20, 0)
    m[5] = ON
    out.append(("row set", m.copy(), (0, 5, 20, 6)))
    m[6] = ON
    out.append(("two rows set", m.copy(), (0, 5, 20, 7)))
    m = _full(10, 20)
    m[:, 5] = 0
    out.append(("empty column", m, (6, 0, 20, 10)))
    m = _full(6, 6)
    for r, c in [(0, 4), (1, 2), (1, 5), (3, 0), (4, 5), (5, 2)]:
        m[r, c] = 0
    out.append(("complex I", m, (1, 2, 5, 5)))
    m = _full(6, 6)
    m[2, 2] = 0
    m[3, 0] = 0
    out.append(("complex II", m, (3, 0, 6, 6)))
    g = np.load(os.path.join(GOLD, "crop_real_life_mask.npz"))
    h, w = g["shape"]
    real = (np.unpackbits(g["bits"])[:h * w].reshape(h, w) * 255).astype(
        np.uint8)
    out.append(("real life", real, tuple(int(v) for v in g["expected"])))
    return out


KATS = _kats()


@pytest.mark.parametrize("kat", KATS, ids=[k[0] for k in KATS])
def test_oracle_reproduces_reference_kats(kat):
    _, mask, want = kat
    assert co.find_largest_crop(mask) == want


def test_empty_matrix():
    assert co.find_largest_crop(np.zeros((0, 0), np.uint8)) is None
    import xpano_b200 as xb
    assert xb.FindLargestCrop(np.zeros((0, 0), np.uint8)) is None


def _random_masks():
    rng = np.random.default_rng(4)
    for k in range(12):
        h, w = int(rng.integers(1, 90)), int(rng.integers(1, 1500))
        m = np.full((h, w), 255, np.uint8)
        # a blobby boundary like a panorama mask, plus holes and 254s
        top = (rng.random(w) * h * 0.4).astype(int)
        bot = h - (rng.random(w) * h * 0.4).astype(int)
        yy = np.arange(h)[:, None]
        m[(yy < top[None, :]) | (yy >= bot[None, :])] = 0
        m[rng.random((h, w)) < 0.01] = int(rng.choice([0, 254, 1]))
        if k % 4 == 0:
            m[:, int(rng.integers(0, w))] = 0
        yield m


def test_host_half_matches_oracle():
    """xp_largest_crop_from_lines (host arithmetic, no device needed)."""
    from xpano_b200.auto_crop import largest_crop_from_lines
    for _, mask, want in KATS:
        lines = co.longest_lines(mask)
        assert largest_crop_from_lines(lines, mask.shape[0]) == want
    for m in _random_masks():
        lines = co.longest_lines(m)
        assert largest_crop_from_lines(lines, m.shape[0]) == \
            co.largest_crop_from_lines(lines, m.shape[0])


@pytest.mark.gpu
def test_device_crop_matches_reference_kats_and_oracle():
    import xpano_b200 as xb
    for name, mask, want in KATS:
        assert xb.FindLargestCrop(mask) == want, name
    for m in _random_masks():
        assert xb.FindLargestCrop(m) == co.find_largest_crop(m)
    # a row-strided mask (cv::Mat ROI)
    big = np.zeros((40, 300), np.uint8)
    big[5:30, 20:250] = 255
    view = big[:, 10:280]
    assert xb.FindLargestCrop(view) == co.find_largest_crop(
        np.ascontiguousarray(view))


@pytest.mark.gpu
def test_composer_auto_crop_on_the_device_mask():
    import math
    import compose_oracle as cmo
    import synth
    import xpano_b200 as xb
    n, w, h, f = 4, 400, 300, 300.0
    hfov = 2 * math.atan(w / 2 / f)
    Ks, Rs = cmo.make_cameras(n, 1, w, h, f, 0.62 * hfov, [0.0], 3, None)
    ims = [synth.texture(h, w, 40 + i, 3, 8) for i in range(n)]
    c = xb.Composer(projection=2, warp_scale=f, fixed_levels=4)
    c.set_images(ims, Ks, Rs)
    c.run()
    _, mask = c.fetch()
    got = c.auto_crop()
    assert got is not None and got == co.find_largest_crop(mask)
    x0, y0, x1, y1 = got
    assert (mask[y0:y1, x0:x1] == 255).all()

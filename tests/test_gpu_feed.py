"""GPU parity tests of the drop-in blender surface exactly as the stitcher
drives it (xpano/algorithm/stitcher.cc:357, 405, 419 through
xpano/algorithm/blenders.cc:116-158): prepare(roi), feed(img CV_8UC3, mask
CV_8U, tl) once per image, blend(dst, dst_mask) -- with row-strided host
buffers (cv::Mat ROIs), masks holding values other than 0/255 (Merge's
compare(mask, 0, CMP_NE), blenders.cc:99-112), 8- and 16-bit samples, and the
same through XP_MEM_DEVICE pointers.

Bars: dst_mask bit-exact; dst within 1 LSB at >= 99.99 % of pixels against
the oracle (mb_oracle.c, pinned to the reference library); device-pointer and
host-pointer paths bit-identical to each other.
"""
import math
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import compose_oracle as co  # noqa: E402
import synth  # noqa: E402
from oracle import mb_oracle  # noqa: E402
from oracle import warp_oracle as wo  # noqa: E402

pytestmark = pytest.mark.gpu


def _split(ims, seed):
    """BGRA test images -> (bgr, mask) pairs as the stitcher holds them: the
    mask is CV_8U with arbitrary non-zero values where the image counts."""
    rng = np.random.default_rng(seed)
    out = []
    for im in ims:
        a = im[..., 3] != 0
        vals = rng.choice(np.array([1, 2, 37, 128, 254, 255], np.uint8),
                          a.shape)
        out.append((np.ascontiguousarray(im[..., :3]),
                    np.where(a, vals, 0).astype(np.uint8)))
    return out


def _strided(a, pad, rng):
    """The same array as a view into a wider allocation (row step > width)."""
    h, w = a.shape[:2]
    big = rng.integers(0, 200, (h, w + pad) + a.shape[2:]).astype(a.dtype)
    big[:, 3:3 + w] = a
    v = big[:, 3:3 + w]
    assert not v.flags["C_CONTIGUOUS"] and np.array_equal(v, a)
    return v


def _expect(ims, pos, **kw):
    maxv = 255 if ims[0].dtype == np.uint8 else 65535
    fed = [np.dstack([im[..., :3],
                      np.where(im[..., 3] != 0, maxv, 0).astype(im.dtype)])
           for im in ims]
    o = mb_oracle.blend(fed, pos, **kw)
    dst = np.where(o.mask[..., None] > 0, np.moveaxis(o.planes, 0, 2), 0)
    return o, dst.astype(o.planes.dtype)


def _roi(ims, pos):
    x0 = min(p[0] for p in pos)
    y0 = min(p[1] for p in pos)
    x1 = max(p[0] + im.shape[1] for im, p in zip(ims, pos))
    y1 = max(p[1] + im.shape[0] for im, p in zip(ims, pos))
    return (x0, y0, x1 - x0, y1 - y0)


def _close(dst, want):
    assert dst.shape == want.shape and dst.dtype == want.dtype
    d = np.abs(dst.astype(np.int64) - want.astype(np.int64))
    assert d.max() <= 1 and (d > 0).mean() < 1e-4, (d.max(), (d > 0).mean())


SURFACE = [
    (41, dict(n=3, mode="ellipse", bpp=8), dict(fixed_levels=4)),
    (42, dict(n=4, mode="blobs", bpp=8, w=260, h=190), dict()),
    (43, dict(n=3, mode="ellipse", bpp=16), dict(fixed_levels=3)),
    (44, dict(n=2, mode="full", bpp=8), dict(fixed_levels=2)),
]


@pytest.mark.parametrize("case", SURFACE, ids=[str(c[0]) for c in SURFACE])
def test_prepare_feed_blend_host_buffers(case):
    import xpano_b200 as xb
    seed, ckw, bkw = case
    ims, pos = synth.blend_case(seed, **ckw)
    o, want = _expect(ims, pos, **bkw)
    rng = np.random.default_rng(seed)
    b = xb.Multiblend(**bkw)
    b.prepare(_roi(ims, pos))                       # stitcher.cc:357
    for k, ((bgr, mask), tl) in enumerate(zip(_split(ims, seed), pos)):
        if k % 2 == 0:                              # cv::Mat ROI: step > width
            bgr, mask = _strided(bgr, 11, rng), _strided(mask, 5, rng)
        b.feed(bgr, mask, tl)                       # stitcher.cc:405
    dst, dst_mask = b.blend()                       # stitcher.cc:419
    assert b.size == (o.width, o.height)
    assert b.min_pos == (o.min_x, o.min_y)
    assert np.array_equal(dst_mask, o.mask)
    assert set(np.unique(dst_mask)) <= {0, 255}
    _close(dst, want)
    assert not dst[dst_mask == 0].any()             # blenders.cc:151-154


@pytest.mark.parametrize("case", SURFACE[:3], ids=[str(c[0])
                                                   for c in SURFACE[:3]])
def test_prepare_feed_blend_device_pointers(case):
    """XP_MEM_DEVICE on every pointer of the surface: pitched device inputs,
    pitched device outputs; identical to the host-pointer result."""
    import torch
    import xpano_b200 as xb
    seed, ckw, bkw = case
    ims, pos = synth.blend_case(seed, **ckw)
    pairs = _split(ims, seed)
    host = xb.Multiblend(**bkw)
    host.prepare(_roi(ims, pos))
    for (bgr, mask), tl in zip(pairs, pos):
        host.feed(bgr, mask, tl)
    want_dst, want_mask = host.blend()

    dev = xb.Multiblend(**bkw)
    dev.prepare(_roi(ims, pos))
    keep = []
    for (bgr, mask), tl in zip(pairs, pos):
        h, w = mask.shape
        it = bgr.dtype.itemsize
        tdt = torch.uint8 if it == 1 else torch.uint16
        pad = 16                                    # samples of row padding
        tb = torch.zeros((h, w * 3 + pad), dtype=tdt, device="cuda")
        tb[:, :w * 3] = torch.from_numpy(bgr.reshape(h, w * 3)).cuda()
        tm = torch.full((h, w + 7), 9, dtype=torch.uint8, device="cuda")
        tm[:, :w] = torch.from_numpy(mask).cuda()
        keep += [tb, tm]
        torch.cuda.synchronize()
        dev.feed_device(tb.data_ptr(), tb.stride(0) * it, tm.data_ptr(),
                        tm.stride(0), w, h, 8 * it, tl)
    W, H = dev.run()
    it = want_dst.dtype.itemsize
    out = torch.full((H, W * 3 + 8), 7, device="cuda",
                     dtype=torch.uint8 if it == 1 else torch.uint16)
    om = torch.full((H, W + 3), 7, dtype=torch.uint8, device="cuda")
    dev.blend_device(out.data_ptr(), out.stride(0) * it, om.data_ptr(),
                     om.stride(0))
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert np.array_equal(got[:, :W * 3].reshape(H, W, 3), want_dst)
    assert (got[:, W * 3:] == 7).all()              # padding untouched
    gm = om.cpu().numpy()
    assert np.array_equal(gm[:, :W], want_mask) and (gm[:, W:] == 7).all()


def test_feed_errors_follow_the_reference():
    import xpano_b200 as xb
    lib = xb._lib.load()
    img = np.zeros((8, 8, 3), np.uint8)
    b = xb.Multiblend()
    # CV_Assert(input_mask.type() == CV_8U): no mask at all is an error
    st = lib.xp_blender_feed(b._h, img.ctypes.data, 0, None, 0, 8, 8, 8, 0, 0,
                             0)
    assert st == 1 and b"mask" in lib.xp_last_error()
    b.feed(np.full((8, 8, 3), 9, np.uint8), np.full((8, 8), 255, np.uint8),
           (0, 0))
    with pytest.raises(xb._lib.XpError):            # multiblend.cpp:100-117
        b.feed(np.zeros((8, 8, 3), np.uint16), np.full((8, 8), 1, np.uint8),
               (4, 0))
    b.feed(np.full((8, 8, 3), 200, np.uint8), np.full((8, 8), 3, np.uint8),
           (4, 0))
    dst, mask = b.blend()
    assert mask.shape == (8, 12) and mask.all()
    with pytest.raises(xb._lib.XpError):            # instances are single-use
        b.feed(img, np.full((8, 8), 255, np.uint8), (0, 0))
    # an all-zero mask: image without an opaque pixel
    b = xb.Multiblend()
    b.feed(img, np.zeros((8, 8), np.uint8), (0, 0))
    with pytest.raises(xb._lib.XpError) as e:
        b.blend()
    assert e.value.code == 4


@pytest.mark.parametrize("bpp", [8, 16])
def test_operator_surface_equals_compose_driver(bpp):
    """What stitcher.cc:358-419 would do through the shim -- warp(image),
    warp(mask), feed, blend -- gives the panorama the fused compose driver
    gives (both on the device), and both match the CPU compose oracle."""
    import xpano_b200 as xb
    n_cols, w_, h_, f = 4, 400, 300, 300.0
    hfov = 2 * math.atan(w_ / 2 / f)
    Ks, Rs = co.make_cameras(n_cols, 1, w_, h_, f, 0.62 * hfov, [0.0], 23,
                             None)
    ims = [synth.texture(h_, w_, 300 + i, 3, bpp) for i in range(len(Ks))]
    kw = dict(fixed_levels=4)
    creator = xb.PickWarper(xb.ProjectionType.kSpherical)
    warper = creator.create(f)
    rois = [warper.warpRoi((w_, h_), K, R) for K, R in zip(Ks, Rs)]
    x0 = min(r[0] for r in rois)
    y0 = min(r[1] for r in rois)
    x1 = max(r[0] + r[2] for r in rois)
    y1 = max(r[1] + r[3] for r in rois)
    blender = xb.Multiblend(**kw)
    blender.prepare((x0, y0, x1 - x0, y1 - y0))
    for im, K, R in zip(ims, Ks, Rs):
        corner, warped = warper.warp(im, K, R, xb.INTER_LINEAR,
                                     xb.BORDER_REFLECT)
        _, mask_warped = warper.warp(np.full((h_, w_), 255, np.uint8), K, R,
                                     xb.INTER_NEAREST, xb.BORDER_CONSTANT)
        blender.feed(warped, mask_warped, corner)
    dst, dst_mask = blender.blend()

    c = xb.Composer(projection=wo.SPHERICAL, warp_scale=f, depth=bpp, **kw)
    c.set_images(ims, Ks, Rs)
    c.run()
    pano, mask = c.fetch()
    assert np.array_equal(dst_mask, mask)
    assert np.array_equal(dst, pano)

    want = co.compose(ims, Ks, Rs, projection=wo.SPHERICAL, warp_scale=f,
                      **kw)
    assert np.array_equal(dst_mask, want["mask"])
    m = dst_mask > 0
    d = np.abs(dst.astype(int) - want["pano"].astype(int))[m]
    # 1 LSB for 8-bit samples, 1e-3 relative for 16-bit ones
    tol = 1 if bpp == 8 else np.maximum(1, 1e-3 * want["pano"].astype(int)[m])
    assert (d <= tol).mean() >= 0.999, (d.max(), (d <= tol).mean())


def test_compose_cancellation_between_images():
    """Stitcher::Cancelled() is polled between images (stitcher.cc:408-412):
    the composer polls a caller-owned flag at the same points."""
    import ctypes
    import xpano_b200 as xb
    n_cols, w_, h_, f = 3, 240, 180, 200.0
    hfov = 2 * math.atan(w_ / 2 / f)
    Ks, Rs = co.make_cameras(n_cols, 1, w_, h_, f, 0.6 * hfov, [0.0], 9, None)
    ims = [synth.texture(h_, w_, 60 + i, 3, 8) for i in range(n_cols)]
    c = xb.Composer(projection=wo.SPHERICAL, warp_scale=f, fixed_levels=3)
    c.set_images(ims, Ks, Rs)
    flag = ctypes.c_int(1)
    c.set_cancel_flag(flag)
    with pytest.raises(xb._lib.XpError) as e:
        c.run()
    assert e.value.code == 6
    flag.value = 0            # the same composer runs once the flag is clear
    info = c.run()
    pano, mask = c.fetch()
    assert pano.shape[:2] == (info["height"], info["width"]) and mask.any()


def test_run_in_two_halves_and_panoramas_in_flight():
    """run_begin / run_end (the batch-export form of xp_composer_run): the
    same panorama bit for bit; fetch completes a pending run; three composers
    driven by three host threads with the compute turn handed over at
    run_begin's return (bench.py's end-to-end leg) each give the panorama of a
    lone run; misuse is rejected."""
    import threading
    import xpano_b200 as xb
    n_cols, w_, h_, f = 4, 400, 300, 300.0
    hfov = 2 * math.atan(w_ / 2 / f)
    Ks, Rs = co.make_cameras(n_cols, 1, w_, h_, f, 0.6 * hfov, [0.0], 9, None)
    sets = [[synth.texture(h_, w_, 700 + 10 * k + i, 3, 8)
             for i in range(n_cols)] for k in range(3)]
    mk = lambda: xb.Composer(projection=wo.SPHERICAL, warp_scale=f,  # noqa
                             fixed_levels=5)
    want = []
    for ims in sets:
        c = mk()
        c.set_images(ims, Ks, Rs)
        c.run()
        want.append(c.fetch())
    c = mk()
    with pytest.raises(xb._lib.XpError):
        c.run_end()                        # nothing begun
    c.set_images(sets[0], Ks, Rs)
    c.run_begin()
    with pytest.raises(xb._lib.XpError):
        c.run_begin()                      # not ended yet
    with pytest.raises(xb._lib.XpError):
        c.set_images(sets[1], Ks, Rs)
    info = c.run_end()
    pano, mask = c.fetch()
    assert (info["width"], info["height"]) == pano.shape[1::-1]
    assert np.array_equal(pano, want[0][0]) and np.array_equal(mask, want[0][1])
    c.set_images(sets[1], Ks, Rs)
    c.run_begin()
    pano, mask = c.fetch()                 # completes the run
    assert np.array_equal(pano, want[1][0]) and np.array_equal(mask, want[1][1])

    comps = [mk() for _ in range(3)]
    turn = threading.Lock()
    got = [[None] * 4 for _ in range(3)]
    errs = []

    def worker(j):
        try:
            for rep in range(4):
                k = (j + rep) % 3
                comps[j].set_images(sets[k], Ks, Rs)
                with turn:
                    comps[j].run_begin()
                comps[j].run_end()
                got[j][rep] = (k,) + comps[j].fetch()
        except Exception as e:  # pragma: no cover
            errs.append(e)
    ts = [threading.Thread(target=worker, args=(j,)) for j in range(3)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    for j in range(3):
        for k, pano, mask in got[j]:
            assert np.array_equal(pano, want[k][0])
            assert np.array_equal(mask, want[k][1])


def test_mask_in_the_reference_rle_format():
    """Result.full_mask as utils::Flex words (mb/flex.h, flex.cpp:78-98),
    decoded the way xpano's ToMat does (blenders.cc:38-62)."""
    import xpano_b200 as xb
    from xpano_b200.blenders import decode_mask_rle
    for seed, ckw in ((51, dict(n=3, mode="ellipse")),
                      (52, dict(n=4, mode="blobs", w=330, h=210))):
        ims, pos = synth.blend_case(seed, **ckw)
        b = xb.Multiblend(fixed_levels=3)
        for im, p in zip(ims, pos):
            b.feed_interleaved(im, p)
        dst, mask = b.blend()
        runs, rows = b.mask_rle()
        H, W = mask.shape
        assert np.array_equal(decode_mask_rle(runs, W, H), mask)
        # maximal runs: colours alternate inside a row; Flex::rows_ offsets
        assert rows[0] == 0 and (np.diff(rows) > 0).all()
        for y in (0, H // 2, H - 1):
            a = rows[y] // 4
            e = rows[y + 1] // 4 if y + 1 < H else len(runs)
            r = runs[a:e]
            assert int((r & 0x7FFFFFFF).sum()) == W
            assert ((r[1:] >> 31) != (r[:-1] >> 31)).all()
        assert runs.nbytes < mask.nbytes / 10

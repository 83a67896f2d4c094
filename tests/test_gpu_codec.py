"""JPEG either side of the path through the C ABI (xp_codec_*)
against cv2's libjpeg-turbo: cv::imread of the sources
(xpano/algorithm/image.cc:83) and cv::imwrite of the panorama
(xpano/pipeline/stitcher_pipeline.cc:94-116).

Bars: baseline streams are decoded by the library's own kernels and equal
cv2.imdecode bit for bit (tests/test_jpeg_sm.py has the sweep); with
XP_JPEG_SM=0 every stream goes to nvJPEG, whose inverse DCT and chroma
upsampling are its own: PSNR >= 40 dB against cv2.imdecode (measured ~47 dB
on the reference's photographs), sizes exact. Encoded streams are valid JPEG
(cv2 decodes them), PSNR >= 38 dB against the source at quality 95, within a
factor 1.5 of cv2.imencode's size at the same settings. Compressed in,
compressed out: the panorama composed from sources decoded on the device equals
the one composed from the same decoded pixels uploaded from the host."""
import math
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle import warp_oracle as wo  # noqa: E402

cv2 = pytest.importorskip("cv2")
pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def psnr(a, b):
    mse = float(((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean())
    return 99.0 if mse == 0 else 10 * math.log10(255.0 ** 2 / mse)


@pytest.fixture(scope="module")
def photos():
    g = np.load(os.path.join(GOLD, "real_photos.npz"))
    return g, [g[f"jpeg_{i}"] for i in range(3)]


def test_decode_against_cv2(photos, monkeypatch):
    import xpano_b200 as xb
    _, jpegs = photos
    codec = xb.JpegCodec()
    for j in jpegs:
        ref = cv2.imdecode(j, cv2.IMREAD_COLOR)
        w, h, nc = codec.info(j)
        assert (h, w) == ref.shape[:2] and nc == 3
        got = codec.decode(j)
        assert got.shape == ref.shape and got.dtype == np.uint8
        assert np.array_equal(got, ref)          # our kernels: exact
    assert codec.stats() == (len(jpegs), 0)
    monkeypatch.setenv("XP_JPEG_SM", "0")        # nvJPEG for everything
    codec = xb.JpegCodec()
    for j in jpegs:
        ref = cv2.imdecode(j, cv2.IMREAD_COLOR)
        got = codec.decode(j)
        assert got.shape == ref.shape
        assert psnr(got, ref) >= 40.0, psnr(got, ref)
    assert codec.stats() == (0, len(jpegs))


def test_decode_errors(photos):
    import xpano_b200 as xb
    codec = xb.JpegCodec()
    with pytest.raises(xb._lib.XpError):
        codec.info(b"not a jpeg")     # tests/data/malformed.jpg's kind
    with pytest.raises(xb._lib.XpError):
        codec.decode(np.zeros(64, np.uint8))


@pytest.mark.parametrize("sub,prog,opt", [(422, False, False),
                                          (444, False, True),
                                          (420, True, False)])
def test_encode_against_cv2(photos, sub, prog, opt):
    import xpano_b200 as xb
    _, jpegs = photos
    src = cv2.imdecode(jpegs[0], cv2.IMREAD_COLOR)[:600, :800]
    codec = xb.JpegCodec()
    data = codec.encode(src, quality=95, subsampling=sub, progressive=prog,
                        optimize=opt)
    back = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
    assert back is not None and back.shape == src.shape
    assert psnr(back, src) >= 38.0, psnr(back, src)
    factor = {444: cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444,
              422: cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422,
              420: cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420}[sub]
    ok, ref = cv2.imencode(".jpg", src, [
        cv2.IMWRITE_JPEG_QUALITY, 95, cv2.IMWRITE_JPEG_PROGRESSIVE, int(prog),
        cv2.IMWRITE_JPEG_OPTIMIZE, int(opt),
        cv2.IMWRITE_JPEG_SAMPLING_FACTOR, factor])
    assert ok
    assert len(ref) / 1.5 <= len(data) <= len(ref) * 1.5, (len(data), len(ref))
    # strided input (a cv::Mat ROI)
    big = np.zeros((src.shape[0] + 2, src.shape[1] + 6, 3), np.uint8)
    big[1:-1, 3:-3] = src
    assert codec.encode(big[1:-1, 3:-3], quality=95, subsampling=sub,
                        progressive=prog, optimize=opt) == data


def test_compressed_in_compressed_out(photos):
    """JPEG sources decoded on the device straight into the buffers the
    composer warps from; the panorama encoded on the device, cropped by the
    auto-crop rectangle: no raw pixel crosses PCIe."""
    import torch
    import xpano_b200 as xb
    g, jpegs = photos
    codec = xb.JpegCodec()
    f = float(g["warped_image_scale"]) / float(g["work_scale"])
    # full-resolution sources: cameras registered on the 1024-px previews,
    # scaled to the 2048-px originals like compose_work_aspect does
    w, h, _ = codec.info(jpegs[0])
    s = max(w, h) / 1024.0
    Ks = []
    for i in range(3):
        k = np.array(g["K"][i], np.float64)
        k[0, 0] *= s / float(g["work_scale"])
        k[1, 1] *= s / float(g["work_scale"])
        k[0, 2] *= s / float(g["work_scale"])
        k[1, 2] *= s / float(g["work_scale"])
        Ks.append(k.astype(np.float32))
    Rs = [g["R"][i] for i in range(3)]
    dev = [torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
           for _ in range(3)]
    codec.decode_batch_device(jpegs, [t.data_ptr() for t in dev])
    torch.cuda.synchronize()
    one = torch.empty_like(dev[1])        # the batch equals one-by-one decodes
    codec.decode_device(jpegs[1], one.data_ptr())
    assert torch.equal(one, dev[1])
    assert np.array_equal(dev[1].cpu().numpy(),
                          cv2.imdecode(jpegs[1], cv2.IMREAD_COLOR))
    c = xb.Composer(projection=wo.SPHERICAL, warp_scale=f * s, depth=8)
    c.set_images_device([t.data_ptr() for t in dev], [(w, h)] * 3, Ks, Rs)
    info = c.run()
    pano, mask = c.fetch()
    # the same pixels through the host path
    c2 = xb.Composer(projection=wo.SPHERICAL, warp_scale=f * s, depth=8)
    c2.set_images([t.cpu().numpy() for t in dev], Ks, Rs)
    c2.run()
    pano2, mask2 = c2.fetch()
    assert np.array_equal(pano, pano2) and np.array_equal(mask, mask2)
    # twice the preview panorama of the reference's test (976 x 1335 +- 2 %)
    assert abs(info["height"] - 2 * 976) <= 0.02 * 2 * 976
    assert abs(info["width"] - 2 * 1335) <= 0.02 * 2 * 1335
    x0, y0, x1, y1 = c.auto_crop()
    data = c.encode_jpeg(codec, crop=(x0, y0, x1 - x0, y1 - y0))
    back = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
    assert back.shape == (y1 - y0, x1 - x0, 3)
    assert psnr(back, pano[y0:y1, x0:x1]) >= 38.0
    assert len(data) < pano[y0:y1, x0:x1].nbytes / 5
    whole = c.encode_jpeg(codec)
    assert cv2.imdecode(np.frombuffer(whole, np.uint8),
                        cv2.IMREAD_COLOR).shape == pano.shape

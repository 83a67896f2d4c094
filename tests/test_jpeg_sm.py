"""The SM JPEG decoder (xpano_b200/csrc/jpeg_sm.{h,cu}, jpeg_parse.h; SURVEY.md
8f2, the cv::imread side, xpano/algorithm/image.cc:83) against cv2.imdecode --
libjpeg-turbo, the decoder cv::imread uses. Bar: bit-exact, every pixel.
And the SM JPEG encoder (jpeg_enc.{h,cu}, jpeg_enc_host.h; the cv::imwrite
side, xpano/pipeline/stitcher_pipeline.cc:94-116) against cv2.imencode. Bar:
the stream is identical, byte for byte.

CPU part: tests/emu/jpeg_emu.cpp compiles the SAME headers as the CUDA kernels
and runs their steps as loops (guess pass, synchronisation sweeps in three
different orders, block scan, write pass, DC scan, inverse DCT, upsampling,
colour conversion). GPU part: the kernels themselves through xp_codec_*, with
XP_JPEG_SM=2 so that a fallback to nvJPEG would be an error."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "emu", "jpeg_emu.cpp")
LIB = os.path.join(HERE, "emu", "libjpeg_emu.so")
GOLD = os.path.join(HERE, "golden")
CSRC = os.path.join(os.path.dirname(HERE), "xpano_b200", "csrc")

SAMPLING = {444: cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444,
            422: cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422,
            420: cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420}


def scene(h, w, seed=0):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.stack([128 + 100 * np.sin(x / 37 + y / 91),
                    128 + 90 * np.cos(x / 13 - y / 29),
                    128 + 110 * np.sin(x * y / 9000)], -1)
    img += rng.normal(0, 12, img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def encode(img, sub=420, quality=90, restart=0, optimize=0, progressive=0):
    par = [cv2.IMWRITE_JPEG_QUALITY, quality,
           cv2.IMWRITE_JPEG_SAMPLING_FACTOR, SAMPLING[sub],
           cv2.IMWRITE_JPEG_OPTIMIZE, optimize,
           cv2.IMWRITE_JPEG_PROGRESSIVE, progressive]
    if restart:
        par += [cv2.IMWRITE_JPEG_RST_INTERVAL, restart]
    ok, buf = cv2.imencode(".jpg", img, par)
    assert ok
    return buf.reshape(-1)


def reference(buf):
    return cv2.imdecode(buf, cv2.IMREAD_COLOR | cv2.IMREAD_IGNORE_ORIENTATION)


CASES = [(h, w, sub, q, rst, opt)
         for (h, w) in [(64, 64), (100, 75), (333, 517), (17, 9), (8, 8),
                        (1, 1), (2, 3), (16, 33)]
         for sub in (444, 422, 420)
         for (q, rst, opt) in [(50, 0, 0), (90, 3, 1), (100, 0, 1),
                               (95, 1, 0)]]


# --------------------------------------------------------------------- CPU
@pytest.fixture(scope="module")
def emu():
    deps = [SRC] + [os.path.join(CSRC, n)
                    for n in ("jpeg_sm.h", "jpeg_parse.h", "hd.h",
                              "jpeg_enc.h", "jpeg_enc_host.h",
                              "jpeg_std_tables.h")]
    if not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB)
                                      for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC",
                               "-o", LIB, SRC])
    lib = ctypes.CDLL(LIB)
    lib.emu_jpeg_decode.argtypes = [ctypes.c_void_p, ctypes.c_size_t,
                                    ctypes.c_int, ctypes.c_void_p,
                                    ctypes.c_size_t, ctypes.c_void_p,
                                    ctypes.c_int]
    lib.emu_jpeg_why.restype = ctypes.c_char_p
    lib.emu_jpeg_encode.restype = ctypes.c_size_t
    lib.emu_jpeg_encode.argtypes = [ctypes.c_void_p, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_size_t,
                                    ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_void_p, ctypes.c_size_t]
    return lib


def emu_decode(lib, buf, sub_bytes=128, order=0):
    """-> (status, info (w, h, sweeps, subsequences) or reason, pixels)"""
    buf = np.ascontiguousarray(buf, np.uint8)
    info = (ctypes.c_int * 4)()
    r = lib.emu_jpeg_decode(buf.ctypes.data, buf.size, sub_bytes, None, 0,
                            info, order)
    if r:
        return r, lib.emu_jpeg_why().decode(), None
    out = np.zeros((info[1], info[0], 3), np.uint8)
    r = lib.emu_jpeg_decode(buf.ctypes.data, buf.size, sub_bytes,
                            out.ctypes.data, out.strides[0], info, order)
    return r, tuple(info), out


def test_emulated_decoder_equals_cv2(emu):
    for k, (h, w, sub, q, rst, opt) in enumerate(CASES):
        buf = encode(scene(h, w, k), sub, q, rst, opt)
        want = reference(buf)
        for sub_bytes, order in ((16, 0), (128, 1), (48, 2)):
            r, info, got = emu_decode(emu, buf, sub_bytes, order)
            assert r == 0, (h, w, sub, q, rst, opt, info)
            assert np.array_equal(got, want), (h, w, sub, q, rst, opt,
                                               sub_bytes, order)


def test_emulated_decoder_grey_and_photos(emu):
    grey = scene(90, 130, 5)[..., 0]
    ok, buf = cv2.imencode(".jpg", grey, [cv2.IMWRITE_JPEG_QUALITY, 85])
    r, info, got = emu_decode(emu, buf.reshape(-1))
    assert r == 0 and np.array_equal(got, reference(buf.reshape(-1)))
    g = np.load(os.path.join(GOLD, "real_photos.npz"))
    for i in range(3):       # the reference's own photographs (4:2:0)
        j = g[f"jpeg_{i}"]
        r, info, got = emu_decode(emu, j, 128, i)
        assert r == 0 and info[3] > 1000, info
        assert np.array_equal(got, reference(j))


def test_synchronisation_is_order_independent(emu):
    """Jacobi sweeps, ascending and descending in-place sweeps end in the same
    states (the sequential decoder's); short subsequences need many sweeps."""
    buf = encode(scene(333, 517, 9), 420, 98)
    want = reference(buf)
    sweeps = []
    for order in (0, 1, 2):
        r, info, got = emu_decode(emu, buf, 8, order)
        assert r == 0 and np.array_equal(got, want)
        sweeps.append(info[2])
    assert sweeps[1] <= 2 and sweeps[0] >= sweeps[1]


def test_uncovered_streams_are_reported(emu):
    img = scene(64, 64)
    r, why, _ = emu_decode(emu, encode(img, progressive=1))
    assert r == 1 and "sequential" in why
    r, why, _ = emu_decode(emu, np.frombuffer(b"not a jpeg at all", np.uint8))
    assert r == 1 and "not a JPEG" in why
    buf = encode(img, 420, 90).copy()
    r, why, _ = emu_decode(emu, buf[:200])          # truncated header
    assert r == 1


def test_damaged_streams_end_cleanly(emu):
    """Flipped bytes in the scan or the headers and truncated files (the kind
    of tests/data/malformed.jpg): the decoder reports the stream as not
    covered or decodes what is there; its synchronisation always converges.
    (The same loop ran clean under -fsanitize=address,undefined.)"""
    rng = np.random.default_rng(7)
    bases = [encode(scene(40, 56, k), sub, q, rst)
             for k, (sub, q, rst) in enumerate([(420, 90, 0), (422, 95, 2),
                                                (444, 50, 0), (420, 100, 1)])]
    seen = set()
    for it in range(400):
        b = bases[it % len(bases)].copy()
        if it % 3 == 0:
            for _ in range(rng.integers(1, 6)):
                b[rng.integers(300, b.size)] = rng.integers(0, 256)
        elif it % 3 == 1:
            for _ in range(rng.integers(1, 4)):
                b[rng.integers(2, 620)] = rng.integers(0, 256)
        else:
            b = b[:rng.integers(4, b.size)].copy()
        info = (ctypes.c_int * 4)()
        r = emu.emu_jpeg_decode(b.ctypes.data, b.size, 32, None, 0, info, 0)
        if r == 0 and info[0] * info[1] <= 1 << 22:
            out = np.zeros((info[1], info[0], 3), np.uint8)
            r = emu.emu_jpeg_decode(b.ctypes.data, b.size, 32, out.ctypes.data,
                                    out.strides[0], info, it % 3)
        assert r in (0, 1), (it, r)
        seen.add(r)
    assert seen == {0, 1}


ENC_CASES = [(h, w, sub, q)
             for (h, w) in [(64, 64), (16, 16), (8, 8), (100, 75), (333, 517),
                            (17, 9), (1, 1), (2, 3), (16, 33), (24, 40),
                            (9, 17)]
             for sub in (444, 422, 420) for q in (10, 50, 95, 100)]


def emu_encode(lib, img, quality, sub, optimize=0):
    img = np.ascontiguousarray(img)
    h, w = img.shape[:2]
    n = lib.emu_jpeg_encode(img.ctypes.data, w, h, img.strides[0], quality,
                            sub, optimize, None, 0)
    out = np.zeros(n, np.uint8)
    assert lib.emu_jpeg_encode(img.ctypes.data, w, h, img.strides[0], quality,
                               sub, optimize, out.ctypes.data, n) == n
    return out


def test_emulated_encoder_equals_cv2(emu):
    """The encoder's arithmetic (jpeg_enc.h / jpeg_enc_host.h): the stream is
    cv2.imencode's byte for byte -- headers, quantisation tables, edge
    expansion, dummy blocks of partial MCUs, Huffman coding, padding."""
    for k, (h, w, sub, q) in enumerate(ENC_CASES):
        img = scene(h, w, 100 + k)
        for opt in (0, 1):   # standard tables / jpeg_gen_optimal_table
            assert np.array_equal(emu_encode(emu, img, q, sub, opt),
                                  encode(img, sub, q, optimize=opt)), \
                (h, w, sub, q, opt)
    g = np.load(os.path.join(GOLD, "real_photos.npz"))
    photo = reference(g["jpeg_0"])[:700, :900]
    for sub in (444, 422, 420):
        for opt in (0, 1):
            assert np.array_equal(emu_encode(emu, photo, 95, sub, opt),
                                  encode(photo, sub, 95, optimize=opt))


def test_emulated_codec_random_sweep(emu):
    """Seeded random sizes (1..259), qualities (1..100), samplings, noise /
    flat / smooth content, standard and optimised tables: the encoder's stream
    equals cv2.imencode's, and the decoder decodes it to cv2.imdecode's
    pixels."""
    rng = np.random.default_rng(123)
    for it in range(90):
        h, w = int(rng.integers(1, 260)), int(rng.integers(1, 260))
        sub, q, opt = (444, 422, 420)[it % 3], int(rng.integers(1, 101)), it % 2
        if it % 4 == 0:
            img = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
        elif it % 4 == 1:
            img = np.full((h, w, 3), rng.integers(0, 256, 3), np.uint8)
        else:
            img = scene(h, w, it)
        got = emu_encode(emu, img, q, sub, opt)
        assert np.array_equal(got, encode(img, sub, q, optimize=opt)), \
            (h, w, sub, q, opt)
        r, info, pixels = emu_decode(emu, got, 64, it % 3)
        assert r == 0 and np.array_equal(pixels, reference(got)), (h, w, sub, q)


def test_standard_tables_header_is_current():
    subprocess.check_call([sys.executable,
                           os.path.join(os.path.dirname(HERE), "tools",
                                        "make_jpeg_tables.py"), "--check"])


# --------------------------------------------------------------------- GPU
@pytest.fixture()
def sm_only(monkeypatch):
    monkeypatch.setenv("XP_JPEG_SM", "2")


@pytest.mark.gpu
def test_gpu_decoder_equals_cv2(sm_only):
    import xpano_b200 as xb
    codec = xb.JpegCodec()
    for k, (h, w, sub, q, rst, opt) in enumerate(CASES):
        buf = encode(scene(h, w, k), sub, q, rst, opt)
        got = codec.decode(buf)
        assert np.array_equal(got, reference(buf)), (h, w, sub, q, rst, opt)
    assert codec.stats() == (len(CASES), 0)


@pytest.mark.gpu
def test_gpu_decoder_photos_grey_and_large(sm_only):
    import torch
    import xpano_b200 as xb
    codec = xb.JpegCodec()
    g = np.load(os.path.join(GOLD, "real_photos.npz"))
    jpegs = [g[f"jpeg_{i}"] for i in range(3)]
    for j in jpegs:      # portrait photographs: EXIF orientation applied
        want = cv2.imdecode(j, cv2.IMREAD_COLOR)
        assert want.shape[0] > want.shape[1]
        assert np.array_equal(codec.decode(j), want)
    grey = scene(90, 130, 5)[..., 0]
    ok, buf = cv2.imencode(".jpg", grey, [cv2.IMWRITE_JPEG_QUALITY, 85])
    assert np.array_equal(codec.decode(buf.reshape(-1)),
                          reference(buf.reshape(-1)))
    # a 12-MP source (BASELINE cfg2's size), with and without restart markers
    big = cv2.resize(reference(jpegs[0]), (4000, 3000),
                     interpolation=cv2.INTER_CUBIC)
    for rst in (0, 250):
        buf = encode(big, 420, 95, rst)
        assert np.array_equal(codec.decode(buf), reference(buf))
    # the batch entry point into device memory, strided rows
    bufs = [encode(big[:1500 + 8 * i, :2000 + 8 * i], 422, 92) for i in range(5)]
    dev, want = [], []
    for b in bufs:
        ref = reference(b)
        want.append(ref)
        dev.append(torch.zeros((ref.shape[0], ref.shape[1] * 3 + 32),
                               dtype=torch.uint8, device="cuda"))
    codec.decode_batch_device(bufs, [t.data_ptr() for t in dev],
                              [t.shape[1] for t in dev])
    torch.cuda.synchronize()
    for t, ref in zip(dev, want):
        got = t.cpu().numpy()
        assert np.array_equal(got[:, :ref.shape[1] * 3].reshape(ref.shape), ref)
        assert not got[:, ref.shape[1] * 3:].any()
    assert codec.stats()[1] == 0


@pytest.mark.gpu
def test_gpu_decoder_orientation_and_fallback(monkeypatch):
    import xpano_b200 as xb
    img = scene(120, 200, 3)
    buf = encode(img, 420, 90)
    # EXIF orientation 6 (rotate 90 clockwise), applied like cv::imread does
    exif = (b"Exif\0\0MM\0*\0\0\0\x08\0\x01\x01\x12\0\x03\0\0\0\x01\0\x06\0\0"
            b"\0\0\0\0")
    app1 = b"\xff\xe1" + (len(exif) + 2).to_bytes(2, "big") + exif
    rot = np.frombuffer(buf[:2].tobytes() + app1 + buf[2:].tobytes(), np.uint8)
    want = cv2.imdecode(rot, cv2.IMREAD_COLOR)
    assert want.shape == (200, 120, 3)
    monkeypatch.setenv("XP_JPEG_SM", "2")
    codec = xb.JpegCodec()
    assert np.array_equal(codec.decode(rot), want)
    prog = encode(img, 420, 90, progressive=1)
    with pytest.raises(xb._lib.XpError, match="SM decoder"):
        codec.decode(prog)
    # default mode: what the kernels do not cover goes to nvJPEG
    monkeypatch.setenv("XP_JPEG_SM", "1")
    codec = xb.JpegCodec()
    got = codec.decode(prog)
    assert got.shape == img.shape
    assert np.array_equal(codec.decode(buf), reference(buf))
    assert codec.stats() == (1, 1)
    monkeypatch.setenv("XP_JPEG_SM", "0")
    codec = xb.JpegCodec()
    codec.decode(buf)
    assert codec.stats() == (0, 1)


@pytest.mark.gpu
def test_gpu_encoder_equals_cv2(sm_only):
    """xp_codec_jpeg_encode through our own kernels: byte-identical to
    cv2.imencode (= cv::imwrite's stream) at every size, sampling, quality."""
    import xpano_b200 as xb
    codec = xb.JpegCodec()
    for k, (h, w, sub, q) in enumerate(ENC_CASES):
        img = scene(h, w, 100 + k)
        opt = k % 2          # standard tables / tables optimised for the image
        got = np.frombuffer(codec.encode(img, quality=q, subsampling=sub,
                                         optimize=bool(opt)), np.uint8)
        assert np.array_equal(got, encode(img, sub, q, optimize=opt)), \
            (h, w, sub, q, opt)
    g = np.load(os.path.join(GOLD, "real_photos.npz"))
    big = cv2.resize(reference(g["jpeg_0"]), (4000, 3000),
                     interpolation=cv2.INTER_CUBIC)
    for sub, opt in ((422, 0), (420, 1)):
        got = np.frombuffer(codec.encode(big, quality=95, subsampling=sub,
                                         optimize=bool(opt)), np.uint8)
        assert np.array_equal(got, encode(big, sub, 95, optimize=opt))
    view = big[5:1205, 7:1800]            # a strided view (a cv::Mat ROI)
    got = np.frombuffer(codec.encode(view, quality=90, subsampling=444),
                        np.uint8)
    assert np.array_equal(got, encode(np.ascontiguousarray(view), 444, 90))
    st = codec.all_stats()
    assert st[2] == len(ENC_CASES) + 3 and st[3] == 0
    # what the kernels do not cover is an error here, nvJPEG's by default
    with pytest.raises(xb._lib.XpError, match="SM encoder"):
        codec.encode(big[:64, :64], progressive=True)


@pytest.mark.gpu
def test_gpu_damaged_streams_end_cleanly():
    """The kernels on damaged streams: an error or some pixels, never a hang
    or a fault -- the next valid stream still decodes exactly."""
    import xpano_b200 as xb
    codec = xb.JpegCodec()
    rng = np.random.default_rng(11)
    bases = [encode(scene(120, 168, k), sub, q, rst)
             for k, (sub, q, rst) in enumerate([(420, 90, 0), (422, 95, 2),
                                                (444, 50, 0)])]
    outcomes = set()
    for it in range(45):
        b = bases[it % len(bases)].copy()
        if it % 3 == 0:
            for _ in range(rng.integers(1, 6)):
                b[rng.integers(700, b.size)] = rng.integers(0, 256)
        elif it % 3 == 1:
            b[rng.integers(2, 620)] = rng.integers(0, 256)
        else:
            b = b[:rng.integers(700, b.size)].copy()
        try:
            got = codec.decode(b)
            assert got.ndim == 3 and got.shape[2] == 3
            outcomes.add("pixels")
        except xb._lib.XpError:
            outcomes.add("error")
    assert "pixels" in outcomes
    good = bases[0]
    assert np.array_equal(codec.decode(good), reference(good))


@pytest.mark.gpu
def test_gpu_codec_random_sweep(sm_only):
    """The kernels on seeded random sizes, qualities, samplings and content:
    encoded streams equal cv2.imencode's, decoded pixels cv2.imdecode's."""
    import xpano_b200 as xb
    codec = xb.JpegCodec()
    rng = np.random.default_rng(321)
    for it in range(60):
        h, w = int(rng.integers(1, 700)), int(rng.integers(1, 900))
        sub, q, opt = (444, 422, 420)[it % 3], int(rng.integers(1, 101)), it % 2
        if it % 4 == 0:
            img = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
        elif it % 4 == 1:
            img = np.full((h, w, 3), rng.integers(0, 256, 3), np.uint8)
        else:
            img = scene(h, w, it)
        got = np.frombuffer(codec.encode(img, quality=q, subsampling=sub,
                                         optimize=bool(opt)), np.uint8)
        assert np.array_equal(got, encode(img, sub, q, optimize=opt)), \
            (h, w, sub, q, opt)
        rst = int(rng.integers(0, 4))
        src = encode(img, sub, q, rst, opt)
        assert np.array_equal(codec.decode(src), reference(src)), \
            (h, w, sub, q, rst)
    assert codec.all_stats() == (60, 0, 60, 0)

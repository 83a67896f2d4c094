"""The reference-side binding compiles and runs: include/xpano_b200_shim.hpp
(B200Warper / B200WarperCreator / B200Blender / B200GainCompensator, the
classes INTEGRATION.md
tells a maintainer to add) is compiled against a stand-in for OpenCV's
declarations (tests/cv_stub/, OpenCV's C++ headers are not in this image) and
driven polymorphically like stitcher.cc:144-169, 357-419 drives the originals;
and a plain-C program (gcc) uses include/xpano_b200.h directly."""
import math
import os
import struct
import subprocess
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
BUILD = os.path.join(HERE, "build")
LIBDIR = os.path.join(ROOT, "xpano_b200")


def build_native():
    """-> (shim_driver, c_abi_smoke) paths; compiled when missing or stale."""
    os.makedirs(BUILD, exist_ok=True)
    link = ["-L", LIBDIR, "-lxpano_b200", "-Wl,-rpath," + LIBDIR]
    jobs = [
        ("shim_driver", ["g++", "-std=c++17", "-O1", "-Wall", "-Werror",
                         "-I", os.path.join(HERE, "cv_stub"),
                         "-I", os.path.join(ROOT, "include"),
                         os.path.join(HERE, "shim_driver.cpp")],
         ["shim_driver.cpp", "../include/xpano_b200_shim.hpp",
          "../include/xpano_b200.h", "cv_stub/opencv2/core.hpp",
          "cv_stub/opencv2/stitching/detail/exposure_compensate.hpp"]),
        ("c_abi_smoke", ["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror",
                         "-O1", "-I", os.path.join(ROOT, "include"),
                         os.path.join(HERE, "c_abi_smoke.c")],
         ["c_abi_smoke.c", "../include/xpano_b200.h"]),
    ]
    out = []
    for name, cmd, deps in jobs:
        exe = os.path.join(BUILD, name)
        newest = max(os.path.getmtime(os.path.join(HERE, d)) for d in deps)
        if not os.path.exists(exe) or os.path.getmtime(exe) < newest:
            subprocess.check_call(cmd + ["-o", exe] + link)
        out.append(exe)
    return out


def test_shim_and_c_program_compile():
    if not os.path.exists(os.path.join(LIBDIR, "libxpano_b200.so")):
        import __graft_entry__ as ge
        ge.build()
    shim, csmoke = build_native()
    assert os.access(shim, os.X_OK) and os.access(csmoke, os.X_OK)
    import xpano_b200
    if xpano_b200._lib.load().xp_device_count() == 0:
        # no CPU fallback: the C program reports the missing device and stops
        r = subprocess.run([csmoke], capture_output=True, text=True)
        assert r.returncode == 77 and "no CUDA device" in r.stderr


def _rot_y(a):
    return np.array([[math.cos(a), 0, math.sin(a)], [0, 1, 0],
                     [-math.sin(a), 0, math.cos(a)]], np.float32)


@pytest.mark.gpu
def test_shim_driver_equals_python_surface(tmp_path):
    import xpano_b200 as xb
    shim, _ = build_native()
    n, w, h, scale = 4, 420, 300, 310.0
    K = np.array([[320, 0, w / 2], [0, 320, h / 2], [0, 0, 1]], np.float32)
    Rs = [_rot_y(0.5 * (i - 1.5)) for i in range(n)]
    ims = [np.clip(synth.texture(h, w, 500 + i, 3, 8).astype(np.float32)
                   * (0.8 + 0.1 * i), 0, 255).astype(np.uint8)
           for i in range(n)]  # exposures differ: the compensator has work
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    with open(fin, "wb") as f:
        f.write(struct.pack("<4if", n, w, h, int(xb.ProjectionType.kSpherical),
                            scale))
        for R in Rs:
            f.write(K.tobytes() + np.ascontiguousarray(R).tobytes())
        for im in ims:
            f.write(im.tobytes())
    r = subprocess.run([shim, str(fin), str(fout)], capture_output=True,
                       text=True)
    assert r.returncode == 0 and "shim ok" in r.stdout, r.stderr
    raw = np.fromfile(fout, np.uint8)
    W, H, cx, cy = (int(v) for v in raw[:16].view(np.int32))
    pano = raw[16:16 + W * H * 3].reshape(H, W, 3)
    mask = raw[16 + W * H * 3:16 + W * H * 4].reshape(H, W)
    tail = raw[16 + W * H * 4:]
    gw, gh = (int(v) for v in tail[:8].view(np.int32))
    gain0 = tail[8:8 + 4 * gw * gh].view(np.float32).reshape(gh, gw)

    warper = xb.PickWarper(xb.ProjectionType.kSpherical).create(scale)
    blender = xb.Multiblend(output_bpp=8)
    blender.prepare((0, 0, 1, 1))
    corners, warped, masks = [], [], []
    for im, R in zip(ims, Rs):
        corner, wi = warper.warp(im, K, R)
        _, mw = warper.warp(np.full((h, w), 255, np.uint8), K, R,
                            xb.INTER_NEAREST, xb.BORDER_CONSTANT)
        corners.append(corner)
        warped.append(wi)
        masks.append(mw)
    comp = xb.BlocksGainCompensator()
    comp.feed(corners, warped, masks)
    assert np.array_equal(comp.getMatGains()[0], gain0)
    assert gain0.min() < 0.99 or gain0.max() > 1.01
    for i in range(n):
        blender.feed(comp.apply(i, corners[i], warped[i], masks[i]), masks[i],
                     corners[i])
    dst, dst_mask = blender.blend()
    assert (W, H) == dst.shape[1::-1] and (cx, cy) == corners[0]
    assert np.array_equal(mask, dst_mask)
    assert np.array_equal(pano, dst)

    # the same loop with the BlendingMethod::kOpenCV blender installed
    r = subprocess.run([shim, str(fin), str(fout), "opencv"],
                       capture_output=True, text=True)
    assert r.returncode == 0 and "shim ok" in r.stdout, r.stderr
    raw = np.fromfile(fout, np.uint8)
    W2, H2 = (int(v) for v in raw[:8].view(np.int32))
    pano2 = raw[16:16 + W2 * H2 * 3].reshape(H2, W2, 3)
    mask2 = raw[16 + W2 * H2 * 3:16 + W2 * H2 * 4].reshape(H2, W2)
    x0 = min(c[0] for c in corners)
    y0 = min(c[1] for c in corners)
    x1 = max(c[0] + w_.shape[1] for c, w_ in zip(corners, warped))
    y1 = max(c[1] + w_.shape[0] for c, w_ in zip(corners, warped))
    mb = xb.MultiBandOpenCV()
    mb.prepare((x0, y0, x1 - x0, y1 - y0))
    for i in range(n):
        mb.feed(comp.apply(i, corners[i], warped[i], masks[i]), masks[i],
                corners[i])
    dst2, dst_mask2 = mb.blend()
    assert (W2, H2) == dst2.shape[1::-1]
    assert np.array_equal(mask2, dst_mask2) and np.array_equal(pano2, dst2)


@pytest.mark.gpu
def test_plain_c_program_runs_the_path():
    import xpano_b200 as xb
    _, csmoke = build_native()
    r = subprocess.run([csmoke], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    tok = r.stdout.split()
    assert tok[:2] == ["c_abi", "ok"]
    Wc, Hc, mx, my, bpp, csum, msum = (int(v) for v in tok[2:9])
    # the same inputs through the Python mirror
    W, H, N = 320, 200, 3
    K = np.array([[260, 0, W / 2], [0, 260, H / 2], [0, 0, 1]], np.float32)
    y, x, c = np.mgrid[0:H, 0:W, 0:3]
    warper = xb.RotationWarper(xb.ProjectionType.kSpherical, 250.0)
    blender = xb.Multiblend(fixed_levels=4)
    f32 = np.float32
    for i in range(N):
        img = (40 + 60 * i + ((x * 3 + y * 5 + c * 17) % 90)).astype(np.uint8)
        a = f32(0.45) * f32(i - 1)
        ca = f32(1) - a * a / f32(2) + a * a * a * a / f32(24)
        sa = a - a * a * a / f32(6)
        R = np.array([[ca, 0, sa], [0, 1, 0], [-sa, 0, ca]], np.float32)
        corner, wi = warper.warp(img, K, R)
        _, wm = warper.warp(np.full((H, W), 255, np.uint8), K, R,
                            xb.INTER_NEAREST, xb.BORDER_CONSTANT)
        blender.feed(wi, wm, corner)
    dst, dmask = blender.blend()
    assert (Wc, Hc) == blender.size and (mx, my) == blender.min_pos
    assert bpp == 8 and msum == int((dmask != 0).sum())
    s = 0  # the C program's polynomial checksum, mod 2^64
    for v in dst.reshape(-1).tolist():
        s = (s * 31 + v) & 0xFFFFFFFFFFFFFFFF
    assert s == csum

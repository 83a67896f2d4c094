"""Seeded synthetic inputs shared by the parity tests and golden generator.

Image content follows SURVEY.md section 8(d): band-limited sinusoids plus
small hash noise and a per-image gain, so that warp parity is not dominated by
1/32-px bucket flips on white noise.
"""
from __future__ import annotations

import numpy as np


def texture(h, w, seed, channels=3, bpp=8):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    planes = []
    for c in range(channels):
        phi = rng.uniform(0, 1)
        base = (128 + 60 * np.sin(2 * np.pi * (xx / 97.0 + phi))
                * np.cos(2 * np.pi * (yy / 131.0))
                + 40 * np.sin(2 * np.pi * ((xx + yy) / 211.0)))
        planes.append(base)
    img = np.stack(planes, -1) * rng.uniform(0.9, 1.1)
    img += rng.integers(-4, 5, img.shape)
    img = np.clip(img, 0, 255)
    if bpp == 16:
        return (img * 257).astype(np.uint16)
    return img.astype(np.uint8)


def alpha_mask(h, w, seed, mode):
    rng = np.random.default_rng(seed + 7919)
    yy, xx = np.mgrid[0:h, 0:w]
    if mode == "full":
        return np.ones((h, w), bool)
    if mode == "ellipse":
        cx, cy = w / 2 + rng.uniform(-10, 10), h / 2 + rng.uniform(-10, 10)
        return (((xx - cx) / (w * rng.uniform(0.4, 0.6))) ** 2
                + ((yy - cy) / (h * rng.uniform(0.4, 0.6))) ** 2) < 1
    if mode == "blobs":
        return (np.sin(xx / 9.0 + rng.uniform(0, 6))
                + np.cos(yy / 7.0 + rng.uniform(0, 6))) > -0.6
    raise ValueError(mode)


def blend_case(seed, n=3, w=200, h=150, bpp=8, mode="ellipse", spp=4):
    """n overlapping BGRA images with positions, deterministic in `seed`."""
    rng = np.random.default_rng(seed)
    ims, pos = [], []
    maxv = 255 if bpp == 8 else 65535
    dt = np.uint8 if bpp == 8 else np.uint16
    for i in range(n):
        ww = int(w * rng.uniform(0.7, 1.3))
        hh = int(h * rng.uniform(0.7, 1.3))
        img = texture(hh, ww, seed * 100 + i, 3, bpp)
        if spp == 4:
            a = alpha_mask(hh, ww, seed * 100 + i, mode)
            # arbitrary colour under transparent pixels: the inpaint stage
            # must overwrite it
            img[~a] = rng.integers(0, maxv, (int((~a).sum()), 3)).astype(dt)
            img = np.dstack([img, (a * maxv).astype(dt)])
        ims.append(np.ascontiguousarray(img))
        pos.append((int(rng.integers(-50, int(w * 0.8) * max(1, n // 2))),
                    int(rng.integers(-30, h // 2))))
    return ims, pos

"""Synthetic registered cameras for tests and bench.py.
Registration is outside the export path, so cameras are given: pin-hole K and
R = Ry(yaw) Rx(pitch) Rz(roll) with small seeded jitter."""
from __future__ import annotations

import math

import numpy as np


def make_cameras(n_cols, n_rows, w, h, focal, yaw_step, pitches, seed,
                 extra=None):
    """Synthetic registered cameras: pin-hole K,
    R = Ry(yaw) Rx(pitch) Rz(roll) with small seeded jitter."""
    rng = np.random.default_rng(seed)
    Ks, Rs = [], []
    for r in range(n_rows):
        for c in range(n_cols):
            yaw = (c - (n_cols - 1) / 2) * yaw_step
            pitch = pitches[r] + math.radians(rng.uniform(-1, 1))
            roll = math.radians(rng.uniform(-0.5, 0.5))
            R = rot(yaw, pitch, roll)
            if extra is not None:  # algorithm::Rotate: R <- M * R
                R = (np.asarray(extra, np.float32) @ R).astype(np.float32)
            Ks.append(np.array([[focal, 0, w / 2], [0, focal, h / 2],
                                [0, 0, 1]], np.float32))
            Rs.append(R)
    return Ks, Rs


def rot(yaw, pitch, roll):
    cy, sy = math.cos(yaw), math.sin(yaw)
    cp, sp = math.cos(pitch), math.sin(pitch)
    cr, sr = math.cos(roll), math.sin(roll)
    ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    rz = np.array([[cr, -sr, 0], [sr, cr, 0], [0, 0, 1]])
    return (ry @ rx @ rz).astype(np.float32)

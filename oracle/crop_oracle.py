"""TEST INFRASTRUCTURE ONLY -- CPU restatement of xpano's auto-crop
(xpano/algorithm/auto_crop.cc:118-143 FindLargestCrop, :32-63
FindLongestLineInColumn, :65-113 the seeded expansion; constants.h:89
kAutoCropSamplingDistance = 512). Pinned by the reference's own known-answer
tests (tests/auto_crop_test.cc, incl. the "Real life example" on
tests/data/mask.png) in tests/test_crop.py.

Only tests/ may import this module; the product (xpano_b200) never does."""
from __future__ import annotations

import numpy as np

K_MASK_ON = 255                      # xpano/constants.h kMaskValueOn
K_SAMPLING_DISTANCE = 512            # xpano/constants.h:89


def longest_lines(mask: np.ndarray) -> np.ndarray:
    """(W, 2) int32 [start, end) of the longest run of 255 in every column;
    the FIRST of equally long runs (auto_crop.cc:51-53 replaces only on a
    strictly longer one); columns without a set pixel give (rows, 0), the
    reference's invalid_line."""
    h, w = mask.shape
    on = mask == K_MASK_ON
    lines = np.empty((w, 2), np.int32)
    lines[:, 0] = h
    lines[:, 1] = 0
    best = np.zeros(w, np.int64)
    cur_start = np.zeros(w, np.int64)
    run = np.zeros(w, np.int64)
    for y in range(h):
        row = on[y]
        cur_start = np.where(row & (run == 0), y, cur_start)
        run = np.where(row, run + 1, 0)
        better = run > best
        best = np.where(better, run, best)
        lines[:, 0] = np.where(better, cur_start, lines[:, 0])
        lines[:, 1] = np.where(better, y + 1, lines[:, 1])
    return lines


def _area(r):
    return (r[2] - r[0]) * (r[3] - r[1])


def _expand(lines, invalid, seed):
    """auto_crop.cc:65-113; rect = (x0, y0, x1, y1)."""
    n = len(lines)
    valid = lambda l: l[0] < l[1]  # noqa: E731
    if not valid(lines[seed]):
        return None
    cur = (seed, int(lines[seed][0]), seed + 1, int(lines[seed][1]))
    largest = cur
    if n == 1:
        return largest
    left = seed - 1
    right = seed + n % 2
    ll = tuple(int(v) for v in lines[left])
    rl = tuple(int(v) for v in lines[right])
    while valid(ll) or valid(rl):
        lrect = (left, max(ll[0], cur[1]), cur[2], min(ll[1], cur[3]))
        rrect = (cur[0], max(rl[0], cur[1]), right + 1, min(rl[1], cur[3]))
        if _area(lrect) > _area(rrect):
            cur = lrect
            if left == 0:
                ll = invalid
            else:
                left -= 1
                ll = tuple(int(v) for v in lines[left])
        else:
            cur = rrect
            if right == n - 1:
                rl = invalid
            else:
                right += 1
                rl = tuple(int(v) for v in lines[right])
        if _area(cur) >= _area(largest):
            largest = cur
    return largest


def largest_crop_from_lines(lines, rows):
    """auto_crop.cc:118-143 given the per-column lines."""
    n = len(lines)
    if n == 0 or rows == 0:
        return None
    invalid = (rows, 0)
    best = None
    num_samples = 1 + n // K_SAMPLING_DISTANCE
    for i in range(num_samples):
        start = (i + 1) * n // (num_samples + 1)
        r = _expand(lines, invalid, start)
        if r is not None and (best is None or _area(r) >= _area(best)):
            best = r
    return best


def find_largest_crop(mask: np.ndarray):
    """-> (x0, y0, x1, y1) (RectPPi start / end) or None."""
    if mask is None or mask.size == 0:
        return None
    return largest_crop_from_lines(longest_lines(mask), mask.shape[0])

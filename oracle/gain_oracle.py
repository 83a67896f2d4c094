"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the exposure step of
Stitcher::EstimateSeams (xpano/algorithm/stitcher.cc:273-278):
cv::detail::BlocksGainCompensator::feed / apply, as constructed by
xpano/algorithm/stitcher.cc:200-201 (32x32 blocks, one feed, two gain
filtering iterations, similarity threshold 1).

OpenCV is a third-party dependency of the reference and is not vendored under
/root/reference (CI pins 4.11.0, .github/workflows/test.yml:16); this file
restates the published algorithm (modules/stitching/src/exposure_compensate.cpp:
BlocksCompensator::feed -> GainCompensator::singleFeed -> cv::solve ->
sepFilter2D([.25 .5 .25]) x iterations) and is pinned against cv2 4.13.0:
live in tests/test_oracle_gain.py and through the recorded vectors of
tests/golden/gain_cv2_goldens.npz (tests/golden/make_golden.py gain_goldens).
The product never imports this module.

Pinned details (each was checked against cv2, see the test):
  * block grid: ceil(size / 32) blocks per axis, block size = ceil(size / count)
  * a pair of blocks whose rectangles overlap gets N = max(1, #intersecting
    mask pixels); with no intersecting pixel the pair keeps I = 0 and does not
    un-skip its blocks
  * I(i, j) = mean over the intersection of sqrt(b^2 + g^2 + r^2) (double)
  * blocks that intersect no other block keep gain 1 and leave the system
  * A, b as in singleFeed (alpha 0.01, beta 100), solved in double
  * gain maps are float32, filtered with BORDER_REFLECT_101
"""
import numpy as np

ALPHA = 0.01
BETA = 100.0


def block_grid(w, h, bw=32, bh=32):
    """-> (nx, ny, block_w, block_h) of BlocksCompensator::feed."""
    nx, ny = (w + bw - 1) // bw, (h + bh - 1) // bh
    return nx, ny, (w + nx - 1) // nx, (h + ny - 1) // ny


def make_blocks(corners, sizes, bw=32, bh=32):
    """-> list of (image, x0, y0, x1, y1) in panorama coordinates, raster
    order per image, images in order; and the per-image (nx, ny)."""
    blocks, per = [], []
    for k, ((cx, cy), (w, h)) in enumerate(zip(corners, sizes)):
        nx, ny, sw, sh = block_grid(w, h, bw, bh)
        per.append((nx, ny))
        for by in range(ny):
            for bx in range(nx):
                x0, y0 = bx * sw, by * sh
                blocks.append((k, cx + x0, cy + y0, cx + min(x0 + sw, w),
                               cy + min(y0 + sh, h)))
    return blocks, per


def pair_stats(corners, images, masks, blocks):
    """For every pair i <= j of blocks with overlapping rectangles:
    (i, j, count, sum1, sum2)."""
    norms = [np.sqrt((im.astype(np.int64) ** 2).sum(-1).astype(np.float64))
             for im in images]
    on = [m == 255 for m in masks]
    by_image = {}
    for idx, b in enumerate(blocks):
        by_image.setdefault(b[0], []).append(idx)
    out = []
    n_img = len(images)
    for a in range(n_img):
        for b in range(a, n_img):
            for i in by_image[a]:
                _, ax0, ay0, ax1, ay1 = blocks[i]
                for j in by_image[b]:
                    if j < i:
                        continue
                    _, bx0, by0, bx1, by1 = blocks[j]
                    x0, y0 = max(ax0, bx0), max(ay0, by0)
                    x1, y1 = min(ax1, bx1), min(ay1, by1)
                    if x0 >= x1 or y0 >= y1:
                        continue
                    sa = (slice(y0 - corners[a][1], y1 - corners[a][1]),
                          slice(x0 - corners[a][0], x1 - corners[a][0]))
                    sb = (slice(y0 - corners[b][1], y1 - corners[b][1]),
                          slice(x0 - corners[b][0], x1 - corners[b][0]))
                    inter = on[a][sa] & on[b][sb]
                    cnt = int(inter.sum())
                    out.append((i, j, cnt, float(norms[a][sa][inter].sum()),
                                float(norms[b][sb][inter].sum())))
    out.sort()
    return out


def solve_gains(n_blocks, stats):
    """GainCompensator::singleFeed's linear system -> gains (float64)."""
    N, I = {}, {}
    skip = np.ones(n_blocks, bool)
    for i, j, cnt, s1, s2 in stats:
        N[(i, j)] = N[(j, i)] = max(1, cnt)
        I.setdefault((i, j), 0.0)
        I.setdefault((j, i), 0.0)
        if cnt == 0:
            continue
        if i != j:
            skip[i] = skip[j] = False
        I[(i, j)] = s1 / cnt
        I[(j, i)] = s2 / cnt
    idx = [i for i in range(n_blocks) if not skip[i]]
    pos = {i: k for k, i in enumerate(idx)}
    A = np.zeros((len(idx), len(idx)))
    b = np.zeros(len(idx))
    for (i, j) in sorted(N):
        if skip[i] or skip[j]:
            continue
        nij = N[(i, j)]
        ki, kj = pos[i], pos[j]
        b[ki] += BETA * nij
        A[ki, ki] += BETA * nij
        if i != j:
            A[ki, ki] += 2 * ALPHA * I[(i, j)] * I[(i, j)] * nij
            A[ki, kj] -= 2 * ALPHA * I[(i, j)] * I[(j, i)] * nij
    g = np.ones(n_blocks)
    if idx:
        g[idx] = np.linalg.solve(A, b)
    return g


def reflect101(i, n):
    if n == 1:
        return 0
    while i < 0 or i >= n:
        i = -i if i < 0 else 2 * (n - 1) - i
    return i


def smooth(gm):
    """sepFilter2D(gm, CV_32F, [.25 .5 .25], [.25 .5 .25]) in float32,
    BORDER_REFLECT_101: rows first, then columns, each tap sum evaluated as
    centre * 0.5 + (left + right) * 0.25."""
    gm = np.asarray(gm, np.float32)
    h, w = gm.shape
    k0, k1 = np.float32(0.5), np.float32(0.25)
    xl = [reflect101(x - 1, w) for x in range(w)]
    xr = [reflect101(x + 1, w) for x in range(w)]
    rows = (gm * k0 + (gm[:, xl] + gm[:, xr]) * k1).astype(np.float32)
    yu = [reflect101(y - 1, h) for y in range(h)]
    yd = [reflect101(y + 1, h) for y in range(h)]
    return (rows * k0 + (rows[yu] + rows[yd]) * k1).astype(np.float32)


def blocks_gain_feed(corners, images, masks, bw=32, bh=32, iters=2):
    """-> (gain maps [ny x nx float32 per image], per-block gains float64)."""
    sizes = [(im.shape[1], im.shape[0]) for im in images]
    blocks, per = make_blocks(corners, sizes, bw, bh)
    g = solve_gains(len(blocks), pair_stats(corners, images, masks, blocks))
    maps, k = [], 0
    for nx, ny in per:
        gm = g[k:k + nx * ny].reshape(ny, nx).astype(np.float32)
        k += nx * ny
        for _ in range(iters):
            gm = smooth(gm)
        maps.append(gm)
    return maps, g

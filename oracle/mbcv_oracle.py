"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of xpano's second
blender, blenders::MultiBandOpenCV (xpano/algorithm/blenders.h:41-50,
blenders.cc:160-175; selected by BlendingMethod::kOpenCV, algorithm.cc:135-136):
cv::detail::MultiBandBlender(try_gpu = false, num_bands = 5, weight_type =
CV_32F) fed with CV_16SC3 images, result converted to CV_8U. SURVEY.md 8(f4).

OpenCV is a third-party dependency of the reference, not vendored under
/root/reference (CI pins 4.11.0, .github/workflows/test.yml:16); this file
restates the published algorithm (modules/stitching/src/blenders.cpp:
MultiBandBlender::prepare / feed / blend, createLaplacePyr,
restoreImageFromLaplacePyr, normalizeUsingWeightMap; modules/imgproc/src/
pyramids.cpp: pyrDown / pyrUp for CV_16S and CV_32F) and is pinned against cv2
4.13.0: live in tests/test_oracle_mbcv.py and through the recorded vectors of
tests/golden/mbcv_cv2_goldens.npz (tests/golden/make_golden.py mbcv_goldens).
The product never imports this module.

Pinned exactly (bit for bit against cv2): pyrDown / pyrUp on CV_16SC3
(5-tap [1 4 6 4 1], BORDER_REFLECT_101, (v + 128) >> 8 and (v + 32) >> 6),
the region arithmetic of feed, saturating subtract / add.
Not bit-pinned: pyrDown on the CV_32F weight maps. cv2 sums the five taps in
an order that depends on the SIMD width of its build and on whether a column
is in the vector body, its tail or the border (observed on the 4.13.0 wheel);
this file and the device use the scalar order of pyramids.cpp everywhere. A
weight then differs from cv2's in its last bit (1.2e-7 relative) at some
positions, which moves the truncation short(src * weight) by one in rare
samples: the blended 8-bit panorama agrees with cv2's within 1 LSB on all but
a few pixels in 10^4 (bar in the tests: >= 99.9 % within 1 LSB).
"""
import numpy as np

WEIGHT_EPS = np.float32(1e-5)


def r101(i, n):
    if n == 1:
        return 0
    while i < 0 or i >= n:
        i = -i if i < 0 else 2 * (n - 1) - i
    return i


def pyr_down(src):
    """cv::pyrDown, BORDER_REFLECT_101: int16 (H, W, C) or float32 (H, W)."""
    isf = src.dtype == np.float32
    h, w = src.shape[:2]
    dh, dw = (h + 1) // 2, (w + 1) // 2
    a = src if isf else src.astype(np.int32)

    def cols(off):
        return [r101(2 * x + off, w) for x in range(dw)]

    def rows(off):
        return [r101(2 * y + off, h) for y in range(dh)]
    if isf:
        six, four = np.float32(6), np.float32(4)
        row = a[:, cols(0)] * six + (a[:, cols(-1)] + a[:, cols(1)]) * four \
            + a[:, cols(-2)] + a[:, cols(2)]
        v = row[rows(0)] * six + (row[rows(-1)] + row[rows(1)]) * four \
            + row[rows(-2)] + row[rows(2)]
        return (v * np.float32(1 / 256.)).astype(np.float32)
    row = a[:, cols(0)] * 6 + (a[:, cols(-1)] + a[:, cols(1)]) * 4 \
        + a[:, cols(-2)] + a[:, cols(2)]
    v = row[rows(0)] * 6 + (row[rows(-1)] + row[rows(1)]) * 4 \
        + row[rows(-2)] + row[rows(2)]
    return ((v + 128) >> 8).astype(np.int16)


def pyr_up(src):
    """cv::pyrUp to exactly twice the size, int16 (H, W, C): left / top
    neighbours reflect (101), right / bottom ones replicate."""
    h, w = src.shape[:2]
    a = src.astype(np.int32)
    xl = [r101(x - 1, w) for x in range(w)]
    xr = [min(x + 1, w - 1) for x in range(w)]
    row = np.zeros((h, 2 * w) + a.shape[2:], np.int32)
    row[:, 0::2] = a[:, xl] + a * 6 + a[:, xr]
    row[:, 1::2] = (a + a[:, xr]) * 4
    yu = [r101(y - 1, h) for y in range(h)]
    yd = [min(y + 1, h - 1) for y in range(h)]
    out = np.zeros((2 * h, 2 * w) + a.shape[2:], np.int32)
    out[0::2] = row * 6 + row[yu] + row[yd]
    out[1::2] = (row + row[yd]) * 4
    return ((out + 32) >> 6).astype(np.int16)


def sat16(v):
    return np.clip(v, -32768, 32767).astype(np.int16)


def pad_reflect(a, top, bottom, left, right):
    """copyMakeBorder(..., BORDER_REFLECT): edge pixel repeated (fedcba|abc)."""
    pad = ((top, bottom), (left, right)) + ((0, 0),) * (a.ndim - 2)
    return np.pad(a, pad, mode="symmetric")


class MultiBandBlender:
    """cv::detail::MultiBandBlender(false, num_bands, CV_32F) as
    blenders::MultiBandOpenCV drives it."""

    def __init__(self, num_bands=5):
        self.actual_num_bands = num_bands

    def prepare(self, roi):
        x, y, w, h = roi
        self.final = (x, y, w, h)
        max_len = float(max(w, h))
        nb = min(self.actual_num_bands,
                 int(np.ceil(np.log(max_len) / np.log(2.0))))
        self.nb = nb
        m = 1 << nb
        w += (m - w % m) % m
        h += (m - h % m) % m
        self.roi = (x, y, w, h)
        self.lap = [np.zeros((h, w, 3), np.int16)]
        self.wts = [np.zeros((h, w), np.float32)]
        for _ in range(nb):
            ph, pw = self.lap[-1].shape[:2]
            self.lap.append(np.zeros(((ph + 1) // 2, (pw + 1) // 2, 3),
                                     np.int16))
            self.wts.append(np.zeros(((ph + 1) // 2, (pw + 1) // 2),
                                     np.float32))

    def region(self, tl, size):
        """-> (tl_new, br_new) of MultiBandBlender::feed."""
        nb = self.nb
        rx, ry, rw, rh = self.roi
        gap = 3 * (1 << nb)
        tlx, tly = max(rx, tl[0] - gap), max(ry, tl[1] - gap)
        brx = min(rx + rw, tl[0] + size[0] + gap)
        bry = min(ry + rh, tl[1] + size[1] + gap)
        tlx = rx + (((tlx - rx) >> nb) << nb)
        tly = ry + (((tly - ry) >> nb) << nb)
        m = 1 << nb
        width, height = brx - tlx, bry - tly
        width += (m - width % m) % m
        height += (m - height % m) % m
        brx, bry = tlx + width, tly + height
        dy = max(bry - (ry + rh), 0)
        dx = max(brx - (rx + rw), 0)
        return (tlx - dx, tly - dy), (brx - dx, bry - dy)

    def feed(self, img, mask, tl):
        """img: (h, w, 3) uint8 (xpano converts to CV_16S, blenders.cc:164),
        mask: (h, w) uint8."""
        img = np.asarray(img).astype(np.int16)
        h, w = img.shape[:2]
        (tnx, tny), (bnx, bny) = self.region(tl, (w, h))
        top, left = tl[1] - tny, tl[0] - tnx
        bottom, right = bny - tl[1] - h, bnx - tl[0] - w
        pyr = [pad_reflect(img, top, bottom, left, right)]
        for _ in range(self.nb):
            pyr.append(pyr_down(pyr[-1]))
        for i in range(self.nb):
            up = pyr_up(pyr[i + 1])
            pyr[i] = sat16(pyr[i].astype(np.int32) - up.astype(np.int32))
        wmap = (mask.astype(np.float32) * np.float32(1. / 255.)) \
            .astype(np.float32)
        # convertTo(CV_32F, 1/255): saturate_cast<float>(v * (double)(1/255.))
        wmap = (mask.astype(np.float64) * (1. / 255.)).astype(np.float32)
        wp = [np.pad(wmap, ((top, bottom), (left, right)))]
        for _ in range(self.nb):
            wp.append(pyr_down(wp[-1]))
        rx, ry = self.roi[0], self.roi[1]
        x0, y0, x1, y1 = tnx - rx, tny - ry, bnx - rx, bny - ry
        for i in range(self.nb + 1):
            d = self.lap[i][y0:y1, x0:x1]
            prod = (pyr[i].astype(np.float32) * wp[i][..., None])
            d += prod.astype(np.int16)          # static_cast<short>: truncation
            self.wts[i][y0:y1, x0:x1] += wp[i]
            x0, y0, x1, y1 = x0 // 2, y0 // 2, x1 // 2, y1 // 2

    def blend(self):
        """-> (dst (h, w, 3) uint8, dst_mask (h, w) uint8) as
        MultiBandOpenCV::blend returns them."""
        for i in range(self.nb + 1):
            q = self.lap[i].astype(np.float32) / (self.wts[i]
                                                  + WEIGHT_EPS)[..., None]
            self.lap[i] = q.astype(np.int16)    # truncation
        for i in range(self.nb, 0, -1):
            up = pyr_up(self.lap[i])
            self.lap[i - 1] = sat16(up.astype(np.int32)
                                    + self.lap[i - 1].astype(np.int32))
        _, _, w, h = self.final
        dst = self.lap[0][:h, :w].copy()
        mask = np.where(self.wts[0][:h, :w] > WEIGHT_EPS, 255, 0) \
            .astype(np.uint8)
        dst[mask == 0] = 0
        return np.clip(dst, 0, 255).astype(np.uint8), mask

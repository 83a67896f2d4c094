"""Times the MultiBandOpenCV blender on the cfg2 panorama:
the device implementation (xp_multiband_*, host buffers in and out) next to
cv2.detail_MultiBandBlender on the host cores, same warped images."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import bench  # noqa: E402
import xpano_b200 as xb  # noqa: E402


def main():
    import cv2
    cfg = bench.CONFIGS["cfg2"]
    wl = bench.make_workload(cfg, 0, "graphcut")
    warper = xb.RotationWarper(cfg["projection"], cfg["focal"])
    feeds = []
    for im, K, R in zip(wl["ims"], wl["Ks"], wl["Rs"]):
        c, wi = warper.warp(im, K, R)
        _, wm = warper.warp(np.full(im.shape[:2], 255, np.uint8), K, R,
                            xb.INTER_NEAREST, xb.BORDER_CONSTANT)
        feeds.append((wi, wm, c))
    x0 = min(f[2][0] for f in feeds)
    y0 = min(f[2][1] for f in feeds)
    x1 = max(f[2][0] + f[0].shape[1] for f in feeds)
    y1 = max(f[2][1] + f[0].shape[0] for f in feeds)
    roi = (x0, y0, x1 - x0, y1 - y0)
    mp = roi[2] * roi[3] / 1e6

    def device():
        b = xb.MultiBandOpenCV()
        b.prepare(roi)
        for f in feeds:
            b.feed(*f)
        return b.blend()
    device()
    t0 = time.perf_counter()
    for _ in range(3):
        dst, mask = device()
    t_dev = (time.perf_counter() - t0) / 3

    t0 = time.perf_counter()
    b = cv2.detail_MultiBandBlender(0, 5, cv2.CV_32F)
    b.prepare(roi)
    for img, m, tl in feeds:
        b.feed(img.astype(np.int16), m, tl)
    ref, rmask = b.blend(None, None)
    ref = np.clip(ref, 0, 255).astype(np.uint8)
    t_cpu = time.perf_counter() - t0
    d = np.abs(dst.astype(int) - ref.astype(int))
    print(f"MultiBandOpenCV on {roi[2]}x{roi[3]} ({mp:.1f} MP, 8 images): "
          f"device {1e3 * t_dev:.1f} ms through host buffers "
          f"({mp / t_dev:.0f} MP/s), cv2 on {os.cpu_count()} cores "
          f"{t_cpu:.2f} s ({mp / t_cpu:.1f} MP/s); mask equal "
          f"{np.array_equal(mask, rmask)}, max diff {int(d.max())}, "
          f"differing pixels {float((d > 0).any(-1).mean()):.2e}")


if __name__ == "__main__":
    main()

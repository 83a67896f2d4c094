"""One 12-MP source (cfg2's first image, 4:2:2, quality 95) through the SM JPEG
decoder a few times: wall time per decode; run under
`ncu --metrics gpu__time_duration.sum` for the per-kernel times."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cv2  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import xpano_b200 as xb  # noqa: E402


def main():
    cfg = bench.CONFIGS["cfg2"]
    wl = bench.make_workload(cfg, 0, "none")
    im = wl["ims"][0]
    sub = {"444": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444,
           "422": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422,
           "420": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420}[
               os.environ.get("XP_PROBE_SUB", "422")]
    buf = cv2.imencode(".jpg", im, [cv2.IMWRITE_JPEG_QUALITY, 95,
                                    cv2.IMWRITE_JPEG_SAMPLING_FACTOR, sub])[1]
    codec = xb.JpegCodec()
    h, w = im.shape[:2]
    dev = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
    n = int(os.environ.get("XP_PROBE_N", "5"))
    codec.decode_device(buf, dev.data_ptr())
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        codec.decode_device(buf, dev.data_ptr())
    dt = (time.perf_counter() - t0) / n
    print(f"{w}x{h}, {buf.size / 1e6:.2f} MB: {1e3 * dt:.2f} ms per decode "
          f"(one stream, host steps included); stats {codec.stats()}")


if __name__ == "__main__":
    main()

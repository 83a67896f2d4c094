"""The cfg2 panorama-sized image (20534 x 2962, synthetic) through the SM JPEG
encoder from device memory a few times: wall time per encode; run under
`ncu --metrics gpu__time_duration.sum` for the per-kernel times."""
import ctypes
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import xpano_b200 as xb  # noqa: E402
from xpano_b200 import _lib  # noqa: E402


def main():
    w, h = 20534, 2962
    tile = bench.make_image(4000, 3000, 2000, 8)
    img = np.tile(tile, (1, 6, 1))[:h, :w]
    dev = torch.from_numpy(np.ascontiguousarray(img)).cuda()
    codec = xb.JpegCodec()
    lib = _lib.load()
    n = ctypes.c_size_t()
    out = np.empty(w * h, np.uint8)
    reps = int(os.environ.get("XP_PROBE_N", "5"))

    def once():
        _lib.check(lib.xp_codec_jpeg_encode(
            codec._h, ctypes.c_void_p(dev.data_ptr()), w, h, w * 3,
            _lib.MEM_DEVICE, 95, 422, 0, 0, out.ctypes.data, out.size,
            ctypes.byref(n)))
    once()
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    dt = (time.perf_counter() - t0) / reps
    print(f"{w}x{h}: {1e3 * dt:.2f} ms per encode -> {n.value / 1e6:.1f} MB; "
          f"stats {codec.all_stats()}")


if __name__ == "__main__":
    main()

"""Times the JPEG steps either side of the path on the cfg2
workload: eight 12-MP sources decoded on the device, the 60-MP panorama
encoded on the device, next to cv2 (libjpeg-turbo) on the host cores; and the
whole compressed-in / compressed-out panorama."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import xpano_b200 as xb  # noqa: E402


def main():
    import cv2
    cfg = bench.CONFIGS["cfg2"]
    wl = bench.make_workload(cfg, 0, "graphcut")
    params = [cv2.IMWRITE_JPEG_QUALITY, 95, cv2.IMWRITE_JPEG_SAMPLING_FACTOR,
              cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422]
    jpegs = [cv2.imencode(".jpg", im, params)[1] for im in wl["ims"]]
    mb = sum(j.size for j in jpegs) / 1e6
    h, w = wl["ims"][0].shape[:2]
    codec = xb.JpegCodec()
    dev = [torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
           for _ in jpegs]

    def decode_all():
        if os.environ.get("XP_DECODE_ONE_BY_ONE"):
            for j, t in zip(jpegs, dev):
                codec.decode_device(j, t.data_ptr())
        else:
            codec.decode_batch_device(jpegs, [t.data_ptr() for t in dev])
    decode_all()
    t0 = time.perf_counter()
    for _ in range(3):
        decode_all()
    t_dec = (time.perf_counter() - t0) / 3
    t0 = time.perf_counter()
    ref = [cv2.imdecode(j, cv2.IMREAD_COLOR) for j in jpegs]
    t_dec_cpu = time.perf_counter() - t0
    d = np.abs(dev[0].cpu().numpy().astype(int) - ref[0].astype(int))

    comp = xb.Composer(projection=cfg["projection"], warp_scale=cfg["focal"],
                       max_pano_mpx=cfg["max_pano_mpx"], depth=8,
                       fixed_levels=cfg["fixed_levels"])

    def whole():
        decode_all()
        comp.set_images_device([t.data_ptr() for t in dev], [(w, h)] * len(dev),
                               wl["Ks"], wl["Rs"], wl["gains"], wl["seams"])
        info = comp.run()
        return info, comp.encode_jpeg(codec)
    info, out = whole()
    t0 = time.perf_counter()
    for _ in range(3):
        info, out = whole()
    t_all = (time.perf_counter() - t0) / 3
    t0 = time.perf_counter()
    for _ in range(3):
        out = comp.encode_jpeg(codec)
    t_enc = (time.perf_counter() - t0) / 3
    pano, _ = comp.fetch()
    t0 = time.perf_counter()
    ok, cpu_out = cv2.imencode(".jpg", pano, params)
    t_enc_cpu = time.perf_counter() - t0
    mp = info["width"] * info["height"] / 1e6
    print(f"sources: 8 x {w}x{h} JPEG, {mb:.1f} MB in all (raw: "
          f"{8 * w * h * 3 / 1e6:.0f} MB); decode on the device "
          f"{1e3 * t_dec:.1f} ms ({8 * w * h / 1e6 / t_dec:.0f} MP/s), cv2 "
          f"{1e3 * t_dec_cpu:.0f} ms; max |diff| {int(d.max())}, mean "
          f"{float(d.mean()):.3f}")
    print(f"panorama {info['width']}x{info['height']} ({mp:.1f} MP): encode on "
          f"the device {1e3 * t_enc:.1f} ms -> {len(out) / 1e6:.1f} MB, cv2 "
          f"{1e3 * t_enc_cpu:.0f} ms -> {cpu_out.size / 1e6:.1f} MB")
    print(f"compressed in -> compressed out: {1e3 * t_all:.1f} ms per panorama "
          f"({mp / t_all:.0f} MP/s), {mb:.1f} MB up, {len(out) / 1e6:.1f} MB down")


if __name__ == "__main__":
    main()

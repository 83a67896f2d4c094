"""Development probe: which stage of run() stretches while pinned-memory
copies are in flight on other streams."""
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
import xpano_b200 as xb  # noqa: E402


def main():
    cfg = bench.CONFIGS["cfg2"]
    wl = bench.make_workload(cfg, 0, "graphcut")
    keep = bench.pin_images(torch, wl["ims"])
    host = [t.numpy() for t in keep]
    comp = xb.Composer(projection=cfg["projection"], warp_scale=cfg["focal"],
                       max_pano_mpx=cfg["max_pano_mpx"], depth=cfg["bpp"],
                       fixed_levels=cfg["fixed_levels"], device=0)
    comp.set_images(host, wl["Ks"], wl["Rs"], wl["gains"], wl["seams"])
    for _ in range(3):
        comp.run()
    n = 256 << 20
    hin = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    hout = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    din = torch.empty(n, dtype=torch.uint8, device="cuda")
    dout = torch.empty(n, dtype=torch.uint8, device="cuda")
    stop = threading.Event()

    def bg(up, down):
        torch.cuda.set_device(0)
        s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
        while not stop.is_set():
            if up:
                with torch.cuda.stream(s1):
                    din.copy_(hin, non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    hout.copy_(dout, non_blocking=True)
            s1.synchronize()
            s2.synchronize()

    for name, up, down in (("none", 0, 0), ("H2D", 1, 0), ("D2H", 0, 1),
                           ("both", 1, 1)):
        stop.clear()
        t = None
        if up or down:
            t = threading.Thread(target=bg, args=(up, down))
            t.start()
            time.sleep(0.05)
        acc = {}
        t0 = time.perf_counter()
        R = 10
        for _ in range(R):
            info = comp.run()
            for k, v in info["timing"].items():
                acc[k] = acc.get(k, 0.0) + v / R
            acc["warp"] = acc.get("warp", 0.0) + info["warp_ms"] / R
        wall = 1e3 * (time.perf_counter() - t0) / R
        stop.set()
        if t:
            t.join()
        print(f"{name:5s} wall {wall:6.2f} ms/run  " +
              " ".join(f"{k}={v:.2f}" for k, v in acc.items()))


if __name__ == "__main__":
    main()

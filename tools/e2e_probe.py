"""Development probe: (1) pinned-memory copy rates of the box, each direction
alone and both at once; (2) per-phase wall times of the pipelined end-to-end
steps of bench.py (upload / wait for the compute turn / run / fetch)."""
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import xpano_b200 as xb  # noqa: E402


SPLIT = False


def copy_rates():
    n = 256 << 20
    hin = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    hout = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    din = torch.empty(n, dtype=torch.uint8, device="cuda")
    dout = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(up, down, reps=8):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            if up:
                with torch.cuda.stream(s1):
                    din.copy_(hin, non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    hout.copy_(dout, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        return reps * n / dt / 1e9
    run(True, True, 2)
    print(f"pinned copies: H2D alone {run(True, False):.1f} GB/s, "
          f"D2H alone {run(False, True):.1f} GB/s, both at once "
          f"{run(True, True):.1f} GB/s each")


def main():
    NF = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    global SPLIT
    SPLIT = len(sys.argv) > 2 and sys.argv[2] == "split"
    copy_rates()
    cfg = bench.CONFIGS["cfg2"]
    wl = bench.make_workload(cfg, 0, "graphcut")
    keep = bench.pin_images(torch, wl["ims"])
    host = [t.numpy() for t in keep]
    Ks, Rs = wl["Ks"], wl["Rs"]
    mk = lambda: xb.Composer(projection=cfg["projection"],  # noqa: E731
                             warp_scale=cfg["focal"],
                             max_pano_mpx=cfg["max_pano_mpx"],
                             depth=cfg["bpp"],
                             fixed_levels=cfg["fixed_levels"], device=0)
    comps = [mk() for _ in range(NF)]
    comps[0].set_images(host, Ks, Rs, wl["gains"], wl["seams"])
    info = comps[0].run()
    W, H = info["width"], info["height"]
    outs = [(torch.empty((H, W, 3), dtype=torch.uint8, pin_memory=True),
             torch.empty((H, W), dtype=torch.uint8, pin_memory=True))
            for _ in range(NF)]
    log = []
    turn = threading.Lock()
    STEPS = 4

    def worker(j):
        for k in range(STEPS):
            a = time.perf_counter()
            comps[j].set_images(host, Ks, Rs, wl["gains"], wl["seams"])
            b = time.perf_counter()
            with turn:
                b2 = time.perf_counter()
                if SPLIT:
                    comps[j].run_begin()
                else:
                    inf = comps[j].run()
                c = time.perf_counter()
            if SPLIT:
                inf = comps[j].run_end()
            comps[j].fetch(outs[j][0].numpy(), outs[j][1].numpy())
            d = time.perf_counter()
            tm = inf["timing"]
            log.append((j, k, a, b, c, d, b2, inf["warp_ms"], tm["total"],
                        tm["seam"], tm["pyramid_tail"]))
    for rep in range(2):
        log.clear()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ts = [threading.Thread(target=worker, args=(j,)) for j in range(NF)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        torch.cuda.synchronize()
        t1 = time.perf_counter()
    print("in flight", NF, "ms per pano", 1e3 * (t1 - t0) / (STEPS * NF))
    for j, k, a, b, c, d, b2, wm, tot, seam, tail in sorted(
            log, key=lambda r: r[2]):
        print(f"thr {j} step {k}: start {1e3 * (a - t0):7.2f} upload "
              f"{1e3 * (b - a):6.2f} wait {1e3 * (b2 - b):6.2f} run "
              f"{1e3 * (c - b2):6.2f} (device: warp {wm:.2f} + blend "
              f"{tot:.2f}, seam {seam:.2f}, tail {tail:.2f}) fetch "
              f"{1e3 * (d - c):6.2f}")
    # each phase alone
    for name, fn in (("upload", lambda: comps[0].set_images(
            host, Ks, Rs, wl["gains"], wl["seams"])),
                     ("run", comps[0].run),
                     ("fetch", lambda: comps[0].fetch(outs[0][0].numpy(),
                                                      outs[0][1].numpy()))):
        fn()
        t0 = time.perf_counter()
        for _ in range(5):
            fn()
        print(f"{name} alone: {1e3 * (time.perf_counter() - t0) / 5:.2f} ms")


if __name__ == "__main__":
    main()

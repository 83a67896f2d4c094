"""Development probe: per-phase wall times of pipelined end-to-end steps."""
import sys, os, time, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, xpano_b200 as xb

cfg = bench.CONFIGS["cfg2"]
ims, Ks, Rs = bench.make_workload(cfg, 0)
pinned = []
for im in ims:
    t = torch.empty(im.shape, dtype=torch.uint8, pin_memory=True); t.numpy()[...] = im; pinned.append(t)
host = [t.numpy() for t in pinned]
NF = int(sys.argv[1]) if len(sys.argv) > 1 else 2
comps = [xb.Composer(projection=cfg["projection"], warp_scale=cfg["focal"], max_pano_mpx=cfg["max_pano_mpx"],
                     depth=cfg["bpp"], fixed_levels=cfg["fixed_levels"], device=0) for _ in range(NF)]
comps[0].set_images(host, Ks, Rs); info = comps[0].run()
W, H = info["width"], info["height"]
outs = [(torch.empty((H, W, 3), dtype=torch.uint8, pin_memory=True), torch.empty((H, W), dtype=torch.uint8, pin_memory=True)) for _ in range(NF)]
log = []
turn = threading.Lock()
def worker(j, n):
    for k in range(n):
        a = time.perf_counter(); comps[j].set_images(host, Ks, Rs)
        b = time.perf_counter()
        with turn:
            b2 = time.perf_counter(); comps[j].run()
        c = time.perf_counter(); comps[j].fetch(outs[j][0].numpy(), outs[j][1].numpy())
        d = time.perf_counter(); log.append((j, k, a, b, c, d, b2))
for rep in range(2):
    log.clear()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    ts = [threading.Thread(target=worker, args=(j, 4)) for j in range(NF)]
    [t.start() for t in ts]; [t.join() for t in ts]
    torch.cuda.synchronize(); t1 = time.perf_counter()
print("in flight", NF, "ms per pano", 1e3 * (t1 - t0) / (4 * NF))
for j, k, a, b, c, d, b2 in sorted(log, key=lambda r: r[2]):
    print(f"thr {j} step {k}: start {1e3*(a-t0):7.2f} upload {1e3*(b-a):6.2f} wait {1e3*(b2-b):6.2f} run {1e3*(c-b2):6.2f} fetch {1e3*(d-c):6.2f}")

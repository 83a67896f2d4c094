"""Development probe: aggregate pinned host-to-device rate of several GPUs of
one box driven from one process (which links share an uplink, where the host
side saturates). Run under `gpurun --gpus 8`."""
import time

import torch

n = 256 << 20
bufs = {}
for dev in range(torch.cuda.device_count()):
    torch.cuda.set_device(dev)
    h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    d = torch.empty(n, dtype=torch.uint8, device=f"cuda:{dev}")
    d.copy_(h, non_blocking=True)
    torch.cuda.synchronize(dev)
    bufs[dev] = (h, d)


def run(devs, reps=6):
    t0 = time.perf_counter()
    for _ in range(reps):
        for dev in devs:
            h, d = bufs[dev]
            with torch.cuda.device(dev):
                d.copy_(h, non_blocking=True)
    for dev in devs:
        torch.cuda.synchronize(dev)
    return reps * len(devs) * n / (time.perf_counter() - t0) / 1e9


for devs in ([0], [0, 1], [0, 1, 2, 3], list(range(8)), [0, 4], [0, 2, 4, 6]):
    if max(devs) < len(bufs):
        print(devs, round(run(devs), 1), "GB/s aggregate H2D")

"""Development probe: latency of a host<->GPU round trip (null kernel + sync,
event record + sync) on an idle stream while pinned-memory copies of various
chunk sizes are in flight on other streams."""
import threading
import time

import torch

n = 256 << 20
hin = torch.empty(n, dtype=torch.uint8, pin_memory=True)
hout = torch.empty(n, dtype=torch.uint8, pin_memory=True)
din = torch.empty(n, dtype=torch.uint8, device="cuda")
dout = torch.empty(n, dtype=torch.uint8, device="cuda")
x = torch.zeros(32, device="cuda")
stop = threading.Event()


def bg(up, down, chunk):
    torch.cuda.set_device(0)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    while not stop.is_set():
        for o in range(0, n, chunk):
            if up:
                with torch.cuda.stream(s1):
                    din[o:o + chunk].copy_(hin[o:o + chunk], non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    hout[o:o + chunk].copy_(dout[o:o + chunk],
                                            non_blocking=True)
        s1.synchronize()
        s2.synchronize()


def measure():
    s = torch.cuda.Stream()
    ev = torch.cuda.Event()
    a, b = [], []
    for _ in range(200):
        t0 = time.perf_counter()
        with torch.cuda.stream(s):
            ev.record()
        s.synchronize()
        a.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        with torch.cuda.stream(s):
            x.add_(1)
        s.synchronize()
        b.append(time.perf_counter() - t0)
        time.sleep(0.0005)
    f = lambda v: (1e6 * sorted(v)[len(v) // 2], 1e6 * sorted(v)[int(len(v) * 0.9)], 1e6 * max(v))  # noqa
    return f(a), f(b)


for name, up, down, chunk in (("none", 0, 0, n), ("H2D 256M", 1, 0, n),
                              ("D2H 256M", 0, 1, n), ("both 256M", 1, 1, n),
                              ("both 32M", 1, 1, 32 << 20),
                              ("both 4M", 1, 1, 4 << 20),
                              ("both 1M", 1, 1, 1 << 20)):
    stop.clear()
    t = None
    if up or down:
        t = threading.Thread(target=bg, args=(up, down, chunk))
        t.start()
        time.sleep(0.05)
    ea, kb = measure()
    stop.set()
    if t:
        t.join()
    print(f"{name:10s} event+sync us (median, p90, max) = "
          f"{ea[0]:7.0f} {ea[1]:7.0f} {ea[2]:7.0f}   kernel+sync = "
          f"{kb[0]:7.0f} {kb[1]:7.0f} {kb[2]:7.0f}")

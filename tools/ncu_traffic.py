#!/usr/bin/env python3
"""DRAM traffic of the pyramid/blend group from an ncu metrics capture.

usage: tools/ncu_traffic.py capture.csv "<workload name>" [profiles/traffic.json]

capture.csv comes from
    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
        --clock-control none --csv --log-file capture.csv python bench.py ...
Sums dram read + write bytes over the kernels of the group (SURVEY.md 8d,
K9-K16: image/mask pyramids, composite, collapse, offset + out), divides by
the number of steps in the capture (one out_kernel launch per step) and
records the per-step figure bench.py reports as `roofline.traffic`, with the
per-kernel breakdown next to it.
"""
import collections
import csv
import json
import os
import re
import sys

GROUP = ("shrink_tiled_kernel", "shrink_batch_kernel", "mask_shrink", "lapcomp_kernel",
         "composite_kernel", "pad_kernel", "laplace_kernel",
         "xor_sum", "offset_kernel", "out_kernel")
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9,
        "ns": 1.0, "us": 1e3, "ms": 1e6}


def main():
    path, workload = sys.argv[1], sys.argv[2]
    out = sys.argv[3] if len(sys.argv) > 3 else os.path.join(
        os.path.dirname(os.path.abspath(__file__)), "..", "profiles",
        "traffic.json")
    rows = [r for r in csv.reader(open(path))
            if len(r) >= 15 and r[0].isdigit()]
    short = lambda r: re.sub(r"^void ", "", re.sub(r"[<(].*", "", r[4]))
    # whole steps only: a step ends with its out_kernel launch
    ends = sorted({int(r[0]) for r in rows if short(r) == "out_kernel"})
    last = ends[-1] if ends else -1
    steps = len(ends)
    per = collections.OrderedDict()
    for r in rows:
        if int(r[0]) > last:
            continue
        name = short(r)
        metric, unit, val = r[12], r[13], float(r[14].replace(",", ""))
        if not name.startswith(GROUP):
            continue
        k = per.setdefault(name, {"launches": set(), "dram_bytes": 0.0,
                                  "time_ns": 0.0})
        k["launches"].add(r[0])
        if metric.startswith("dram__bytes"):
            k["dram_bytes"] += val * UNIT.get(unit, 1.0)
        elif metric.startswith("gpu__time_duration"):
            k["time_ns"] += val * UNIT.get(unit, 1.0)
    steps = max(steps, 1)
    rows = {n: {"launches_per_step": len(k["launches"]) / steps,
                "dram_bytes_per_step": k["dram_bytes"] / steps,
                "us_per_step": k["time_ns"] / steps / 1e3}
            for n, k in per.items()}
    total = sum(v["dram_bytes_per_step"] for v in rows.values())
    try:
        with open(out) as f:
            doc = json.load(f)
    except Exception:
        doc = {}
    doc[workload] = {"blend_group_dram_bytes_per_step": total,
                     "steps_in_capture": steps, "source": os.path.basename(path),
                     "kernels": rows}
    with open(out, "w") as f:
        json.dump(doc, f, indent=1)
    print(f"{workload}: {total / 1e9:.3f} GB of DRAM traffic per step over "
          f"{steps} steps")
    for n, v in rows.items():
        print(f"  {n:28s} {v['launches_per_step']:5.1f} launches "
              f"{v['dram_bytes_per_step'] / 1e6:9.1f} MB {v['us_per_step']:8.1f} us")


if __name__ == "__main__":
    main()

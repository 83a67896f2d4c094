#!/usr/bin/env python3
"""Per-kernel totals of an ncu launch list (gpu__time_duration.sum CSV).
usage: tools/launch_summary.py launches.csv [steps]
(without steps: the list is cut after the last out_kernel and divided by the
number of out_kernel launches, i.e. whole steps only)"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
steps = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
if steps == 0.0:
    # whole steps only: a step ends with its out_kernel launch
    ends = [k for k, r in enumerate(rows) if "out_kernel" in r[4]]
    if ends:
        rows = rows[:ends[-1] + 1]
        steps = float(len(ends))
    else:
        steps = 1.0
agg = collections.OrderedDict()
for r in rows:
    name = re.sub(r"^void ", "", re.sub(r"\(.*", "", r[4]))
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += float(r[-1]) / 1e3
tot = sum(v[1] for v in agg.values())
print(f"{len(rows)} launches, {tot / steps:.1f} us per step ({steps:g} steps)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[1] / steps:10.1f} us {v[0] / steps:7.1f}  {k[:100]}")

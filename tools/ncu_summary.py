#!/usr/bin/env python3
"""Summarise an .ncu-rep (ncu --set full) into a small markdown table.

usage: tools/ncu_summary.py report.ncu-rep > profiles/rN_<name>.md
Reads the report with `ncu -i ... --page raw --csv`; one row per captured
launch with the counters the roofline discussion in DESIGN.md uses.
"""
import csv
import io
import subprocess
import sys

COLS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram rd"),
    ("dram__bytes_write.sum", "dram wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("smsp__inst_executed.sum", "warp insts"),
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"],
                         check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print(f"# ncu --set full summary of `{rep.split('/')[-1]}`\n")
    print("| # | kernel | " + " | ".join(n for _, n in COLS) + " |")
    print("|---|---|" + "---|" * len(COLS))
    for k, r in enumerate(rows[2:]):
        name = r[idx["Kernel Name"]].split("(")[0]
        cells = []
        for key, _ in COLS:
            if key not in idx:
                cells.append("-")
                continue
            v, u = r[idx[key]], units[idx[key]]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:.4g}"
            except ValueError:
                pass
            cells.append(f"{v} {u}".strip())
        print(f"| {k} | `{name}` | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main()

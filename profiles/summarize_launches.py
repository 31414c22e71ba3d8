#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and mean time.
usage: python profiles/summarize_launches.py profiles/rNN_launches.csv [skip_first_n_launches]"""
import csv
import re
import sys
from collections import OrderedDict


def main():
    path = sys.argv[1]
    skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        rows.append((int(r["ID"]), r["Kernel Name"], r["Grid Size"], r["Block Size"], float(r["Metric Value"])))
    rows = [r for r in rows if r[0] >= skip]
    agg = OrderedDict()
    for _, name, grid, block, ns in rows:
        short = re.sub(r"\(.*", "", name)
        short = re.sub(r"^void ", "", short)
        if len(short) > 90:
            short = short[:87] + "..."
        a = agg.setdefault(short, [0, 0.0, grid, block])
        a[0] += 1
        a[1] += ns
    total = sum(a[1] for a in agg.values())
    print(f"| kernel | launches | total us | mean us | share | grid | block |")
    print(f"|---|---:|---:|---:|---:|---|---|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {a[0]} | {a[1] / 1e3:.1f} | {a[1] / a[0] / 1e3:.2f} | {100 * a[1] / total:.1f}% | {a[2]} | {a[3]} |")
    print(f"\ntotal {total / 1e3:.1f} us over {len(rows)} launches")


if __name__ == "__main__":
    main()

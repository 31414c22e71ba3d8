#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv` (one kernel): stall-reason totals and the hottest SASS lines.
usage: python profiles/stall_summary.py source_page.csv [top_n] [kernel_index | all out_prefix]
With `all`, one summary per captured launch is written to <out_prefix><index>_<kernel>.txt."""
import contextlib
import csv
import re
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    if len(sys.argv) > 3 and sys.argv[3] == "all":
        for which, st in enumerate(starts):
            name = re.sub(r"[^A-Za-z0-9_]+", "_", re.sub(r"[<(].*", "", rows[st][1].replace("void ", "")))[:40]
            with open(f"{sys.argv[4]}{which:02d}_{name}.txt", "w") as f, contextlib.redirect_stdout(f):
                summarize(rows[st:], topn)
        return
    which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    summarize(rows[starts[which]:], topn)


def summarize(rows, topn):
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    ix = {n: i for i, n in enumerate(hdr)}
    end = next((i for i in range(h + 1, len(rows)) if rows[i] and rows[i][0] == "Kernel Name"), len(rows))
    data = [r for r in rows[h + 1:end] if len(r) == len(hdr)]   # this kernel only
    print(rows[0][:2])
    S = ix["# Samples"]
    tot = sum(int(r[S]) for r in data)
    print("total samples", tot, "| SASS instructions", len(data), "| warp instructions executed",
          sum(int(r[ix["Instructions Executed"]]) for r in data))
    stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
    agg = {s: sum(int(r[ix[s]]) for r in data) for s in stalls}
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]:
        print(f"  {k:28s} {v:8d} {100 * v / max(tot, 1):5.1f}%")
    top = sorted(enumerate(data), key=lambda kv: -int(kv[1][S]))[:topn]
    print("hottest SASS lines (index, samples, executed, instruction, top stalls)")
    for i, r in sorted(top):
        st = sorted(((s, int(r[ix[s]])) for s in stalls if int(r[ix[s]]) > 0), key=lambda kv: -kv[1])[:3]
        print(f"  {i:5d} {r[S]:>6s} {r[ix['Instructions Executed']]:>8s}  {r[ix['Source']].strip()[:72]:72s} {st}")


if __name__ == "__main__":
    main()

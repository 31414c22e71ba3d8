#!/usr/bin/env python
"""Pull the metrics the write-up cites out of `ncu -i X.ncu-rep --page raw --csv` and append them to a long-format CSV.
usage: ncu -i X.ncu-rep --page raw --csv | python profiles/extract_metrics.py "<capture label>" >> profiles/rNN_ncu_full_metrics.csv"""
import csv
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def main():
    label = sys.argv[1] if len(sys.argv) > 1 else "capture"
    rows = list(csv.reader(sys.stdin))
    hdr, units = rows[0], rows[1]
    ix = {n: i for i, n in enumerate(hdr)}
    w = csv.writer(sys.stdout)
    for r in rows[2:]:
        if len(r) != len(hdr):
            continue
        name = r[ix["Kernel Name"]]
        for m in WANT:
            if m in ix:
                w.writerow([label, name, m, units[ix[m]], r[ix[m]]])


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""From the raw page of ONE `ncu --set full` pass over `profiles/tools/prof_all.py all`: the warm (last) instance of every
kernel per workload -> profiles/rNN_ncu_full_metrics.csv, and DRAM bytes per launch -> profiles/rNN_traffic.json.
usage: python profiles/tools/extract_prof_all.py raw.csv profiles/r02"""
import csv, json, re, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from extract_metrics import WANT

raw, prefix = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(raw)))
hdr, units = rows[0], rows[1]
ix = {n: i for i, n in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) == len(hdr)]
names = [re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "") for r in data]
# workloads in the order prof_all.py runs them; a workload starts at its first characteristic kernel
bounds, labels = [], ["C2 fp32 NCHW maps, F16X3", "C2 fp16 NHWC maps, F16X3", "configs[4] shape (B=32,P=6840,K=171) fp32 maps, F16X3",
                      "push scan, 16 maps 256x129x257 fp32 NCHW", "push scan, 16 maps fp16 NHWC",
                      "stage 2 (C2 shape, dense head fwd + grouping fwd/bwd)", "eval epilogue, 760x129x257 -> 1024x2048"]
def first(pred, start=0):
    return next(i for i in range(start, len(names)) if pred(names[i]))
b0 = 0
b1 = first(lambda n: "head_fwd_tma_kernel" in n and n.rstrip(">").endswith(", 0"), 0)   # 16-bit maps: F32 flag 0 (or absent)
b2 = first(lambda n: "z_pack_kernel" in n, b1)
b3 = first(lambda n: "push_scan_tc_kernel" in n, b2)
b4 = first(lambda n: "push_scan_tc_kernel" in n and n.rstrip(">").endswith(", 1"), b3)
b5 = first(lambda n: "group_fwd_kernel" in n, b4) - 1
b6 = first(lambda n: "upsample_argext" in n, b5)
starts = [b0, b1, b2, b3, b4, b5, b6, len(names)]
def label(i):
    return next(labels[k] for k in range(7) if starts[k] <= i < starts[k + 1])
last = {}
for i, r in enumerate(data):
    last[(label(i), names[i])] = (i, r)
extra = ["sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
         "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.max", "smsp__cycles_active.avg"]
with open(prefix + "_ncu_full_metrics.csv", "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["capture", "launch_id", "kernel", "metric", "unit", "value"])
    for (lab, name), (i, r) in last.items():
        for m in WANT + extra:
            if m in ix:
                w.writerow([lab, r[ix["ID"]], name.replace("<unnamed>::", ""), m, units[ix[m]], r[ix[m]]])
assert units[ix["dram__bytes_read.sum"]] == "Mbyte", units[ix["dram__bytes_read.sum"]]
def dram(lab, key):
    i, r = next(v for (l, n), v in last.items() if l == lab and key in n)
    return int(round((float(r[ix["dram__bytes_read.sum"]]) + float(r[ix["dram__bytes_write.sum"]])) * 1e6))
t = {"backward": dram(labels[0], "head_bwd_tc_kernel"), "forward": dram(labels[0], "head_fwd_"),
     "backward_fp16_nhwc": dram(labels[1], "head_bwd_tc_kernel"), "forward_fp16_nhwc": dram(labels[1], "head_fwd_tma_kernel"),
     "coco_forward": dram(labels[2], "head_fwd_tma_kernel"), "coco_wide_dz": dram(labels[2], "wide_dz_kernel"),
     "coco_wide_dpw": dram(labels[2], "wide_dpw_kernel"), "push_16_maps": dram(labels[3], "push_scan_tc_kernel"),
     "push_16_maps_fp16_nhwc": dram(labels[4], "push_scan_tc_kernel"),
     "note": "dram__bytes_read.sum + dram__bytes_write.sum per launch (warm instance), one ncu --set full pass over "
             "profiles/tools/prof_all.py all (profiles/r02_ncu_full_metrics.csv); writes are understated where the output still sits in "
             "the 126 MB L2 when the kernel ends (the C2 backward writes 69 MB of dz, 27 MB reach DRAM inside the kernel); reads: z once "
             "(69.2 MB fp32 / 34.6 MB fp16) + dlogits"}
json.dump(t, open(prefix + "_traffic.json", "w"), indent=1)
print(t)
for (lab, name), (i, r) in last.items():
    print(f"{lab[:40]:40s} {name[:58]:58s} {float(r[ix['gpu__time_duration.sum']]) * 1e3:8.1f} us  tensor {r[ix['sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed']][:5]:>5s}  "
          f"issue {r[ix['smsp__issue_active.avg.pct_of_peak_sustained_active']][:5]:>5s}  xu {r[ix['sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active']][:5]:>5s}  "
          f"regs {r[ix['launch__registers_per_thread']]}  smem {r[ix['launch__shared_mem_per_block_dynamic']][:6]}")

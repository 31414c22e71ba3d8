"""Reads the clock stamps of CTA 0 of head_fwd_tma_kernel (debug build -DSPS_TIMELINE, SCALEPROTO_B200_LIB pointing at it)
and prints, per role, the mean time between events.  usage: python profiles/tools/timeline_tma.py [f16|f16x3]"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
import scaleprotoseg_b200 as sps
from scaleprotoseg_b200 import _lib
from oracle import head_oracle as ho

dev = "cuda:0"
math = {"f16": sps.MATH_F16, "f16x3": sps.MATH_F16X3}[sys.argv[1] if len(sys.argv) > 1 else "f16"]
B, H, W, P, K = 16, 65, 65, 760, 19
g = torch.Generator(device=dev).manual_seed(1)
z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev, generator=g)).half().contiguous(memory_format=torch.channels_last)
protos = torch.rand(P, 64, device=dev, generator=g)
W_ = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous().to(dev)
cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=math)
for _ in range(3):
    sps.head_forward_raw(z, protos, W_, cfg, False, False)
torch.cuda.synchronize()
lib = _lib.load()
buf = (C.c_longlong * (64 * 1024))()
lib.sps_debug_timeline2.argtypes = [C.c_void_p, C.c_int]
assert lib.sps_debug_timeline2(buf, 64 * 1024) == 0
t = np.frombuffer(buf, dtype=np.int64).reshape(64, 1024).astype(np.float64)
U = 48
names = {0: "mma1 start", 1: "mma1 A ready", 2: "mma1 B ready", 3: "mma1 S free", 4: "mma1 issued", 8: "mma2 start", 9: "mma2 W ready",
         10: "mma2 P ready", 11: "mma2 issued", 16: "prod B start", 17: "prod B slot free", 18: "prod W start", 19: "prod W slot free",
         24: "z wait tma", 25: "z tma landed", 26: "z norm done", 27: "z tma issue", 32: "epi wait S", 33: "epi S ready", 34: "epi ld0 done",
         35: "epi ld1 done", 36: "epi P free", 37: "epi packed", 38: "epi arrived"}
t0 = t[0, 0]


def seg(a, b, idx):
    d = t[b, idx] - t[a, idx]
    return f"{names[a]:>18s} -> {names[b]:<18s} mean {d.mean():8.0f}  max {d.max():8.0f}"


units = np.arange(6, U)
print("kernel span (mma1 first start .. last issue):", t[4, U - 1] - t0, "clk for", U, "units =", (t[4, U - 1] - t0) / U, "clk/unit")
for a, b in ((0, 1), (1, 2), (2, 3), (3, 4)):
    print(seg(a, b, units))
print("mma1 unit period", np.diff(t[4, units]).mean())
for a, b in ((8, 9), (9, 10), (10, 11)):
    print(seg(a, b, units))
for a, b in ((16, 17), (18, 19)):
    print(seg(a, b, units))
pairs = np.arange(2, 15)
for a, b in ((27, 25), (24, 25), (25, 26)):
    print(seg(a, b, pairs))
for s_ in range(3):
    us = np.arange(6 + s_, U, 3)
    print(f"set {(6 + s_) % 3}:")
    for a, b in ((32, 33), (33, 34), (34, 35), (35, 37), (37, 38)):
        print("   ", seg(a, b, us))
    print("    unit period of the set", np.diff(t[38, us]).mean())
print("first 14 units: S ready (epi) / mma1 issued / epi arrived, relative clk")
for u in range(14):
    print(u, int(t[4, u] - t0), int(t[33, u] - t0), int(t[38, u] - t0), int(t[11, u] - t0))
print("pairs: tma issue / z wait start / landed / norm done | mma1 first unit start, A ready, last unit issued   (relative clk)")
for p_ in range(10):
    u0 = 3 * p_
    print(p_, int(t[27, p_] - t0), int(t[24, p_] - t0), int(t[25, p_] - t0), int(t[26, p_] - t0), "|", int(t[0, u0] - t0), int(t[1, u0] - t0), int(t[4, u0 + 2] - t0))
print("unit: B(i) bulk issued / mma1 starts waiting for B(i) / B(i) seen ready / W(i) issued / mma2 W ready")
for u in range(8, 26):
    print(u, int(t[17, u] - t0), int(t[1, u] - t0), int(t[2, u] - t0), int(t[19, u] - t0), int(t[9, u] - t0))

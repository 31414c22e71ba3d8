"""In-kernel clock stamps of CTA 0 of the push scan (SCALEPROTO_B200_LIB -> a -DSPS_PUSH_TIMELINE build).
usage: timeline_push.py [fp32|fp16] [f16|f16x3]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import scaleprotoseg_b200 as sps
from scaleprotoseg_b200 import _lib, push as spush
from oracle import head_oracle as ho
dev = "cuda:0"
fp16 = len(sys.argv) > 1 and sys.argv[1] == "fp16"
math = sps.MATH_F16 if (len(sys.argv) > 2 and sys.argv[2] == "f16") else sps.MATH_F16X3
P, K, n, h, w = 760, 19, 16, 129, 257
g = torch.Generator(device=dev).manual_seed(1)
ident = ho.make_identity(P, K, 4)
protos = torch.rand(P, 64, device=dev, generator=g)
pc = spush.proto_class_of(ident).to(dev)
z = torch.sigmoid(torch.randn(n, 256, h, w, device=dev, generator=g))
if fp16:
    z = z.half().contiguous(memory_format=torch.channels_last)
labels = torch.stack([ho.make_labels(i, h, w, K, absent_class=3) for i in range(n)]).to(torch.int16).to(dev)
cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=math)
for _ in range(3):
    spush.scan_maps(z, labels, protos, pc, cfg)
torch.cuda.synchronize()
lib = _lib.load()
NB = 16 * 1024
buf = (ctypes.c_longlong * NB)()
lib.sps_debug_push_timeline.argtypes = [ctypes.c_void_p, ctypes.c_int]
print("rc", lib.sps_debug_push_timeline(buf, NB))
g_ = lambda r, i: buf[r * 1024 + i]
t0 = min(v for v in buf if v > 0)
f = lambda v: f"{v - t0:7d}" if v else "     -1"
names = ["mma top", "mma zfull", "mma dempty", "mma issued", "tma issue", "ld top", "ld ready", "ld zfull-arrive", "epi top", "epi dfull", "epi done"]
ntile = max(i for i in range(1024) if g_(3, i)) + 1
EPI2 = True
print("tiles of CTA 0:", ntile, " span", g_(3, ntile - 1) - t0, "clk ->", (g_(3, ntile - 1) - g_(3, 9)) / (ntile - 10), "clk per tile (steady)")
def mean(a, b, lo=10):
    v = [g_(b, i) - g_(a, i) for i in range(lo, ntile) if g_(a, i) and g_(b, i)]
    return (sum(v) / max(len(v), 1), max(v) if v else 0)
for a, b in ((0, 1), (1, 3), (4, 6), (5, 6), (6, 7), (7, 1)):
    m, mx = mean(a, b)
    print(f"{names[a]:>16s} -> {names[b]:<16s} mean {m:8.0f}  max {mx:8d}")
for a, b in ((8, 9), (9, 10), (3, 9)):
    m, mx = mean(a, b)
    print(f"{names[a]:>16s} -> {names[b]:<16s} mean {m:8.0f}  max {mx:8d}   (epilogue set 0)")
print("tile: " + " | ".join(names[:8]))
for i in range(20, 36):
    print(f"{i:4d} " + " ".join(f(g_(r, i)) for r in range(8)) + "  | epi set 0 top/dfull/done " + f(g_(8, i)) + f(g_(9, i)) + f(g_(10, i)))

st = [g_(12, b) for b in range(148) if g_(12, b)]
en = [g_(13, b) for b in range(148) if g_(13, b)]
if st:
    k0 = min(st)
    dur = sorted(e - s_ for s_, e in zip(st, en))
    print(f"CTAs {len(st)}: kernel span {max(en) - k0} ns; CTA start spread {max(st) - k0} ns; CTA duration min/median/max {dur[0]} / {dur[len(dur) // 2]} / {dur[-1]} ns")
    print("per-CTA duration (ns) by block:", [e - s_ for s_, e in zip(st, en)][::6])

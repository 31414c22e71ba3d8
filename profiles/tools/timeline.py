import sys, os, ctypes, torch
os.environ["SCALEPROTO_B200_LIB"] = "/root/repo/build/libsps_tl.so"
sys.path.insert(0, "/root/repo")
import scaleprotoseg_b200 as sps
from scaleprotoseg_b200 import _lib
from oracle import head_oracle as ho
dev = "cuda:0"
math = int(sys.argv[1]) if len(sys.argv) > 1 else sps.MATH_F16
B, H, W, P, K = 16, 65, 65, 760, 19
z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev)); protos = torch.rand(P, 64, device=dev)
W_ = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous().to(dev)
cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=math)
for _ in range(3): sps.head_forward_raw(z, protos, W_, cfg, False, False)
torch.cuda.synchronize()
lib = _lib.load()
buf = (ctypes.c_longlong * 8192)()
lib.sps_debug_timeline.argtypes = [ctypes.c_void_p, ctypes.c_int]
print("rc", lib.sps_debug_timeline(buf, 8192))
g = lambda r, e: buf[r * 1024 + e]
t0 = min(v for v in buf if v > 0)
f = lambda v: f"{v - t0:6d}" if v else "    -1"
for u in range(24, 50):
    print(u, "mma1: start", f(g(5, u)), "afull", f(g(6, u)), "bfull", f(g(4, u)), "sempty", f(g(7, u)), "issued", f(g(0, u)),
          "| mma2 pfull", f(g(1, u)), "issued", f(g(2, u)), "| epi top/sfull/ld0/st0/ld1/st1/done", " ".join(f(g(3, 8*u+k)) for k in (0,1,2,3,4,5,7)))

ent = [buf[5 * 1024 + 512 + i] for i in range(148)]; end = [buf[6 * 1024 + 512 + i] for i in range(148)]; ex = [buf[7 * 1024 + 512 + i] for i in range(148)]
t0g = min(ent)
print("entry min/max (us)", 0.0, (max(ent) - t0g) / 1e3, "| loop end min/max", (min(end) - t0g) / 1e3, (max(end) - t0g) / 1e3, "| exit min/max", (min(ex) - t0g) / 1e3, (max(ex) - t0g) / 1e3)
for i in (0, 1, 83, 84, 147):
    print(f"cta {i:3d}: entry {(ent[i]-t0g)/1e3:7.2f} loop end {(end[i]-t0g)/1e3:7.2f} exit {(ex[i]-t0g)/1e3:7.2f}")
print("cta0 first epilogue sfull (clk from first stamp):", g(3, 1) - t0, " first mma1 issued:", g(0, 0) - t0)

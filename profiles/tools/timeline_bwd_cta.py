import sys, os, ctypes, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))   # SCALEPROTO_B200_LIB -> a -DSPS_TIMELINE build
import scaleprotoseg_b200 as sps
from scaleprotoseg_b200 import _lib
from oracle import head_oracle as ho
dev = "cuda:0"
math = int(sys.argv[1]) if len(sys.argv) > 1 else sps.MATH_F16X3
nhwc = len(sys.argv) > 2 and sys.argv[2] == "nhwc16"
B, H, W, P, K = 16, 65, 65, 760, 19
z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev)); protos = torch.rand(P, 64, device=dev)
if nhwc: z = z.half().contiguous(memory_format=torch.channels_last)
W_ = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous().to(dev)
N = B * H * W
dl = torch.randn(N, K, device=dev) / N
cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=math)
for _ in range(3): sps.head_backward_raw(z, protos, W_, cfg, dl, None, None, True, True, True)
torch.cuda.synchronize()
lib = _lib.load()
buf = (ctypes.c_longlong * 8192)()
lib.sps_debug_timeline_bwd.argtypes = [ctypes.c_void_p, ctypes.c_int]
print("rc", lib.sps_debug_timeline_bwd(buf, 8192))
ent = [buf[5 * 1024 + i] for i in range(148)]; end = [buf[6 * 1024 + i] for i in range(148)]; ex = [buf[7 * 1024 + i] for i in range(148)]
t0 = min(ent)
print("entry  min/max (us)", 0.0, (max(ent) - t0) / 1e3)
print("loop end min/max (us)", (min(end) - t0) / 1e3, (max(end) - t0) / 1e3)
print("exit   min/max (us)", (min(ex) - t0) / 1e3, (max(ex) - t0) / 1e3)
for i in (0, 1, 36, 37, 38, 74, 111, 147):
    print(f"cta {i:3d}: entry {(ent[i]-t0)/1e3:7.2f} loop end {(end[i]-t0)/1e3:7.2f} exit {(ex[i]-t0)/1e3:7.2f}")

g = lambda r, e: buf[r * 1024 + e]
c0 = g(4, 960)
print("cta0 final section clk rel. to loop end: acc wait/got/bar1/bar2/synced", [g(4, 960 + k) - c0 for k in range(5)])
print("warp arrival at final barrier (clk rel.):", [g(4, 900 + w) - c0 for w in range(15)])

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
g = lambda r, e: buf[r * 1024 + e]
t0 = min(v for v in buf if v > 0)
f = lambda v: f"{v - t0:6d}" if v else "    -1"
print("last stamp", max(buf) - t0)
for u in range(12, 30):
    ti = u // 3
    print(u, "I1 top/waited/issued", f(g(0, 4*u)), f(g(0, 4*u+1)), f(g(0, 4*u+2)), "| I2 top/waited/issued", f(g(1, 4*u)), f(g(1, 4*u+1)), f(g(1, 4*u+2)),
          "| epi top/sfull/ld/math/gempty/gfull", " ".join(f(g(3, 8*u+k)) for k in range(6)))
    if u % 3 == 2:
        print("   tile", ti, "loader wait/got/done", f(g(2, 4*ti)), f(g(2, 4*ti+1)), f(g(2, 4*ti+2)), "[half0", f(g(6, 512+2*ti)), "half1", f(g(6, 513+2*ti)), "norm", f(g(2, 4*ti+3)), "] | dz readout start/dzfull/done", f(g(4, 4*ti)), f(g(4, 4*ti+1)), f(g(4, 4*ti+2)))

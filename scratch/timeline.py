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
tl = [[buf[r * 1024 + e] for e in range(64)] for r in range(8)]
t0 = min(v for r in tl for v in r if v > 0)
names = ["mma1_issued", "mma2_pfull_seen", "load_start", "load_done", "epi_sfull", "epi_ld_done", "epi_math_done", "epi_unit_done"]
for e in range(48):
    print(e, " ".join(f"{names[r][:12]}={tl[r][e]-t0 if tl[r][e] else -1:7d}" for r in range(8)))

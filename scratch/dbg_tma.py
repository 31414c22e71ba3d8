import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import scaleprotoseg_b200 as sps
from scaleprotoseg_b200 import _lib, head as sh
from oracle import head_oracle as ho
dev = "cuda:0"
which, mathname, shape = sys.argv[1], sys.argv[2], sys.argv[3]
math = {"f16": sps.MATH_F16, "f16x3": sps.MATH_F16X3, "bf16": sps.MATH_BF16}[mathname]
B, H, W, P, K = {"small": (2, 9, 7, 228, 19), "c2": (16, 65, 65, 760, 19), "one": (1, 8, 16, 228, 19)}[shape]
g = torch.Generator().manual_seed(1)
z = torch.sigmoid(torch.randn(B, 256, H, W, generator=g))
protos = torch.rand(P, 64, generator=g)
w = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous()
ranges = ho.uniform_ranges(P, 4)
dl = torch.randn(B * H * W, K, generator=g) / (B * H * W)
cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
zt = z.half().to(dev).contiguous(memory_format=torch.channels_last)
pc, wc, dlc = protos.to(dev), w.to(dev), dl.to(dev)
if which == "fwd":
    l, _, _ = sps.head_forward_raw(zt, pc, wc, cfg, False, False)
    torch.cuda.synchronize()
    lr, _, _ = ho.head_stage1(z.half().float(), protos.view(P, 64, 1, 1), ranges, w)
    print("fwd ok, max err", (l.cpu().view(lr.shape) - lr).abs().max().item())
elif which == "fwdd":
    l, d, a = sps.head_forward_raw(zt, pc, wc, cfg, True, True)
    torch.cuda.synchronize()
    print("fwd dense ok")
else:
    dz, dp, dw = sps.head_backward_raw(zt, pc, wc, cfg, dlc, None, None, True, True, True)
    torch.cuda.synchronize()
    print("bwd ok", dz.float().abs().max().item(), dp.abs().max().item())

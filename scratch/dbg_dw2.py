import sys, torch
sys.path.insert(0, "/root/repo")
import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
dev = "cuda:0"
B, H, W = 1, 9, 9
torch.manual_seed(0)
P, K, S, D = 228, 19, 4, 64
z = torch.sigmoid(torch.randn(B, S * D, H, W)).to(dev); protos = torch.rand(P, D).to(dev)
ident = ho.make_identity(P, K, S); W_ = ho.make_last_layer(ident).to(dev); ranges = ho.uniform_ranges(P, S)
cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_FP32)
lg, dg, ag = sps.head_forward_raw(z, protos, W_, cfg, True, True)
N = B*H*W
dl = (torch.randn(N, K) / N).to(dev)
ref = dl.t() @ ag
for name, dd, da in [("none", None, None), ("ddist0", torch.zeros_like(dg), None), ("dact0", None, torch.zeros_like(ag))]:
    for need_dz, need_dp in [(True, True), (False, True), (False, False), (True, False)]:
        dz, dp, dw = sps.head_backward_raw(z, protos, W_, cfg, dl, dd, da, need_dz, need_dp, True)
        print(name, need_dz, need_dp, "err", (dw - ref).abs().max().item(), "refmax", ref.abs().max().item(),
              "ratio", (dw / ref).flatten()[:4].tolist())

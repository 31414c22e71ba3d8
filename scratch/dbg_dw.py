import sys, torch
sys.path.insert(0, "/root/repo")
import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
dev = "cuda:0"
for (B, H, W) in [(1, 9, 9), (2, 9, 7), (1, 9, 14), (2, 8, 8), (1,1,126)]:
    torch.manual_seed(0)
    P, K, S, D = 228, 19, 4, 64
    z = torch.sigmoid(torch.randn(B, S * D, H, W)); protos = torch.rand(P, D, 1, 1)
    ident = ho.make_identity(P, K, S); W_ = ho.make_last_layer(ident); ranges = ho.uniform_ranges(P, S)
    zr, pr, wr = z.clone().requires_grad_(True), protos.clone().requires_grad_(True), W_.clone().requires_grad_(True)
    lo, do, ao = ho.head_stage1(zr, pr, ranges, wr)
    dl = torch.randn(lo.shape) / (B * H * W)
    (lo * dl).sum().backward()
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_FP32)
    zc, pc, wc = (t.to(dev).requires_grad_(True) for t in (z, protos, W_))
    lg, dg, ag = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, True, True)
    (lg.view(B, H, W, K) * dl.to(dev)).sum().backward()
    e = (wc.grad.cpu() - wr.grad).abs()
    print((B, H, W), "dW max err", e.max().item(), "ref max", wr.grad.abs().max().item(),
          "bad rows k", (e.max(1).values > 1e-5).nonzero().flatten().tolist()[:20],
          "bad cols", (e.max(0).values > 1e-5).nonzero().flatten().tolist()[:12],
          "dz err", (zc.grad.cpu() - zr.grad).abs().max().item(), "dp err", (pc.grad.cpu() - pr.grad).abs().max().item())

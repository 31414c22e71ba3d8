import sys, torch, time
sys.path.insert(0, "/root/repo")
import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
dev = "cuda:0"
def run(B, H, W, P, K, math, ranges=None, want=True):
    torch.manual_seed(0)
    S, D = 4, 64
    z = torch.sigmoid(torch.randn(B, S * D, H, W)); protos = torch.rand(P, D)
    ident = ho.make_identity(P, K, S) if ranges is None else None
    W_ = ho.make_last_layer(ident).contiguous() if ident is not None else torch.randn(K, P)
    ranges = ranges or ho.uniform_ranges(P, S)
    lo, do, ao = ho.head_stage1(z, protos.view(P, D, 1, 1), ranges, W_)
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    lg, dg, ag = sps.head_forward_raw(z.to(dev), protos.to(dev), W_.to(dev), cfg, want, want)
    torch.cuda.synchronize()
    def rel(a, b): return ((a.cpu().double() - b.double()).abs() / b.double().abs().clamp_min(1.0)).max().item()
    out = {"logits": rel(lg.view(lo.shape), lo)}
    if want: out.update(dist=rel(dg, do), act=rel(ag, ao))
    print((B, H, W, P, K), "math", math, {k: f"{v:.2e}" for k, v in out.items()}, flush=True)
for math in (sps.MATH_F16, sps.MATH_F16X3, sps.MATH_BF16, sps.MATH_BF16X3):
    run(1, 9, 9, 24, 2, math)
    run(2, 9, 7, 228, 19, math)
    run(1, 33, 40, 760, 19, math)
    run(3, 65, 65, 760, 19, math, want=False)
    run(2, 20, 20, 221, 19, math, ranges=[(0, 50), (50, 121), (121, 130), (130, 221)])
# timing at C2
B, H, W, P, K = 16, 65, 65, 760, 19
z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev)); protos = torch.rand(P, 64, device=dev)
W_ = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous().to(dev)
for math in (sps.MATH_FP32, sps.MATH_F16, sps.MATH_F16X3, sps.MATH_BF16):
    for dense in (False, True):
        cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=math)
        for _ in range(3): sps.head_forward_raw(z, protos, W_, cfg, dense, dense)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): sps.head_forward_raw(z, protos, W_, cfg, dense, dense)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print("fwd C2 math", math, "dense", dense, f"{ms*1e3:.1f} us  {B*H*W*P/ms/1e6:.1f} G pos*proto/s", flush=True)

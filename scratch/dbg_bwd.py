import sys, torch
sys.path.insert(0, "/root/repo")
import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
dev = "cuda:0"
def nerr(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30)).item()
def run(B, H, W, P, K, math, ranges=None, scale=1.0):
    torch.manual_seed(0)
    S, D = 4, 64
    z = torch.sigmoid(torch.randn(B, S * D, H, W, device=dev)); protos = torch.rand(P, D, device=dev)
    Wl = (ho.make_last_layer(ho.make_identity(P, K, S)).contiguous() if ranges is None else torch.randn(K, P)).to(dev)
    ranges = ranges or ho.uniform_ranges(P, S)
    N = B * H * W
    dl = (torch.randn(N, K, device=dev) / N * scale)
    c32 = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_FP32)
    ctc = sps.HeadConfig.from_ranges(ranges, K, math=math)
    dz0, dp0, dw0 = sps.head_backward_raw(z, protos, Wl, c32, dl, None, None, True, True, True)
    dz1, dp1, dw1 = sps.head_backward_raw(z, protos, Wl, ctc, dl, None, None, True, True, True)
    torch.cuda.synchronize()
    print((B, H, W, P, K), "math", math, f"dz {nerr(dz1, dz0):.2e} dp {nerr(dp1, dp0):.2e} dW {nerr(dw1, dw0):.2e}", flush=True)
for math in (sps.MATH_F16, sps.MATH_F16X3, sps.MATH_BF16X3):
    run(1, 9, 9, 24, 2, math)
    run(2, 9, 7, 228, 19, math)
    run(1, 33, 40, 760, 19, math)
    run(2, 20, 20, 221, 19, math, ranges=[(0, 50), (50, 121), (121, 130), (130, 221)])
    run(2, 20, 20, 1100, 19, math, ranges=[(0, 500), (500, 700), (700, 900), (900, 1100)])
    run(3, 65, 65, 760, 19, math, scale=1e4)
import time
B, H, W, P, K = 16, 65, 65, 760, 19
z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev)); protos = torch.rand(P, 64, device=dev)
Wl = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous().to(dev)
dl = torch.randn(B*H*W, K, device=dev) / (B*H*W)
for math in (sps.MATH_F16, sps.MATH_F16X3):
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=math)
    for _ in range(3): sps.head_backward_raw(z, protos, Wl, cfg, dl, None, None, True, True, True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): sps.head_backward_raw(z, protos, Wl, cfg, dl, None, None, True, True, True)
    e1.record(); torch.cuda.synchronize()
    print("bwd C2 math", math, f"{e0.elapsed_time(e1)/10*1e3:.1f} us", flush=True)

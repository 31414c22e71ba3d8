"""Quick device-timed comparison of the feature-map sources at the C2 shape (fwd / bwd, fused mode)."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import scaleprotoseg_b200 as sps
from scaleprotoseg_b200 import _lib, head as sh
from oracle import head_oracle as ho

dev = torch.device("cuda:0")
B, H, W, S, D, P, K = 16, 65, 65, 4, 64, 760, 19
if len(sys.argv) > 1 and sys.argv[1] == "c5":
    B, P, K = 32, 6840, 171
g = torch.Generator(device=dev).manual_seed(1)
NS = 4
zs = [torch.sigmoid(torch.randn(B, S * D, H, W, device=dev, generator=g)) for _ in range(NS)]
protos = torch.rand(P, D, device=dev, generator=g)
w = ho.make_last_layer(ho.make_identity(P, K, S)).contiguous().to(dev)
N = B * H * W
dls = [torch.randn(N, K, device=dev, generator=g) / N for _ in range(NS)]
ranges = ho.uniform_ranges(P, S)


def ev(fn, iters=50):
    for i in range(5):
        fn(i)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def run(name, math, zlist, flags=0, relay=False):
    base = sps.HeadConfig.from_ranges(ranges, K, math=math)
    cfg = sps.HeadConfig(base.scale_lo, base.num_classes, base.act_kind, base.eps, base.math, flags)
    out = {"mode": name}
    try:
        rel = [sh.z_relay(z, protos, cfg) for z in zlist] if relay else [None] * NS
        out["fwd_us"] = ev(lambda i: sps.head_forward_raw(zlist[i % NS], protos, w, cfg, False, False, relay=rel[i % NS]))
        out["bwd_us"] = ev(lambda i: sps.head_backward_raw(zlist[i % NS], protos, w, cfg, dls[i % NS], None, None, True, True, True, relay=rel[i % NS]))
        if relay:
            out["relay_us"] = ev(lambda i: sh.z_relay(zlist[i % NS], protos, cfg))
        out["step_us"] = out["fwd_us"] + out["bwd_us"] + out.get("relay_us", 0.0)
    except RuntimeError as e:
        out["error"] = str(e)[:300]
    print(json.dumps(out), flush=True)


z16 = [z.half().contiguous(memory_format=torch.channels_last) for z in zs]
zb16 = [z.bfloat16().contiguous(memory_format=torch.channels_last) for z in zs]
run("fp32 NCHW loaders f16x3", sps.MATH_F16X3, zs)
run("fp32 NCHW loaders f16", sps.MATH_F16, zs)
run("fp32 NCHW zpack(internal) f16x3", sps.MATH_F16X3, zs, _lib.FLAG_ZPACK)
run("fp32 NCHW relay shared f16x3", sps.MATH_F16X3, zs, _lib.FLAG_ZPACK, relay=True)
run("fp32 NCHW relay shared f16", sps.MATH_F16, zs, _lib.FLAG_ZPACK, relay=True)
run("fp16 NHWC tma f16x3", sps.MATH_F16X3, z16)
run("fp16 NHWC tma f16", sps.MATH_F16, z16)
run("bf16 NHWC tma bf16", sps.MATH_BF16, zb16)
print("fp32 reroutes:", int(_lib.load().sps_fp32_reroute_count()))

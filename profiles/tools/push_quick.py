"""Device-timed push scan on resident maps (16 maps of 256x129x257 per launch), fp32 NCHW and fp16 channels-last.
SCALEPROTO_B200_LIB may point at a debug build (-DSPS_PUSH_DBG_NOEPI / -DSPS_PUSH_DBG_NOMMA)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import scaleprotoseg_b200 as sps
from scaleprotoseg_b200 import push as spush
from oracle import head_oracle as ho
dev = "cuda:0"
P, K, n, h, w = 760, 19, 16, 129, 257
if len(sys.argv) > 1 and sys.argv[1] == "coco":
    P, K, n, h, w = 6840, 171, 16, 65, 65
g = torch.Generator(device=dev).manual_seed(1)
ident = ho.make_identity(P, K, 4)
protos = torch.rand(P, 64, device=dev, generator=g)
pc = spush.proto_class_of(ident).to(dev)
NS = 4   # 4 x 16 maps x 34 MB fp32 = 2.2 GB >> L2
zs = [torch.sigmoid(torch.randn(n, 256, h, w, device=dev, generator=g)) for _ in range(NS)]
labels = torch.stack([ho.make_labels(i, h, w, K, absent_class=3) for i in range(n)]).to(torch.int16).to(dev)
for name, math, conv in (("fp32 NCHW f16x3", sps.MATH_F16X3, None), ("fp32 NCHW f16", sps.MATH_F16, None),
                         ("fp16 NHWC f16x3", sps.MATH_F16X3, torch.float16), ("fp16 NHWC f16", sps.MATH_F16, torch.float16)):
    zl = zs if conv is None else [z.to(conv).contiguous(memory_format=torch.channels_last) for z in zs]
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=math)
    for i in range(3):
        spush.scan_maps(zl[i % NS], labels, protos, pc, cfg)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    IT = 20
    for i in range(IT):
        spush.scan_maps(zl[i % NS], labels, protos, pc, cfg)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / IT * 1e3
    print(json.dumps({"mode": name, "us_per_16_maps": round(us, 1), "ms_per_3000_maps": round(us * 3000 / n / 1e3, 2)}), flush=True)

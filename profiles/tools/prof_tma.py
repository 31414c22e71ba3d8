import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
dev = "cuda:0"
math = {"f16": sps.MATH_F16, "f16x3": sps.MATH_F16X3}[sys.argv[1] if len(sys.argv) > 1 else "f16"]
B, H, W, P, K = 16, 65, 65, 760, 19
g = torch.Generator(device=dev).manual_seed(1)
z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev, generator=g)).half().contiguous(memory_format=torch.channels_last)
protos = torch.rand(P, 64, device=dev, generator=g)
W_ = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous().to(dev)
dl = torch.randn(B * H * W, K, device=dev, generator=g) / (B * H * W)
cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=math)
for _ in range(4):
    sps.head_forward_raw(z, protos, W_, cfg, False, False)
    sps.head_backward_raw(z, protos, W_, cfg, dl, None, None, True, True, True)
torch.cuda.synchronize()
print("ok")

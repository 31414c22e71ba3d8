import sys, torch
sys.path.insert(0, "/root/repo")
import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
dev = "cuda:0"
math = int(sys.argv[1]) if len(sys.argv) > 1 else sps.MATH_F16
dense = len(sys.argv) > 2 and sys.argv[2] == "dense"
B, H, W, P, K = 16, 65, 65, 760, 19
z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev)); protos = torch.rand(P, 64, device=dev)
W_ = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous().to(dev)
cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=math)
for _ in range(3): sps.head_forward_raw(z, protos, W_, cfg, dense, dense)
torch.cuda.synchronize()
print("ok")

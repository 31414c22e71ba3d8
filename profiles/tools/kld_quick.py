"""KLD loss steps (dense and feature mode) at the C2 batch, for ncu launch lists."""
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
from scaleprotoseg_b200.kld import KldTables, kld_loss
B, H, W, P, K, S = 16, 65, 65, 760, 19, 4
dev = 'cuda:0'
ranges = ho.uniform_ranges(P, S)
tables = KldTables(ho.make_identity(P, K, S), ranges, dev)
labels = torch.stack([ho.make_labels(i, H, W, K) for i in range(B)]).to(dev)
dist = (torch.rand(B, P, H, W, device=dev) * 6).requires_grad_(True)
z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev)).requires_grad_(True)
pv = torch.rand(P, 64, 1, 1, device=dev).requires_grad_(True)
cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_FP32)
for i in range(3):
    dist.grad = None
    kld_loss(dist, labels, tables).backward()
    z.grad = pv.grad = None
    sps.kld_loss_from_features(z, pv, labels, tables, cfg).backward()
torch.cuda.synchronize()

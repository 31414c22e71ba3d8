"""One pass over the kernels the write-up cites, at the BASELINE shapes (for ncu --set full)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import scaleprotoseg_b200 as sps
from scaleprotoseg_b200 import push as spush, head as sh
from oracle import head_oracle as ho
dev = "cuda:0"
ITERS = int(os.environ.get("PROF_ITERS", "2"))


def run(which):
    g = torch.Generator(device=dev).manual_seed(1)
    B, H, W, P, K = 16, 65, 65, 760, 19
    ranges = ho.uniform_ranges(P, 4)
    ident = ho.make_identity(P, K, 4)
    W_ = ho.make_last_layer(ident).contiguous().to(dev)
    protos = torch.rand(P, 64, device=dev, generator=g)
    N = B * H * W
    if which in ("c2", "c2_16"):
        z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev, generator=g))
        if which == "c2_16":
            z = z.half().contiguous(memory_format=torch.channels_last)
        dl = torch.randn(N, K, device=dev, generator=g) / N
        cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_F16X3)
        for _ in range(ITERS):
            sps.head_forward_raw(z, protos, W_, cfg, False, False)
            sps.head_backward_raw(z, protos, W_, cfg, dl, None, None, True, True, True)
    elif which == "c5":
        B, P, K = 32, 6840, 171
        N = B * H * W
        z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev, generator=g))
        protos = torch.rand(P, 64, device=dev, generator=g)
        W_ = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous().to(dev)
        dl = torch.randn(N, K, device=dev, generator=g) / N
        cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=sps.MATH_F16X3)
        for _ in range(ITERS):
            sps.head_forward_raw(z, protos, W_, cfg, False, False)
            sps.head_backward_raw(z, protos, W_, cfg, dl, None, None, True, True, True)
    elif which in ("push", "push16"):
        n, h, w = 16, 129, 257
        z = torch.sigmoid(torch.randn(n, 256, h, w, device=dev, generator=g))
        if which == "push16":
            z = z.half().contiguous(memory_format=torch.channels_last)
        labels = torch.stack([ho.make_labels(i, h, w, K, absent_class=3) for i in range(n)]).to(torch.int16).to(dev)
        cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_F16X3)
        pc = spush.proto_class_of(ident).to(dev)
        for _ in range(ITERS):
            keys = spush.scan_maps(z, labels, protos, pc, cfg)
        spush.topk_keys(torch.randint(0, 2 ** 62, (3000, P), device=dev, dtype=torch.int64), 5)
    elif which == "stage2":
        z = torch.sigmoid(torch.randn(B, 256, H, W, device=dev, generator=g))
        cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_F16X3)
        tables = sh.build_group_tables(ident, 3, dev)
        Gp = torch.rand(3 * P, device=dev, generator=g).requires_grad_(True)
        Wg = torch.randn(K, 3 * tables.Kc, device=dev, generator=g).requires_grad_(True)
        dl = torch.randn(N, K, device=dev, generator=g) / N
        dgr = torch.randn(N, 3 * tables.Kc, device=dev, generator=g) / N
        for _ in range(ITERS):
            _, dist, act = sps.head_forward_raw(z, protos, None, cfg, True, True)
            groups, logits = sh.GroupHeadFn.apply(act, Gp, Wg, cfg, tables, (B, H, W))
            Gp.grad = Wg.grad = None
            torch.autograd.backward([logits, groups], [dl, dgr])
        d = torch.rand(1, P, 129, 257, device=dev, generator=g) * 8.0
        sps.upsample_argmin(d, (1024, 2048))

    torch.cuda.synchronize()


names = sys.argv[1:]
if names == ["all"]:
    names = ["c2", "c2_16", "c5", "push", "push16", "stage2"]
for nm in names:
    run(nm)
    torch.cuda.empty_cache()
print("ok")

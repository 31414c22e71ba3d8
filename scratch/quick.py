"""Quick GPU iteration: time fwd / bwd / push kernels (CUDA events, rotating inputs) and check them vs the fp32 kernels.
usage: python scratch/quick.py [math=4] [what=fwd,bwd,push]"""
import os, sys, torch
sys.path.insert(0, "/root/repo")
import scaleprotoseg_b200 as sps
from scaleprotoseg_b200 import push as spush
from oracle import head_oracle as ho
dev = "cuda:0"
math = int(sys.argv[1]) if len(sys.argv) > 1 else sps.MATH_F16X3
what = sys.argv[2].split(",") if len(sys.argv) > 2 else ["fwd", "bwd", "push"]
B, H, W, P, K = 16, 65, 65, 760, 19
torch.manual_seed(0)
zs = [torch.sigmoid(torch.randn(B, 256, H, W, device=dev)) for _ in range(4)]
protos = torch.rand(P, 64, device=dev)
ident = ho.make_identity(P, K, 4)
W_ = ho.make_last_layer(ident).contiguous().to(dev)
N = B * H * W
dls = [torch.randn(N, K, device=dev) / N for _ in range(4)]
ACT = os.environ.get("ACT", "log")
cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, act=ACT, math=math)
cfg32 = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, act=ACT, math=sps.MATH_FP32)

def timeit(fn, iters=20, warm=5):
    for i in range(warm): fn(i)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for i in range(iters): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3

def rel(a, b, floor=1.0):
    return ((a - b).abs() / b.abs().clamp_min(floor)).max().item()

if "fwd" in what:
    t = timeit(lambda i: sps.head_forward_raw(zs[i % 4], protos, W_, cfg, False, False))
    lg = sps.head_forward_raw(zs[0], protos, W_, cfg, False, False)[0]
    lr = sps.head_forward_raw(zs[0], protos, W_, cfg32, False, False)[0]
    lgd, dd, ad = sps.head_forward_raw(zs[0], protos, W_, cfg, True, True)
    lrd, dr, ar = sps.head_forward_raw(zs[0], protos, W_, cfg32, True, True)
    td = timeit(lambda i: sps.head_forward_raw(zs[i % 4], protos, W_, cfg, True, True))
    print(f"fwd fused {t:.1f} us | dense {td:.1f} us | logits err {rel(lg, lr):.2e} dense: logits {rel(lgd, lrd):.2e} dist {rel(dd, dr):.2e} act {rel(ad, ar):.2e}")
if "bwd" in what:
    t = timeit(lambda i: sps.head_backward_raw(zs[i % 4], protos, W_, cfg, dls[i % 4], None, None, True, True, True))
    g = sps.head_backward_raw(zs[0], protos, W_, cfg, dls[0], None, None, True, True, True)
    r = sps.head_backward_raw(zs[0], protos, W_, cfg32, dls[0], None, None, True, True, True)
    errs = [((a - b).abs().max() / b.abs().max()).item() for a, b in zip(g, r)]
    print(f"bwd {t:.1f} us | grad err (dz, dp, dW) " + " ".join(f"{e:.2e}" for e in errs))
if "push" in what:
    h, w, nm = 129, 257, 16
    zp = torch.sigmoid(torch.randn(nm * 2, 256, h, w, device=dev))
    labels = torch.stack([ho.make_labels(i, h, w, K, absent_class=3) for i in range(nm * 2)]).to(torch.int16).to(dev)
    pc = spush.proto_class_of(ident).to(dev)
    keys = torch.empty(nm, P, device=dev, dtype=torch.int64)
    t = timeit(lambda i: spush.scan_maps(zp[(i % 2) * nm:(i % 2 + 1) * nm], labels[(i % 2) * nm:(i % 2 + 1) * nm], protos, pc, cfg, keys_out=keys), iters=10, warm=3)
    k1 = spush.scan_maps(zp[:4], labels[:4], protos, pc, cfg).clone()
    k0 = spush.scan_maps(zp[:4], labels[:4], protos, pc, cfg32).clone()
    i1, v1 = spush.decode_keys(k1); i0, v0 = spush.decode_keys(k0)
    same = (i1 == i0).float().mean().item()
    print(f"push {t:.1f} us per {nm} maps = {nm*h*w/t*1e6/1e9:.2f} Gpatch/s, {nm*h*w*1026/t*1e6/1e9:.0f} GB/s | idx agree {same:.4f} val err {rel(v1, v0):.2e}")
if "train" in what:
    # what stage-1 training really calls: dense distances out, gradient into distances (KLD) and logits
    t1 = timeit(lambda i: sps.head_forward_raw(zs[i % 4], protos, W_, cfg, True, False))
    dd = torch.randn(B, P, H, W, device=dev) * 1e-3 / N
    t2 = timeit(lambda i: sps.head_backward_raw(zs[i % 4], protos, W_, cfg, dls[i % 4], dd, None, True, True, True), iters=5, warm=2)
    print(f"train-style: fwd (logits + dense dist) {t1:.1f} us | bwd (dlogits + ddist) {t2:.1f} us")
if "stage2" in what:
    from scaleprotoseg_b200 import head as sh
    tables = sh.build_group_tables(ident, 3, dev)
    Gp = torch.rand(3 * P, device=dev).requires_grad_(True)
    Wg = torch.randn(K, 3 * tables.Kc, device=dev).requires_grad_(True)
    def s2(i, train=True):
        act = sps.head_forward_raw(zs[i % 4], protos, None, cfg, False, True)[2]
        groups, lg = sh.GroupHeadFn.apply(act, Gp, Wg, cfg, tables, (B, H, W))
        if train:
            Gp.grad = Wg.grad = None
            lg.backward(dls[i % 4])
    tf = timeit(lambda i: sps.head_forward_raw(zs[i % 4], protos, None, cfg, False, True))
    act0 = sps.head_forward_raw(zs[0], protos, None, cfg, False, True)[2]
    tg = timeit(lambda i: sh.GroupHeadFn.apply(act0, Gp, Wg, cfg, tables, (B, H, W)))
    tt = timeit(lambda i: s2(i))
    print(f"stage2: head fwd (act only) {tf:.1f} us | group fwd {tg:.1f} us | fwd+bwd (frozen prototypes) {tt:.1f} us")

"""GPU parity: the CUDA path (through the C ABI) against (i) golden vectors produced by the unmodified
reference and (ii) the CPU oracle on seeded inputs.

Tolerances (BASELINE.json north_star: rel 1e-3):
  distances   |err| <= 1e-3 * max(|ref|, 1)
  activations |err| <= 1e-3 * max(|ref|, 1)      (log-similarity; slope ~1/(d+eps) is why d gets the floor)
  logits      |err| <= 1e-3 * max(|ref|, 1)
  gradients   |err| <= 1e-3 * max|ref|            (per tensor, normalised by its largest entry)
  push        indices bit-exact where the runner-up is further than the distance tolerance, otherwise the
              distance at the returned index must be within tolerance of the oracle minimum.
SPS_MATH_FP32 (exact operands) is held to 20x tighter bounds.  The tensor-core modes:
  F16X3 / BF16X3  (3-term split products)  the 1e-3 bounds above on every output and gradient
  F16             (fp16 operands, rounded)  distances / logits 2e-3 (measured 6e-4), gradients 2e-3
  BF16            (bf16 operands, rounded)  distances / logits 1.5e-2 (8-bit mantissa), gradients 1.5e-2
"""
import os

import numpy as np
import pytest
import torch

import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
from scaleprotoseg_b200 import push as spush

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
MATHS = [sps.MATH_FP32]


TC_MATHS = [sps.MATH_F16X3, sps.MATH_BF16X3, sps.MATH_F16, sps.MATH_BF16]


def _tol(math):
    return {sps.MATH_FP32: 5e-5, sps.MATH_F16X3: 1e-3, sps.MATH_BF16X3: 1e-3, sps.MATH_F16: 2e-3, sps.MATH_BF16: 1.5e-2}[math]


def assert_rel(got, ref, tol, name, floor=1.0):
    got, ref = torch.as_tensor(got).detach().cpu().double(), torch.as_tensor(ref).detach().cpu().double()
    assert got.shape == ref.shape, f"{name}: shape {tuple(got.shape)} vs {tuple(ref.shape)}"
    err = ((got - ref).abs() / ref.abs().clamp_min(floor)).max().item() if ref.numel() else 0.0
    assert err <= tol, f"{name}: max rel err {err:.3e} > {tol:.1e}"


def assert_grad(got, ref, tol, name):
    ref = torch.as_tensor(ref).detach().cpu().double()
    scale = ref.abs().max().clamp_min(1e-30).item()
    assert_rel(torch.as_tensor(got).detach().cpu().double() / scale, ref / scale, tol, name, floor=1.0)


def _load(golden_dir, name):
    return {k: v for k, v in np.load(os.path.join(golden_dir, name + ".npz")).items()}


def _ranges(g):
    return [tuple(int(v) for v in r) for r in g["ranges"]]


def _cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.mark.parametrize("math", MATHS)
@pytest.mark.parametrize("name", ["stage1_em", "stage1_em_ragged", "stage1_linear", "stage1_city228"])
def test_stage1_golden(built_lib, golden_dir, name, math):
    g = _load(golden_dir, name)
    K = g["last_w"].shape[0]
    cfg = sps.HeadConfig.from_ranges(_ranges(g), K, act=str(g["act_kind"]), math=math)
    z, p, w = (_cu(g[k]).requires_grad_(True) for k in ("z", "protos", "last_w"))
    logits, d, a = sps.PrototypeHeadFn.apply(z, p, w, cfg, True, True)
    tol = _tol(math)
    assert_rel(d, g["dist"], tol, "dist"), assert_rel(a, g["act"], tol, "act")
    assert_rel(logits.view(g["logits"].shape), g["logits"], tol, "logits")
    loss = (logits.view(g["logits"].shape) * _cu(g["dlogits"])).sum() + (d * _cu(g["ddist"])).sum() + \
        (a * _cu(g["dact"])).sum()
    dz, dp, dw = torch.autograd.grad(loss, [z, p, w])
    gt = 20 * tol if math == sps.MATH_FP32 else tol
    assert_grad(dz, g["dz"], gt, "dz"), assert_grad(dp, g["dprotos"], gt, "dprotos")
    assert_grad(dw, g["dlast_w"], gt, "dlast_w")


@pytest.mark.parametrize("math", MATHS + [sps.MATH_F16X3])      # F16X3: what install() feeds the group head with by default
@pytest.mark.parametrize("name", ["stage2_em", "stage2_city228_ragged"])
def test_stage2_golden(built_lib, golden_dir, name, math):
    g = _load(golden_dir, name)
    ident = torch.from_numpy(g["identity"])
    K = ident.shape[1]
    cfg = sps.HeadConfig.from_ranges(_ranges(g), K, math=math)
    n_mod = int(g["n_group_mods"])
    z, p, wg = (_cu(g[k]).requires_grad_(True) for k in ("z", "protos", "last_group_w"))
    gws = [_cu(g[f"group_w_{k}"]).requires_grad_(True) for k in range(n_mod)]
    tables = sps.build_group_tables(ident, 3, DEV)
    assert tables.Kc == n_mod
    _, d, a = sps.PrototypeHeadFn.apply(z, p, None, cfg, True, True)
    G = torch.cat([w_.reshape(-1) for w_ in gws])
    B, _, H, W = z.shape
    groups, logits = sps.GroupHeadFn.apply(a, G, wg, cfg, tables, (B, H, W))
    tol = _tol(math)
    assert_rel(d, g["dist"], tol, "dist"), assert_rel(a, g["act"], tol, "act")
    assert_rel(groups, g["groups"], 4 * tol, "groups"), assert_rel(logits.view(g["logits"].shape), g["logits"], 4 * tol, "logits")
    loss = (logits.view(g["logits"].shape) * _cu(g["dlogits"])).sum() + (d * _cu(g["ddist"])).sum() + \
        (a * _cu(g["dact"])).sum() + (groups * _cu(g["dgroups"])).sum()
    grads = torch.autograd.grad(loss, [z, p, wg] + gws)
    gt = 20 * tol if math == sps.MATH_FP32 else tol
    assert_grad(grads[0], g["dz"], gt, "dz"), assert_grad(grads[1], g["dprotos"], gt, "dprotos")
    assert_grad(grads[2], g["dlast_group_w"], gt, "dlast_group_w")
    allref = np.concatenate([g[f"dgroup_w_{k}"].ravel() for k in range(n_mod)])
    allgot = torch.cat([grads[3 + k].reshape(-1) for k in range(n_mod)])
    assert_grad(allgot, allref, gt, "dgroup_w")


@pytest.mark.parametrize("K,per", [(171, 1), (40, 2)])
def test_stage2_many_classes_vs_oracle(built_lib, K, per):
    """Stage-2 grouping head with a wide last_layer_group (COCO-Stuff class count: it does not fit shared memory and
    is read from global memory; the dWg kernel stages fewer rows per step) against the oracle's autograd."""
    B, H, W, S, D, J = 1, 7, 6, 4, 64, 3
    P = K * per * S
    gen = torch.Generator().manual_seed(ho.SEED + K)
    z = torch.sigmoid(torch.randn(B, S * D, H, W, generator=gen))
    protos = torch.rand(P, D, 1, 1, generator=gen)
    ident = ho.make_identity(P, K, S)
    ranges = ho.uniform_ranges(P, S)
    members = ho.class_members(ident)
    gws = [torch.softmax(torch.randn(J, len(m), generator=gen), dim=1) for m in members]
    wg = torch.randn(K, J * len(members), generator=gen) * 0.1
    N = B * H * W
    dlog = torch.randn(B, H, W, K, generator=gen) / N
    dgr = torch.randn(N, J * len(members), generator=gen) / N
    gr = [w.clone().requires_grad_(True) for w in gws]
    wr = wg.clone().requires_grad_(True)
    lo, _, ao, go = ho.head_stage2(z, protos, ranges, ident, gr, wr)
    ((lo * dlog).sum() + (go * dgr).sum()).backward()
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_FP32)
    tables = sps.build_group_tables(ident, J, DEV)
    _, _, a = sps.PrototypeHeadFn.apply(z.to(DEV), protos.to(DEV), None, cfg, False, True)
    Gc = torch.cat([w.reshape(-1) for w in gws]).to(DEV).requires_grad_(True)
    wc = wg.to(DEV).requires_grad_(True)
    groups, logits = sps.GroupHeadFn.apply(a, Gc, wc, cfg, tables, (B, H, W))
    assert_rel(groups, go, 4e-4, "groups"), assert_rel(logits.view(B, H, W, K), lo, 4e-4, "logits")
    ((logits.view(B, H, W, K) * dlog.to(DEV)).sum() + (groups * dgr.to(DEV)).sum()).backward()
    assert_grad(wc.grad, wr.grad, 1e-3, "dlast_group_w")
    assert_grad(Gc.grad, torch.cat([w.grad.reshape(-1) for w in gr]), 1e-3, "dgroup_w")


def check_push(keys, win_map, z_maps, labels, protos, ranges, ident, K, tol):
    """keys (n,P) int64 from the kernel, win_map (P,) global winners."""
    n = len(z_maps)
    idx_k, val_k = spush.decode_keys(keys.cpu())
    list_val = []
    for i in range(n):
        d = ho.scale_l2_convolution(z_maps[i], protos, ranges)
        idx_o, val_o = ho.push_image_min(d, labels[i], ident, K, 0)
        list_val.append(val_o)
        absent = val_o[0] >= 1e10
        assert torch.equal(val_k[i][absent], val_o[0][absent]) and (idx_k[i][absent] == 0).all()
        # distance at the kernel's index must be (within tolerance) the oracle minimum ...
        dk = d[0].flatten(1)[torch.arange(d.shape[1]), idx_k[i]]
        ok = ~absent
        assert ((dk[ok] - val_o[0][ok]).abs() <= tol * val_o[0][ok].clamp_min(1.0)).all()
        assert ((val_k[i][ok] - val_o[0][ok]).abs() <= tol * val_o[0][ok].clamp_min(1.0)).all()
        # ... and the index is bit-exact wherever the runner-up is further than the tolerance
        masked = d[0].flatten(1).clone()
        cls = ident.argmax(1)
        lab = labels[i].flatten()
        masked[(lab.unsqueeze(0) - 1) != cls.unsqueeze(1)] = float("inf")
        top2 = masked.topk(2, dim=1, largest=False).values
        clear = ok & ((top2[:, 1] - top2[:, 0]) > 2 * tol * top2[:, 0].clamp_min(1.0))
        assert torch.equal(idx_k[i][clear], idx_o[0][clear])
    tot = torch.cat(list_val)
    win_o = tot.argmin(0)
    s2 = tot.topk(min(2, n), dim=0, largest=False).values
    clear = torch.ones_like(win_o, dtype=torch.bool) if n < 2 else \
        ((s2[1] - s2[0]) > 2 * tol * s2[0].clamp_min(1.0)) | (s2[0] >= 1e10)
    assert torch.equal(win_map.cpu()[clear], win_o[clear])
    return win_o


@pytest.mark.parametrize("math", MATHS + [sps.MATH_F16X3, sps.MATH_F16, sps.MATH_BF16X3])
@pytest.mark.parametrize("name", ["push_small", "push_city228"])
def test_push_golden(built_lib, golden_dir, name, math):
    g = _load(golden_dir, name)
    K = int(g["K"])
    ident = torch.from_numpy(g["identity"])
    protos = torch.from_numpy(g["protos"])
    z = torch.sigmoid(torch.from_numpy(g["x"]))
    cfg = sps.HeadConfig.from_ranges(_ranges(g), K, math=math)
    labels = torch.from_numpy(g["labels"])
    wm, pix, wd, out, keys = spush.push_project(z.to(DEV), labels.to(DEV), protos.to(DEV), ident, cfg, batch=2)
    idx_k, val_k = spush.decode_keys(keys.cpu())
    if math == sps.MATH_FP32:
        # exact-operand mode: every index the reference produced, bit for bit (ties included)
        same = idx_k.numpy() == g["idx"]
        assert same.mean() > 0.995
        assert np.array_equal(wm.cpu().numpy()[np.all(same, axis=0)], g["win"][np.all(same, axis=0)])
        agree = (wm.cpu().numpy() == g["win"]) & (pix.cpu().numpy() == g["idx"][g["win"], np.arange(len(g["win"]))])
        got = out.cpu().numpy().reshape(g["new_protos"].shape)
        assert np.array_equal(got[agree], g["new_protos"][agree]) and agree.mean() > 0.99
    check_push(keys, wm, [z[i:i + 1] for i in range(z.shape[0])], [labels[i].long() for i in range(z.shape[0])],
               protos, _ranges(g), ident, K, _tol(math))


# ---------------------------------------------------------------------------------------------
# Tensor-core (tcgen05) modes
# ---------------------------------------------------------------------------------------------
def _grad_tol(math):
    return {sps.MATH_F16X3: 1e-3, sps.MATH_F16: 2e-3, sps.MATH_BF16X3: 1.5e-2, sps.MATH_BF16: 1.5e-2}[math]


@pytest.mark.parametrize("math", TC_MATHS)
@pytest.mark.parametrize("name", ["stage1_em", "stage1_em_ragged", "stage1_city228"])
def test_stage1_golden_tensor_core(built_lib, golden_dir, name, math):
    """Golden vectors of the unmodified reference through the tcgen05 forward (dense mode) and, for the gradient
    that arrives through the logits only, the tcgen05 backward."""
    g = _load(golden_dir, name)
    K = g["last_w"].shape[0]
    cfg = sps.HeadConfig.from_ranges(_ranges(g), K, act=str(g["act_kind"]), math=math)
    z, p, w = (_cu(g[k]).requires_grad_(True) for k in ("z", "protos", "last_w"))
    logits, d, a = sps.PrototypeHeadFn.apply(z, p, w, cfg, True, True)
    tol = _tol(math)
    assert_rel(d, g["dist"], tol, "dist"), assert_rel(a, g["act"], tol, "act")
    assert_rel(logits.view(g["logits"].shape), g["logits"], tol, "logits")
    # all three upstream gradients (dense ones go through the exact-operand kernels)
    loss = (logits.view(g["logits"].shape) * _cu(g["dlogits"])).sum() + (d * _cu(g["ddist"])).sum() + (a * _cu(g["dact"])).sum()
    dz, dp, dw = torch.autograd.grad(loss, [z, p, w])
    assert_grad(dz, g["dz"], 1e-3, "dz"), assert_grad(dp, g["dprotos"], 1e-3, "dprotos"), assert_grad(dw, g["dlast_w"], 1e-3, "dlast_w")


@pytest.mark.parametrize("math", TC_MATHS)
@pytest.mark.parametrize("shape", [(2, 9, 7, 228, 19, None), (1, 33, 40, 760, 19, None),
                                   (4, 65, 65, 760, 19, None),   # more tiles than CTAs per scale: every partial slot written
                                   (1, 12, 11, 320, 40, None),   # 40 classes: the one-pass wide backward (head_bwd_wide.cu)
                                   (1, 9, 10, 684, 171, None),   # COCO-Stuff class count
                                   (2, 20, 20, 221, 19, [(0, 50), (50, 121), (121, 130), (130, 221)]),
                                   (1, 12, 11, 1100, 7, [(0, 500), (500, 700), (700, 900), (900, 1100)])])
def test_tensor_core_fwd_bwd_vs_oracle(built_lib, shape, math):
    """Fused-mode step (logits out, dlogits in) on seeded inputs: tcgen05 forward + tcgen05 backward against the
    CPU oracle's autograd; uniform and ragged scale ranges, a scale wider than one CTA group (500 prototypes),
    a random last layer."""
    B, H, W, P, K, ranges = shape
    gen = torch.Generator().manual_seed(ho.SEED + P)
    z = torch.sigmoid(torch.randn(B, 256, H, W, generator=gen))
    protos = torch.rand(P, 64, 1, 1, generator=gen)
    last_w = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous() if ranges is None else torch.randn(K, P, generator=gen)
    ranges = ranges or ho.uniform_ranges(P, 4)
    N = B * H * W
    dl = torch.randn(B, H, W, K, generator=gen) / N
    zr, pr, wr = z.clone().requires_grad_(True), protos.clone().requires_grad_(True), last_w.clone().requires_grad_(True)
    lo, _, ao = ho.head_stage1(zr, pr, ranges, wr)
    lo.backward(dl)
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    zc, pc, wc = (t.to(DEV).requires_grad_(True) for t in (z, protos, last_w))
    lg, dnone, anone = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, False, False)
    assert dnone is None and anone is None            # fused mode: no (N x P) map exists
    lg.view(B, H, W, K).backward(dl.to(DEV))
    if math in (sps.MATH_F16X3, sps.MATH_BF16X3) and K <= 32:
        assert_rel(lg.view(B, H, W, K), lo, _tol(math), "logits")
    else:
        # rounded 16-bit operands: the error of a P-term sum is bounded relative to sum_p |a W|, not to |logit|.  More than 32
        # classes in a 3-term mode: the distances keep their 3-term products, the logits GEMM runs on one 16-bit term (the
        # hi + lo last-layer tiles do not fit shared memory), so the logits carry the single-term bound of the operand type
        lt = _tol(math) if K <= 32 else _tol({sps.MATH_F16X3: sps.MATH_F16, sps.MATH_BF16X3: sps.MATH_BF16}.get(math, math))
        bound = (ao.detach().abs() @ wr.detach().abs().t()).clamp_min(1.0)
        err = ((lg.detach().cpu().double() - lo.detach().reshape(-1, K).double()).abs() / bound.double()).max().item()
        assert err <= lt, f"logits: {err:.3e} of sum|aW| > {lt:.1e}"
    gt = _grad_tol(math)
    if K > 32:   # the wide backward recomputes the distances with single-term 16-bit products (stated in its header)
        gt = max(gt, _grad_tol(sps.MATH_F16))
    assert_grad(zc.grad, zr.grad, gt, "dz"), assert_grad(pc.grad, pr.grad, gt, "dprotos"), assert_grad(wc.grad, wr.grad, gt, "dlast_w")


@pytest.mark.parametrize("layout", ["fp32_nchw", "fp16_nhwc"])
def test_coco_shape_fwd_bwd_vs_oracle(built_lib, layout):
    """BASELINE configs[4] head shape (P = 6 840 prototypes in 4 scales, K = 171 classes) on 3 tiles of positions (the CPU
    oracle takes seconds at this size): fused forward and the one-pass wide backward - 27 chunks per scale stream through
    the tile-stationary dz kernel, 108 chunk CTAs x tile ranges in the chunk-stationary dp / dW kernel."""
    B, H, W, P, K = 1, 3, 95, 6840, 171
    gen = torch.Generator().manual_seed(ho.SEED + 171)
    z = torch.sigmoid(torch.randn(B, 256, H, W, generator=gen))
    if layout == "fp16_nhwc":
        z = z.half().float()
    protos = torch.rand(P, 64, 1, 1, generator=gen)
    last_w = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous() + 0.02 * torch.randn(K, P, generator=gen)
    ranges = ho.uniform_ranges(P, 4)
    N = B * H * W
    dl = torch.randn(B, H, W, K, generator=gen) / N
    zr, pr, wr = z.clone().requires_grad_(True), protos.clone().requires_grad_(True), last_w.clone().requires_grad_(True)
    lo, _, ao = ho.head_stage1(zr, pr, ranges, wr)
    lo.backward(dl)
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_F16X3)
    zin = z.to(DEV) if layout == "fp32_nchw" else z.half().to(DEV).contiguous(memory_format=torch.channels_last)
    zc, pc, wc = zin.requires_grad_(True), protos.to(DEV).requires_grad_(True), last_w.to(DEV).requires_grad_(True)
    before = int(built_lib.sps_fp32_reroute_count())
    lg, _, _ = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, False, False)
    lg.view(B, H, W, K).backward(dl.to(DEV))
    assert int(built_lib.sps_fp32_reroute_count()) == before, "the COCO shape must stay on the tensor-core kernels"
    bound = (ao.detach().abs() @ wr.detach().abs().t()).clamp_min(1.0)
    err = ((lg.detach().cpu().double() - lo.detach().reshape(-1, K).double()).abs() / bound.double()).max().item()
    assert err <= 1e-3, f"logits: {err:.3e} of sum|aW|"
    assert_grad(zc.grad.float(), zr.grad, 3e-3, "dz"), assert_grad(pc.grad, pr.grad, 2e-3, "dprotos"), assert_grad(wc.grad, wr.grad, 2e-3, "dlast_w")


@pytest.mark.parametrize("math", TC_MATHS)
@pytest.mark.parametrize("depth", [32, 8])
def test_tensor_core_small_depth(built_lib, depth, math):
    """Prototype depth below 64 (channels per scale D = 32, 8): the feature loaders, the dz read-out and the
    partial-gradient write-out all have a 'channels that exist' bound."""
    B, H, W, P, K, S = 2, 13, 11, 120, 5, 4
    gen = torch.Generator().manual_seed(ho.SEED + depth)
    z = torch.sigmoid(torch.randn(B, S * depth, H, W, generator=gen))
    protos = torch.rand(P, depth, 1, 1, generator=gen)
    last_w = torch.randn(K, P, generator=gen)
    ranges = ho.uniform_ranges(P, S)
    dl = torch.randn(B, H, W, K, generator=gen) / (B * H * W)
    zr, pr, wr = z.clone().requires_grad_(True), protos.clone().requires_grad_(True), last_w.clone().requires_grad_(True)
    lo, _, ao = ho.head_stage1(zr, pr, ranges, wr)
    lo.backward(dl)
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    zc, pc, wc = (t.to(DEV).requires_grad_(True) for t in (z, protos, last_w))
    lg, _, _ = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, False, False)
    lg.view(B, H, W, K).backward(dl.to(DEV))
    bound = (ao.detach().abs() @ wr.detach().abs().t()).clamp_min(1.0)
    err = ((lg.detach().cpu().double() - lo.detach().reshape(-1, K).double()).abs() / bound.double()).max().item()
    assert err <= _tol(math), f"logits: {err:.3e} of sum|aW| > {_tol(math):.1e}"
    # 8 channels per scale: distances are small (|z - p|^2 over 8 terms), so many pairs sit where a'(d) = -(1-eps)/((d+1)(d+eps))
    # is steep and the 16-bit gamma tile dominates: twice the tolerance of the 64-channel shapes
    gt = _grad_tol(math) * (2.0 if depth == 8 else 1.0)
    assert_grad(zc.grad, zr.grad, gt, "dz"), assert_grad(pc.grad, pr.grad, gt, "dprotos"), assert_grad(wc.grad, wr.grad, gt, "dlast_w")


def test_tensor_core_pushed_prototype(built_lib):
    """d -> 0 (a prototype equal to a feature vector, as after push): the split mode keeps fp32-class
    distances where the slope of the log similarity is 1/eps."""
    P, K = 24, 2
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K, math=sps.MATH_F16X3)
    z = torch.sigmoid(torch.randn(1, 256, 4, 4, generator=torch.Generator().manual_seed(3)))
    p = torch.rand(P, 64, generator=torch.Generator().manual_seed(4))
    p[7] = z[0, 64:128, 2, 3]
    w = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous()
    lg, dg, ag = sps.head_forward_raw(z.to(DEV), p.to(DEV), w.to(DEV), cfg, True, True)
    lr, dr, ar = ho.head_stage1(z, p.view(P, 64, 1, 1), ho.uniform_ranges(P, 4), w)
    assert dg[0, 7, 2, 3].item() <= 2e-5
    assert_rel(dg, dr, 1e-3, "dist")
    keep = torch.ones_like(ar, dtype=torch.bool)
    keep[2 * 4 + 3, 7] = False                         # the d ~ 1e-6 element itself: |a| > 7 on both sides
    assert ag[2 * 4 + 3, 7].item() > 7.0
    assert_rel(ag[keep], ar[keep], 1e-3, "act")


@pytest.mark.parametrize("math", [sps.MATH_F16X3, sps.MATH_F16])
def test_push_tensor_core_large_map(built_lib, math):
    """Maps whose size is not a multiple of the 128-position tile, several maps per launch, ragged scale ranges
    wider than one CTA group (300 prototypes in a scale), a class absent from every map, a void-only map."""
    K, S = 6, 4
    ranges = [(0, 300), (300, 366), (366, 372), (372, 480)]
    P = 480
    gen = torch.Generator().manual_seed(ho.SEED + 5)
    ident = torch.zeros(P, K)
    ident[torch.arange(P), torch.randint(0, K, (P,), generator=gen)] = 1
    protos = torch.rand(P, 64, 1, 1, generator=gen)
    n, h, w = 5, 23, 31
    z = torch.sigmoid(torch.randn(n, 256, h, w, generator=gen))
    labels = torch.stack([ho.make_labels(i, h, w, K, absent_class=2) for i in range(n)])
    labels[3] = 0
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    pc = spush.proto_class_of(ident)
    keys = spush.scan_maps(z.to(DEV), labels.to(torch.int16).to(DEV), protos.to(DEV), pc, cfg)
    best = spush.reduce_keys(keys, 0)
    win, _ = spush.decode_keys(best)
    check_push(keys, win, [z[i:i + 1] for i in range(n)], [labels[i] for i in range(n)], protos, ranges, ident, K, _tol(math))


@pytest.mark.parametrize("K", [35, 40, 171])
@pytest.mark.parametrize("math", [sps.MATH_F16X3, sps.MATH_BF16X3])
def test_push_tensor_core_many_classes(built_lib, K, math):
    """K = 35 is the largest class count whose mask rides in the GEMM (label values up to 35); K = 40 and the
    COCO-Stuff count take the label-compare epilogue.  Both must give the same answers as the oracle."""
    S, P = 4, 4 * 96
    ranges = ho.uniform_ranges(P, S)
    gen = torch.Generator().manual_seed(ho.SEED + K)
    ident = torch.zeros(P, K)
    ident[torch.arange(P), torch.randint(0, K, (P,), generator=gen)] = 1
    protos = torch.rand(P, 64, 1, 1, generator=gen)
    n, h, w = 3, 19, 27
    z = torch.sigmoid(torch.randn(n, 256, h, w, generator=gen))
    labels = torch.randint(0, K + 1, (n, h, w), generator=gen)          # 0 = void, per-pixel classes: many ties of class
    labels[0, :4] = K                                                     # the top label value is present
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    pc = spush.proto_class_of(ident)
    keys = spush.scan_maps(z.to(DEV), labels.to(torch.int16).to(DEV), protos.to(DEV), pc, cfg)
    best = spush.reduce_keys(keys, 0)
    win, _ = spush.decode_keys(best)
    check_push(keys, win, [z[i:i + 1] for i in range(n)], [labels[i] for i in range(n)], protos, ranges, ident, K, _tol(math))


def test_push_label_values_outside_the_class_range_never_match(built_lib):
    """Label values above K (or negative) belong to no prototype: they behave like void in both scan epilogues."""
    K, S, P = 19, 4, 228
    ranges = ho.uniform_ranges(P, S)
    gen = torch.Generator().manual_seed(ho.SEED + 99)
    ident = ho.make_identity(P, K, S)
    protos = torch.rand(P, 64, 1, 1, generator=gen)
    h, w = 17, 21
    z = torch.sigmoid(torch.randn(2, 256, h, w, generator=gen))
    labels = torch.stack([ho.make_labels(i, h, w, K, absent_class=3) for i in range(2)])
    weird = labels.clone()
    weird[0, :3] = 77          # beyond MASK_MAX_LABEL
    weird[1, :3] = -5
    as_void = labels.clone()
    as_void[:, :3] = 0
    pc = spush.proto_class_of(ident)
    for math in (sps.MATH_F16X3, sps.MATH_FP32):
        cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
        k1 = spush.scan_maps(z.to(DEV), weird.to(torch.int16).to(DEV), protos.to(DEV), pc, cfg)
        k0 = spush.scan_maps(z.to(DEV), as_void.to(torch.int16).to(DEV), protos.to(DEV), pc, cfg)
        assert torch.equal(k1, k0)


@pytest.mark.parametrize("math", TC_MATHS)
@pytest.mark.parametrize("shape", [(2, 9, 7, 228, 19, None), (1, 33, 40, 760, 19, None),
                                   (1, 12, 11, 320, 40, None),   # two class slabs: the distance gradient enters once
                                   (2, 20, 20, 221, 19, [(0, 50), (50, 121), (121, 130), (130, 221)])])
def test_tensor_core_backward_with_distance_gradient(built_lib, shape, math):
    """Stage-1 training step as the reference runs it: the loss reads the logits AND the dense distance map (KLD loss,
    segmentation/module_multiscale.py:242), so a gradient arrives through both.  The tcgen05 backward takes ddist
    next to dlogits; compared with the CPU oracle's autograd."""
    B, H, W, P, K, ranges = shape
    gen = torch.Generator().manual_seed(ho.SEED + 7 * P)
    z = torch.sigmoid(torch.randn(B, 256, H, W, generator=gen))
    protos = torch.rand(P, 64, 1, 1, generator=gen)
    last_w = ho.make_last_layer(ho.make_identity(P, K, 4)).contiguous() if ranges is None else torch.randn(K, P, generator=gen)
    ranges = ranges or ho.uniform_ranges(P, 4)
    N = B * H * W
    dl = torch.randn(B, H, W, K, generator=gen) / N
    dd = torch.randn(B, P, H, W, generator=gen) * (3.0 / (N * P))
    dd[:, ::3] = 0                                     # the KLD loss touches only some (pixel, prototype) pairs
    zr, pr, wr = z.clone().requires_grad_(True), protos.clone().requires_grad_(True), last_w.clone().requires_grad_(True)
    lo, do, _ = ho.head_stage1(zr, pr, ranges, wr)
    ((lo * dl).sum() + (do * dd).sum()).backward()
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    zc, pc, wc = (t.to(DEV).requires_grad_(True) for t in (z, protos, last_w))
    lg, dg, _ = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, True, False)
    ((lg.view(B, H, W, K) * dl.to(DEV)).sum() + (dg * dd.to(DEV)).sum()).backward()
    gt = _grad_tol(math)
    assert_grad(zc.grad, zr.grad, gt, "dz"), assert_grad(pc.grad, pr.grad, gt, "dprotos"), assert_grad(wc.grad, wr.grad, gt, "dlast_w")
    # same answer as the exact-operand kernels on the same inputs (they served this case before)
    cfg32 = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_FP32)
    g_tc = sps.head_backward_raw(z.to(DEV), protos.to(DEV), last_w.to(DEV), cfg, dl.reshape(N, K).to(DEV), dd.to(DEV), None, True, True, True)
    g_32 = sps.head_backward_raw(z.to(DEV), protos.to(DEV), last_w.to(DEV), cfg32, dl.reshape(N, K).to(DEV), dd.to(DEV), None, True, True, True)
    for a, b_, name in zip(g_tc, g_32, ("dz", "dprotos", "dlast_w")):
        assert_grad(a, b_.cpu(), gt, name + " vs fp32 kernels")


@pytest.mark.parametrize("math", [sps.MATH_FP32, sps.MATH_F16X3, sps.MATH_BF16X3])
@pytest.mark.parametrize("shape", [(2, 17, 13, 190, 19), (1, 65, 65, 210, 21)])
def test_single_scale_head_and_push(built_lib, shape, math):
    """S = 1 (the single-scale ProtoSeg baseline, segmentation/model/model.py:250-268: C = D = 64, one range owning every
    prototype; the second shape is its Pascal head, 21 classes x 10): fused step and push scan against the oracle."""
    B, H, W, P, K = shape
    gen = torch.Generator().manual_seed(ho.SEED + P + H)
    z = torch.sigmoid(torch.randn(B, 64, H, W, generator=gen))
    protos = torch.rand(P, 64, 1, 1, generator=gen)
    ident = ho.make_identity(P, K, 1)
    last_w = ho.make_last_layer(ident).contiguous()
    ranges = [(0, P)]
    dl = torch.randn(B, H, W, K, generator=gen) / (B * H * W)
    zr, pr, wr = z.clone().requires_grad_(True), protos.clone().requires_grad_(True), last_w.clone().requires_grad_(True)
    lo, do, _ = ho.head_stage1(zr, pr, ranges, wr)
    lo.backward(dl)
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    zc, pc, wc = (t.to(DEV).requires_grad_(True) for t in (z, protos, last_w))
    lg, dg, _ = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, True, False)
    assert_rel(dg, do, _tol(math), "distances"), assert_rel(lg.view(B, H, W, K), lo, _tol(math), "logits")
    lg.view(B, H, W, K).backward(dl.to(DEV))
    gt = 1e-3 if math == sps.MATH_FP32 else _grad_tol(math)
    if math == sps.MATH_F16X3:
        gt = 2e-3    # single scale: dz is one scale's 190-210 terms, no other scale's gradient to be relative to (measured 1.7e-3)
    assert_grad(zc.grad, zr.grad, gt, "dz"), assert_grad(pc.grad, pr.grad, gt, "dprotos"), assert_grad(wc.grad, wr.grad, gt, "dlast_w")
    labels = [ho.make_labels(i, H, W, K, absent_class=3) for i in range(B)]
    z_maps = [z[i:i + 1] for i in range(B)]
    pcl = spush.proto_class_of(ident).to(DEV)
    keys = spush.scan_maps(z.to(DEV), torch.stack(labels).to(torch.int16).to(DEV), protos.view(P, 64).to(DEV), pcl, cfg)
    best = spush.reduce_keys(keys, 0)
    check_push(keys, (best & 0xFFFFFFFF).cpu(), z_maps, labels, protos, ranges, ident, K, max(_tol(math), 1e-4))


@pytest.mark.parametrize("math", [sps.MATH_FP32, sps.MATH_F16X3])
def test_topk_nearest_vs_oracle(built_lib, math):
    """k nearest maps per prototype (find_nearest.py:89-175): distances within tolerance, map and pixel indices exact
    where the neighbouring distances differ by more than the tolerance; two shards chained through the merge give
    exactly the single-pass lists."""
    n, h, w, P, K, k = 9, 12, 10, 40, 5, 4
    ident, ranges = ho.make_identity(P, K, 4), ho.uniform_ranges(P, 4)
    z = torch.cat([ho.make_feature_map(i, 256, h, w) for i in range(n)])
    labels = torch.stack([ho.make_labels(i, h, w, K, block=3) for i in range(n)])
    labels[4] = 0                                                  # a void-only image never enters a list
    gen = torch.Generator().manual_seed(5)
    protos = torch.rand(P, 64, 1, 1, generator=gen)
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    tol = max(_tol(math), 1e-4)
    dref, iref, pref = ho.topk_nearest([ho.scale_l2_convolution(z[i:i + 1], protos, ranges) for i in range(n)],
                                       list(labels), k)
    lab16 = labels.to(torch.int16).to(DEV)
    dist, img, pix = spush.topk_nearest(z.to(DEV), lab16, protos.to(DEV), cfg, k)
    dist, img, pix = dist.cpu(), img.cpu(), pix.cpu()
    assert ((dist - dref).abs() <= tol * dref.clamp_min(1.0)).all()
    assert (dist[:, 1:] >= dist[:, :-1]).all() and (img != 4).all()
    full = torch.stack([(ho.scale_l2_convolution(z[i:i + 1], protos, ranges)[0] + 10e6 * (labels[i] == 0)).flatten(1).min(1).values
                        for i in range(n)]).t()                  # (P, n) per-image minima
    srt = full.sort(dim=1).values
    gap_ok = torch.ones(P, k, dtype=torch.bool)
    for r in range(k):                                             # rank r is unambiguous when both neighbours are far
        lo = srt[:, r] - (srt[:, r - 1] if r > 0 else -1e30)
        hi = srt[:, r + 1] - srt[:, r]
        gap_ok[:, r] = (lo > 2 * tol * srt[:, r].clamp_min(1.0)) & (hi > 2 * tol * srt[:, r].clamp_min(1.0))
    assert torch.equal(img[gap_ok], iref[gap_ok]) and gap_ok.float().mean() > 0.9
    same_img = img == iref
    d_at = torch.stack([ho.scale_l2_convolution(z[i:i + 1], protos, ranges)[0].flatten(1) for i in range(n)])   # (n,P,hw)
    got_d = d_at[img, torch.arange(P).unsqueeze(1), pix]
    assert ((got_d[same_img] - dref[same_img]).abs() <= 2 * tol * dref[same_img].clamp_min(1.0)).all()
    # class-pure lists: the mask of the push scan instead of the void-only mask
    pc = spush.proto_class_of(ident)
    dcls, icls, _ = spush.topk_nearest(z.to(DEV), lab16, protos.to(DEV), cfg, 1, proto_class=pc)
    keys = spush.scan_maps(z.to(DEV), lab16, protos.view(P, 64).to(DEV), pc.to(DEV), cfg)
    best = spush.reduce_keys(keys, 0)
    assert torch.equal(icls[:, 0].cpu(), (best & 0xFFFFFFFF).cpu())
    # shards: maps 0..4 then 5..8 with map_base, chained through the merge
    pc0 = torch.zeros(P, dtype=torch.int32, device=DEV)
    nv = spush.nonvoid_labels(lab16)
    k0 = spush.scan_maps(z[:5].to(DEV), nv[:5], protos.view(P, 64).to(DEV), pc0, cfg)
    k1 = spush.scan_maps(z[5:].to(DEV), nv[5:], protos.view(P, 64).to(DEV), pc0, cfg)
    top = spush.topk_keys(k0, k, 0)
    top = spush.topk_keys(k1, k, 5, top)
    assert torch.equal((top[0] & 0xFFFFFFFF).cpu(), img) and torch.equal(top[1].cpu().long(), pix)
    a, b = spush.topk_keys(k0, k, 0), spush.topk_keys(k1, k, 5)
    mk, mp = spush.merge_topk([a[0], b[0]], [a[1], b[1]])
    assert torch.equal(mk.cpu(), top[0].cpu()) and torch.equal(mp.cpu(), top[1].cpu())


def test_topk_nearest_golden(built_lib, golden_dir):
    """k nearest images per prototype against the ranking the reference's own find_k_nearest_patches_to_prototypes
    (find_nearest.py:71-352) produced on a small on-disk dataset (tests/golden/topk_small.npz): image order exact wherever
    consecutive ranks are further apart than the distance tolerance; distances and positions against the oracle."""
    g = _load(golden_dir, "topk_small")
    ranges, k, K = _ranges(g), int(g["k"]), int(g["K"])
    z = torch.sigmoid(torch.from_numpy(g["x"]))
    protos = torch.from_numpy(g["protos"])
    labels = torch.from_numpy(g["labels"].astype(np.int64))
    maps = [ho.scale_l2_convolution(z[i:i + 1], protos, ranges) for i in range(z.shape[0])]
    val, img, pix = ho.topk_nearest(maps, list(labels), k)
    assert np.array_equal(img.numpy(), g["rank_img"])
    for math in (sps.MATH_FP32, sps.MATH_F16X3):
        cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
        d, m, px = spush.topk_nearest(z.to(DEV), labels.to(torch.int16).to(DEV), protos.to(DEV), cfg, k)
        tol = _tol(math)
        assert_rel(d, val, tol, "top-k distances")
        gap = (val[:, 1:] - val[:, :-1]) > 4 * tol * val[:, 1:].clamp_min(1.0)
        clear = torch.cat([gap, torch.ones(val.shape[0], 1, dtype=torch.bool)], 1) & torch.cat([torch.ones(val.shape[0], 1, dtype=torch.bool), gap], 1)
        assert torch.equal(m.cpu()[clear], torch.from_numpy(g["rank_img"])[clear]), "ranked images"
        assert torch.equal(px.cpu()[clear], pix[clear]), "positions of the closest patches"

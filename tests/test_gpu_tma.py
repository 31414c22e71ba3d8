"""GPU parity of the TMA-fed kernels: 16-bit channels-last feature maps (SpsHeadDesc.z_dtype / z_layout) and the
16-bit re-lay of fp32 NCHW maps (SPS_FLAG_ZPACK), against the CPU oracle ON THE SAME 16-bit-rounded inputs.

Tolerances (north star: rel 1e-3 with fp32 accumulation, operand types stated):
  fp16 maps, F16X3   the map is exact in one fp16 term, the prototypes / activations / last layer are split hi + lo:
                     distances, activations, logits 1e-3 (measured ~1e-5); gradients 1e-3 of the tensor's max
                     (dz is returned in fp16: 1e-3 + one fp16 rounding)
  fp16 maps, F16     prototypes rounded to fp16: distances / logits 2e-3, gradients 2e-3
  bf16 maps, BF16X3  distances / activations / logits 1e-3; gradients 1.5e-2 (the gradient GEMMs run single-term bf16
                     operands, as in the loader-fed kernels: tests/test_gpu_parity.py::_grad_tol)
  bf16 maps, BF16    1.5e-2 (8-bit mantissa, stated fast mode)
  fp32 maps + ZPACK  the same bounds as the loader-fed tensor-core modes (tests/test_gpu_parity.py)
"""
import pytest
import torch

import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
from scaleprotoseg_b200 import _lib
from scaleprotoseg_b200 import push as spush
from test_gpu_parity import DEV, assert_grad, assert_rel

pytestmark = pytest.mark.gpu

TOL = {sps.MATH_F16X3: 1e-3, sps.MATH_BF16X3: 1e-3, sps.MATH_F16: 2e-3, sps.MATH_BF16: 1.5e-2}
GTOL = {sps.MATH_F16X3: 1e-3, sps.MATH_BF16X3: 1.5e-2, sps.MATH_F16: 2e-3, sps.MATH_BF16: 1.5e-2}
DT = {sps.MATH_F16X3: torch.float16, sps.MATH_F16: torch.float16, sps.MATH_BF16X3: torch.bfloat16, sps.MATH_BF16: torch.bfloat16}


def _case(B, H, W, P, K, seed, S=4, D=64, ragged=False):
    g = torch.Generator().manual_seed(ho.SEED + seed)
    z = torch.sigmoid(torch.randn(B, S * D, H, W, generator=g))
    protos = torch.rand(P, D, generator=g)
    ident = ho.make_identity(P, K, S)
    ranges = ho.uniform_ranges(P, S)
    if ragged:   # what prune_prototypes leaves behind: unequal scale ranges
        n = P // S
        ranges = [(0, n - 3), (n - 3, 2 * n + 2), (2 * n + 2, 3 * n + 1), (3 * n + 1, P)]
    w = ho.make_last_layer(ident) + 0.05 * torch.randn(K, P, generator=g)
    N = B * H * W
    dl = torch.randn(B, H, W, K, generator=g) / N
    return z, protos, ranges, w, dl


def _oracle(z, protos, ranges, w, dl, ddist=None):
    zr, pr, wr = z.clone().requires_grad_(True), protos.clone().view(protos.shape[0], -1, 1, 1).requires_grad_(True), w.clone().requires_grad_(True)
    lo, do, ao = ho.head_stage1(zr, pr, ranges, wr)
    loss = (lo * dl).sum()
    if ddist is not None:
        loss = loss + (do * ddist).sum()
    loss.backward()
    return lo, do, ao, zr.grad, pr.grad.view(protos.shape), wr.grad


@pytest.mark.parametrize("math", [sps.MATH_F16X3, sps.MATH_F16, sps.MATH_BF16X3, sps.MATH_BF16])
@pytest.mark.parametrize("shape", [(2, 9, 7, 228, 19, False), (3, 13, 11, 228, 19, True), (1, 5, 3, 24, 2, False), (2, 23, 17, 760, 19, False)])
def test_nhwc16_forward_backward_vs_oracle(built_lib, math, shape):
    """16-bit channels-last maps through the public autograd API: dense outputs, then the fused training step
    (logits only, gradient into z, the prototypes and the last layer)."""
    B, H, W, P, K, ragged = shape
    z, protos, ranges, w, dl = _case(B, H, W, P, K, seed=B * 100 + H, ragged=ragged)
    dt, tol = DT[math], TOL[math]
    z16 = z.to(dt)
    lo, do, ao, dz_o, dp_o, dw_o = _oracle(z16.float(), protos, ranges, w, dl)
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    zc = z16.to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    pc, wc = protos.to(DEV).requires_grad_(True), w.to(DEV).requires_grad_(True)
    lg, dg, ag = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, True, True)
    assert_rel(dg, do, tol, "dist"), assert_rel(ag, ao, tol, "act"), assert_rel(lg.view(lo.shape), lo, tol, "logits")
    lf, _, _ = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, False, False)     # fused mode: three epilogue sets
    assert_rel(lf.view(lo.shape), lo, tol, "fused logits")
    (lf.view(lo.shape) * dl.to(DEV)).sum().backward()
    assert zc.grad.dtype == dt and zc.grad.is_contiguous(memory_format=torch.channels_last)
    gt = GTOL[math]
    gz = gt + (1e-3 if dt == torch.float16 else 8e-3)      # dz carries one more rounding to the map's element type
    assert_grad(zc.grad.float(), dz_o, gz, "dz"), assert_grad(pc.grad, dp_o, gt, "dprotos"), assert_grad(wc.grad, dw_o, gt, "dW")


@pytest.mark.parametrize("math", [sps.MATH_F16X3, sps.MATH_F16, sps.MATH_BF16X3])
def test_nhwc16_backward_through_distance_map(built_lib, math):
    """Dense training mode on a 16-bit map: a gradient also arrives through the distance map (KLD loss, module_multiscale.py:242)."""
    B, H, W, P, K = 2, 11, 9, 228, 19
    z, protos, ranges, w, dl = _case(B, H, W, P, K, seed=77)
    g = torch.Generator().manual_seed(5)
    dd = torch.randn(B, P, H, W, generator=g) * (1e-3 / (B * H * W))
    dt, tol = DT[math], TOL[math]
    z16 = z.to(dt)
    lo, do, ao, dz_o, dp_o, dw_o = _oracle(z16.float(), protos, ranges, w, dl, dd)
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    zc = z16.to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    pc, wc = protos.to(DEV).requires_grad_(True), w.to(DEV).requires_grad_(True)
    lg, dg, _ = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, True, False)
    ((lg.view(lo.shape) * dl.to(DEV)).sum() + (dg * dd.to(DEV)).sum()).backward()
    gt = GTOL[math]
    gz = gt + (1e-3 if dt == torch.float16 else 8e-3)
    assert_grad(zc.grad.float(), dz_o, gz, "dz"), assert_grad(pc.grad, dp_o, gt, "dprotos"), assert_grad(wc.grad, dw_o, gt, "dW")


@pytest.mark.parametrize("math", [sps.MATH_F16X3, sps.MATH_F16, sps.MATH_BF16X3])
@pytest.mark.parametrize("shape", [(2, 9, 7, 228, 19, False), (3, 13, 11, 228, 19, True), (2, 23, 17, 760, 19, False)])
def test_zpack_fp32_maps_vs_oracle(built_lib, math, shape):
    """fp32 NCHW maps with SPS_FLAG_ZPACK: one 16-bit re-lay per step (shared by forward and backward through the autograd
    Function), TMA-fed kernels, dz in fp32 NCHW.  Also the raw entry points, which re-lay inside their own workspace."""
    B, H, W, P, K, ragged = shape
    z, protos, ranges, w, dl = _case(B, H, W, P, K, seed=B * 10 + W, ragged=ragged)
    lo, do, ao, dz_o, dp_o, dw_o = _oracle(z, protos, ranges, w, dl)
    tol = TOL[math]
    base = sps.HeadConfig.from_ranges(ranges, K, math=math)
    cfg = sps.HeadConfig(base.scale_lo, base.num_classes, base.act_kind, base.eps, base.math, base.flags | _lib.FLAG_ZPACK)
    zc, pc, wc = z.to(DEV).requires_grad_(True), protos.to(DEV).requires_grad_(True), w.to(DEV).requires_grad_(True)
    lg, dg, ag = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, True, True)
    assert_rel(dg, do, tol, "dist"), assert_rel(ag, ao, tol, "act"), assert_rel(lg.view(lo.shape), lo, tol, "logits")
    lf, _, _ = sps.PrototypeHeadFn.apply(zc, pc, wc, cfg, False, False)
    (lf.view(lo.shape) * dl.to(DEV)).sum().backward()
    assert zc.grad.dtype == torch.float32
    gt = GTOL[math]
    assert_grad(zc.grad, dz_o, gt, "dz"), assert_grad(pc.grad, dp_o, gt, "dprotos"), assert_grad(wc.grad, dw_o, gt, "dW")
    # raw calls (no shared re-lay)
    lr, _, _ = sps.head_forward_raw(z.to(DEV), protos.to(DEV), w.to(DEV), cfg, False, False)
    assert torch.equal(lr, lf.detach())
    dz2, dp2, dw2 = sps.head_backward_raw(z.to(DEV), protos.to(DEV), w.to(DEV), cfg, dl.to(DEV), None, None, True, True, True)
    assert torch.equal(dz2, zc.grad) and torch.equal(dp2, pc.grad) and torch.equal(dw2, wc.grad)


def test_nhwc16_full_size_agrees_with_loader_kernels(built_lib):
    """BASELINE configs[1] shape (B=16, 65x65, P=760, K=19).  On fp16-representable maps the TMA-fed kernels and the
    loader-fed kernels see the same operands: logits and parameter gradients agree to accumulation-order noise, and a
    random subset of positions is replayed through the oracle."""
    B, H, W, P, K = 16, 65, 65, 760, 19
    z, protos, ranges, w, dl = _case(B, H, W, P, K, seed=1)
    z16 = z.half()
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_F16X3)
    zl = z16.float().to(DEV)
    zt = z16.to(DEV).contiguous(memory_format=torch.channels_last)
    pc, wc, dlc = protos.to(DEV), w.to(DEV), dl.to(DEV)
    l_ldg, _, _ = sps.head_forward_raw(zl, pc, wc, cfg, False, False)
    l_tma, _, _ = sps.head_forward_raw(zt, pc, wc, cfg, False, False)
    assert_rel(l_tma, l_ldg, 2e-5, "logits tma vs loader")
    dz_l, dp_l, dw_l = sps.head_backward_raw(zl, pc, wc, cfg, dlc, None, None, True, True, True)
    dz_t, dp_t, dw_t = sps.head_backward_raw(zt, pc, wc, cfg, dlc, None, None, True, True, True)
    assert_grad(dp_t, dp_l, 1e-4, "dprotos"), assert_grad(dw_t, dw_l, 1e-4, "dW")
    assert_grad(dz_t.float(), dz_l, 1e-3, "dz (fp16 output)")
    g = torch.Generator().manual_seed(3)
    sel = torch.randint(0, B * H * W, (257,), generator=g)
    zc = z16.float().permute(0, 2, 3, 1).reshape(-1, 256)[sel].t().reshape(1, 256, 1, 257)
    lr, _, _ = ho.head_stage1(zc, protos.view(P, 64, 1, 1), ranges, w)
    assert_rel(l_tma[sel.to(DEV)], lr.view(257, K), 1e-3, "logits subset vs oracle")


@pytest.mark.parametrize("math", [sps.MATH_F16X3, sps.MATH_F16])
@pytest.mark.parametrize("shape", [(3, 17, 9, 228, 19), (2, 33, 29, 760, 19), (2, 9, 11, 160, 40)])
def test_push_scan_nhwc16_vs_oracle(built_lib, math, shape):
    """Push scan fed by TMA from fp16 channels-last maps (K = 40: the label-compare epilogue) against the oracle on the same
    rounded maps: index-exact where the runner-up is further than the tolerance, and the winners' vectors bit-exact."""
    n, h, w, P, K = shape
    g = torch.Generator().manual_seed(ho.SEED + n * h)
    z16 = torch.sigmoid(torch.randn(n, 256, h, w, generator=g)).half()
    protos = torch.rand(P, 64, generator=g)
    ident = ho.make_identity(P, K, 4)
    ranges = ho.uniform_ranges(P, 4)
    labels = torch.stack([ho.make_labels(i, h, w, K, absent_class=3) for i in range(n)])
    win, idx, val, newp = ho.push_full([z16[i:i + 1].float() for i in range(n)], list(labels), protos.view(P, 64, 1, 1), ranges, ident, K)
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    zc = z16.to(DEV).contiguous(memory_format=torch.channels_last)
    wm, pix, wd, out, keys = spush.push_project(zc, labels.to(torch.int16).to(DEV), protos.to(DEV), ident, cfg)
    tol = TOL[math]
    kidx, kval = spush.decode_keys(keys)
    ref_val, ref_idx = torch.cat(val), torch.cat(idx)
    absent = ref_val >= 1e10
    assert torch.equal((kval.cpu() >= 1e10), absent)
    assert_rel(kval.cpu()[~absent], ref_val[~absent], tol, "per-map minimum")
    # index-exact where the masked runner-up is further away than the tolerance
    for i in range(n):
        d = ho.scale_l2_convolution(z16[i:i + 1].float(), protos.view(P, 64, 1, 1), ranges)[0].flatten(1)
        lab = labels[i].flatten()
        cls = ident.argmax(1) + 1
        masked = torch.where(lab.unsqueeze(0) == cls.unsqueeze(1), d, torch.full_like(d, 1e10))
        top2 = masked.topk(2, dim=1, largest=False).values
        clear = (top2[:, 1] - top2[:, 0]) > 4 * tol * top2[:, 0].clamp_min(1.0)
        clear &= ~absent[i]
        assert torch.equal(kidx.cpu()[i][clear], ref_idx[i][clear]), f"map {i}: argmin index"
    ref_pix = torch.stack([idx[int(win[p])][0, p] for p in range(P)])
    same = (wm.cpu() == win) & (pix.cpu() == ref_pix)      # same winner (a near-tie inside the tolerance may pick a neighbour)
    assert same.float().mean() > 0.95
    got = out.cpu()[same]
    assert torch.equal(got, torch.from_numpy(newp).view(P, 64)[same]), "winning vectors (exact copies of the fp16 map values)"


# ---- fp32 NCHW maps by TMA (four residue tensor maps; forward operand in TMEM, push staging) ------------------------------
@pytest.mark.parametrize("math", [sps.MATH_F16X3, sps.MATH_F16, sps.MATH_BF16X3])
@pytest.mark.parametrize("shape", [(2, 9, 7, 228, 19, 0), (3, 13, 11, 228, 19, 1), (1, 16, 16, 228, 19, 3), (2, 23, 17, 760, 19, 2),
                                   (1, 12, 11, 24, 2, 1)])
def test_fp32_maps_by_tma_any_alignment(built_lib, math, shape):
    """The fp32 forward reads the map through tensor maps whose base is the channel row rounded DOWN to 16 bytes: every
    residue of (H*W mod 4) and every 4-byte alignment of the map pointer (a view that starts 1..3 floats into a buffer) has
    to give the oracle's distances, activations and logits; H*W below, equal to and above one 128-position tile; dense and
    fused outputs; the push scan on the same views."""
    B, H, W, P, K, off = shape
    z, protos, ranges, w, dl = _case(B, H, W, P, K, seed=71 + off)
    lo, do, ao, *_ = _oracle(z, protos, ranges, w, dl)
    buf = torch.zeros(z.numel() + 8, device=DEV)
    zv = buf[off:off + z.numel()].view(z.shape)
    zv.copy_(z.to(DEV))
    assert zv.data_ptr() % 16 == (4 * off) % 16 and zv.is_contiguous()
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=math)
    pd, wd = protos.to(DEV), w.contiguous().to(DEV)
    before = _lib.load().sps_fp32_reroute_count()
    logits, dist, act = sps.head_forward_raw(zv, pd, wd, cfg, True, True)
    tol = TOL[math]
    assert_rel(dist.cpu(), do, tol, "distances")
    assert_rel(act.cpu().view(B, H, W, P), ao.view(B, H, W, P), tol, "activations")
    assert_rel(logits.cpu().view(B, H, W, K), lo, tol, "logits")
    fused, dnone, anone = sps.head_forward_raw(zv, pd, wd, cfg, False, False)
    assert dnone is None and anone is None
    assert_rel(fused.cpu().view(B, H, W, K), lo, tol, "logits (fused)")
    assert _lib.load().sps_fp32_reroute_count() == before
    # push scan on the same view: winners must be positions of the right class at the oracle's distance
    ident = ho.make_identity(P, K, 4)
    labels = torch.stack([ho.make_labels(i, H, W, K, absent_class=1 if K > 2 else None) for i in range(B)]).to(torch.int16)
    keys = spush.scan_maps(zv, labels.to(DEV), pd, spush.proto_class_of(ident).to(DEV), cfg).cpu()
    dmin = (keys >> 32).to(torch.int32).view(torch.float32)     # float bits of the minimum distance (1e10 = absent)
    pix = keys & 0xFFFFFFFF
    pc = spush.proto_class_of(ident)
    for b in range(B):
        lab = labels[b].view(-1).to(torch.int64)
        dmap = do[b].detach().view(P, -1)
        for p in range(P):
            m = lab == (int(pc[p]) + 1)
            if not bool(m.any()):
                assert float(dmin[b, p]) >= 1e9
                continue
            best = float(dmap[p][m].min())
            scale = max(float(dmap[p].abs().max()), 1.0)
            assert abs(float(dmin[b, p]) - best) <= 2 * tol * scale, (b, p)
            assert bool(m[int(pix[b, p])]), (b, p)
            assert abs(float(dmap[p, int(pix[b, p])]) - best) <= 2 * tol * scale, (b, p)

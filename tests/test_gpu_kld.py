"""GPU parity of the fused KLD loss (sps_kld_fwd / sps_kld_bwd through the C ABI) against (i) golden vectors produced
by the reference's KLDLoss (segmentation/model/loss.py:53-146, oracle/gen_golden.py) and (ii) the CPU oracle on
seeded inputs.

Tolerances: loss |err| <= 1e-4 * max(|ref|, 1e-3) (fp32 sums in a different order than the reference's);
gradient |err| <= 1e-3 * max|ref| per tensor (the north_star's gradient bound).
"""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
from scaleprotoseg_b200.kld import KldTables, kld_loss
from test_gpu_parity import assert_grad

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _load(golden_dir, name):
    return {k: v for k, v in np.load(os.path.join(golden_dir, name + ".npz")).items()}


def _close_loss(got, ref):
    got, ref = float(torch.as_tensor(got).detach()), float(torch.as_tensor(ref).detach())
    assert abs(got - ref) <= 1e-4 * max(abs(ref), 1e-3), f"loss {got} vs {ref}"


def _close_grad(got, ref, tol=1e-3):
    got, ref = got.detach().cpu().double(), torch.as_tensor(ref).double()
    scale = ref.abs().max().clamp_min(1e-30)
    err = ((got - ref).abs() / scale).max().item()
    assert err <= tol, f"ddist: max err {err:.3e} of the largest entry"


@pytest.mark.parametrize("name", ["kld_em", "kld_small_ragged", "kld_city228"])
def test_kld_golden(built_lib, golden_dir, name):
    g = _load(golden_dir, name)
    ranges = [tuple(int(v) for v in r) for r in g["ranges"]]
    tables = KldTables(torch.from_numpy(g["identity"]), ranges, DEV)
    dist = torch.from_numpy(g["dist"]).to(DEV).requires_grad_(True)
    loss = kld_loss(dist, torch.from_numpy(g["labels"]).to(DEV), tables)
    _close_loss(loss, g["loss"])
    (ddist,) = torch.autograd.grad(loss * float(g["upstream"]), dist)
    _close_grad(ddist, g["ddist"])
    # zero exactly where the pixel's class is not the prototype's
    lab = torch.from_numpy(g["labels"]).reshape(dist.shape[0], 1, -1) - 1
    pc = torch.from_numpy(g["identity"]).argmax(dim=1).reshape(1, -1, 1)
    off_class = (lab != pc).reshape(ddist.shape)
    assert float(ddist.cpu()[off_class].abs().max()) == 0.0


def _seeded(B, H, W, P, K, S, seed, block=8, spread=6.0, absent=None):
    g = torch.Generator().manual_seed(ho.SEED + seed)
    dist = torch.rand(B, P, H, W, generator=g) * spread
    labels = torch.stack([ho.make_labels(1000 * seed + i, H, W, K, absent_class=absent, block=block) for i in range(B)])
    return dist, labels, ho.make_identity(P, K, S), ho.uniform_ranges(P, S)


@pytest.mark.parametrize("shape", [(2, 65, 65, 760, 19, 4), (1, 33, 41, 228, 19, 4), (3, 17, 9, 40, 5, 4), (1, 21, 21, 684, 171, 2)])
def test_kld_vs_oracle(built_lib, shape):
    B, H, W, P, K, S = shape
    dist, labels, ident, ranges = _seeded(B, H, W, P, K, S, seed=P + H, absent=2)
    dref = dist.clone().requires_grad_(True)
    lref = ho.kld_loss(dref, labels, ident, ranges)
    (gref,) = torch.autograd.grad(lref * 0.7, dref)
    d = dist.to(DEV).requires_grad_(True)
    loss = kld_loss(d, labels.to(DEV), KldTables(ident, ranges, DEV))
    _close_loss(loss, lref.detach())
    (gd,) = torch.autograd.grad(loss * 0.7, d)
    _close_grad(gd, gref)


def test_kld_large_distances_do_not_overflow(built_lib):
    """log_softmax of distances in the hundreds (exp would overflow without the running maximum)."""
    dist, labels, ident, ranges = _seeded(1, 16, 16, 40, 5, 4, seed=3, spread=300.0)
    lref = ho.kld_loss(dist.double(), labels, ident, ranges)
    loss = kld_loss(dist.to(DEV), labels.to(DEV), KldTables(ident, ranges, DEV))
    assert torch.isfinite(loss).item()
    _close_loss(loss, lref)


def test_kld_no_contribution(built_lib):
    """Void-only labels and labels above K: loss 0, gradient 0 (loss.py:90-91, 142-146)."""
    dist, labels, ident, ranges = _seeded(2, 8, 8, 40, 5, 4, seed=4)
    labels[0] = 0
    labels[1] = 200
    d = dist.to(DEV).requires_grad_(True)
    loss = kld_loss(d, labels.to(DEV), KldTables(ident, ranges, DEV))
    assert float(loss.detach()) == 0.0
    (gd,) = torch.autograd.grad(loss, d)
    assert float(gd.abs().max()) == 0.0


def test_kld_full_size_properties(built_lib):
    """BASELINE.json configs[1] size (B=16, 65x65, P=760, K=19), where the oracle would take minutes:
    (i) log_softmax is shift-invariant: adding a constant per (image, prototype) row leaves the loss unchanged;
    (ii) the gradient of every row sums to zero over the map (softmax Jacobian), and is zero off-class;
    (iii) identical distance rows inside a group give kld = 0 for its pairs, i.e. loss exactly 1 when all rows agree."""
    B, H, W, P, K, S = 16, 65, 65, 760, 19, 4
    dist, labels, ident, ranges = _seeded(B, H, W, P, K, S, seed=9)
    tables = KldTables(ident, ranges, DEV)
    d = dist.to(DEV).requires_grad_(True)
    lab = labels.to(DEV)
    loss = kld_loss(d, lab, tables)
    shift = torch.rand(B, P, 1, 1, device=DEV) * 3.0
    loss_shift = kld_loss(d.detach() + shift, lab, tables)
    assert abs(float(loss) - float(loss_shift)) <= 1e-4 * max(float(loss), 1e-3)
    (gd,) = torch.autograd.grad(loss, d)
    rows = gd.reshape(B, P, -1)
    assert float(rows.sum(dim=2).abs().max()) <= 1e-3 * float(rows.abs().sum(dim=2).max())
    pc = ident.argmax(dim=1).to(DEV).reshape(1, P, 1)
    off_class = (lab.reshape(B, 1, -1) - 1) != pc
    assert float(rows[off_class].abs().max()) == 0.0
    same = torch.rand(B, 1, H, W, device=DEV).expand(B, P, H, W).contiguous()
    assert abs(float(kld_loss(same, lab, tables)) - 1.0) <= 1e-5


class _KLDLossStandin(nn.Module):
    """Same attributes as the reference's KLDLoss (loss.py:56-64); its own forward is the CPU oracle."""

    def __init__(self, prototype_class_identity, num_scales, scale_num_prototypes):
        super().__init__()
        self.prototype_class_identity = prototype_class_identity
        self.num_scales = num_scales
        self.scale_num_prototypes = scale_num_prototypes

    def forward(self, prototype_distances, target_labels):
        ranges = [self.scale_num_prototypes[s] for s in range(self.num_scales)]
        return ho.kld_loss(prototype_distances, target_labels, self.prototype_class_identity, ranges)


def test_kld_dropin_through_install(built_lib):
    """install_kld binds the kernels over KLDLoss.forward: CUDA distances run the fused path, CPU distances keep the
    original loop; a stage-1 style step (head forward in dense mode -> KLD -> backward into z and prototypes) matches."""
    B, H, W, P, K, S, D = 2, 13, 11, 228, 19, 4, 64
    ident, ranges = ho.make_identity(P, K, S), ho.uniform_ranges(P, S)
    labels = torch.stack([ho.make_labels(50 + i, H, W, K, block=4) for i in range(B)])
    torch.manual_seed(ho.SEED)
    z = torch.sigmoid(torch.randn(B, S * D, H, W))
    protos = torch.rand(P, D, 1, 1)
    last_w = ho.make_last_layer(ident)
    mod = _KLDLossStandin(ident, S, {s: r for s, r in enumerate(ranges)})
    zr, pr = z.clone().requires_grad_(True), protos.clone().requires_grad_(True)
    _, dist_ref, _ = ho.head_stage1(zr, pr, ranges, last_w)
    loss_ref = mod(prototype_distances=dist_ref, target_labels=labels)
    loss_ref.backward()
    sps.install_kld(_KLDLossStandin)
    try:
        assert hasattr(_KLDLossStandin.forward, "__sps_original__")
        _close_loss(mod(prototype_distances=dist_ref.detach(), target_labels=labels), loss_ref.detach())   # CPU: original
        cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_FP32)
        zc, pc = z.to(DEV).requires_grad_(True), protos.to(DEV).requires_grad_(True)
        _, dist, _ = sps.PrototypeHeadFn.apply(zc, pc, last_w.to(DEV), cfg, True, True)
        loss = mod(prototype_distances=dist, target_labels=labels.to(DEV))
        assert loss.is_cuda
        _close_loss(loss, loss_ref.detach())
        loss.backward()
        _close_grad(zc.grad, zr.grad, tol=2e-3)
        _close_grad(pc.grad, pr.grad, tol=2e-3)
    finally:
        sps.uninstall()
    assert not hasattr(_KLDLossStandin.forward, "__sps_original__")


# ---- feature mode: the loss straight from z and the prototypes (no distance map) --------------------------------
def _feature_case(B, H, W, P, K, S, D, seed, prune=(), block=4):
    gen = torch.Generator().manual_seed(ho.SEED + seed)
    z = torch.sigmoid(torch.randn(B, S * D, H, W, generator=gen))
    protos = torch.rand(P, D, 1, 1, generator=gen)
    ident, ranges = ho.make_identity(P, K, S), ho.uniform_ranges(P, S)
    if prune:
        keep = [p for p in range(P) if p not in set(prune)]
        new, lo = [], 0
        for a, b in ranges:
            n = sum(1 for p in keep if a <= p < b)
            new.append((lo, lo + n))
            lo += n
        ident, protos, ranges = ident[keep], protos[keep].contiguous(), new
    labels = torch.stack([ho.make_labels(77 * seed + i, H, W, K, absent_class=1, block=block) for i in range(B)])
    return z, protos, ident, ranges, labels


@pytest.mark.parametrize("case", [dict(B=2, H=13, W=11, P=228, K=19, S=4, D=64, seed=1),
                                  dict(B=1, H=33, W=20, P=760, K=19, S=4, D=64, seed=2, block=8),
                                  dict(B=3, H=9, W=10, P=40, K=5, S=4, D=64, seed=3, prune=(0, 1, 11, 39)),
                                  dict(B=2, H=8, W=9, P=48, K=4, S=2, D=24, seed=4),
                                  dict(B=1, H=12, W=12, P=190, K=19, S=1, D=64, seed=5)])
def test_kld_from_features_vs_oracle(built_lib, case):
    """kld_loss_from_features(z, prototypes, labels) against the oracle's head distances fed to the oracle's KLD loss,
    with the gradients autograd takes back into z and the prototypes (ragged groups, S = 1 and 2, D = 24)."""
    z, protos, ident, ranges, labels = _feature_case(**case)
    K = ident.shape[1]
    zr, pr = z.clone().requires_grad_(True), protos.clone().requires_grad_(True)
    lref = ho.kld_loss(ho.scale_l2_convolution(zr, pr, ranges), labels, ident, ranges)
    lref.backward()
    cfg = sps.HeadConfig.from_ranges(ranges, K, math=sps.MATH_FP32)
    tables = KldTables(ident, ranges, DEV)
    zc, pc = z.to(DEV).requires_grad_(True), protos.to(DEV).requires_grad_(True)
    loss = sps.kld_loss_from_features(zc, pc, labels.to(DEV), tables, cfg)
    _close_loss(loss, lref.detach())
    (loss * 1.0).backward()
    _close_grad(zc.grad, zr.grad), _close_grad(pc.grad, pr.grad)


def test_kld_from_features_matches_dense_path_at_full_size(built_lib):
    """configs[1] size: feature mode against the dense path of this library (head distances -> sps_kld_fwd/bwd ->
    head backward through ddist), loss and both gradients; fused-mode training step = head (no map) + this loss."""
    z, protos, ident, ranges, labels = _feature_case(16, 65, 65, 760, 19, 4, 64, seed=6, block=8)
    cfg = sps.HeadConfig.from_ranges(ranges, 19, math=sps.MATH_FP32)
    tables = KldTables(ident, ranges, DEV)
    lab = labels.to(DEV)
    z1, p1 = z.to(DEV).requires_grad_(True), protos.to(DEV).requires_grad_(True)
    _, dist, _ = sps.PrototypeHeadFn.apply(z1, p1, None, cfg, True, False)
    l1 = kld_loss(dist, lab, tables)
    l1.backward()
    z2, p2 = z.to(DEV).requires_grad_(True), protos.to(DEV).requires_grad_(True)
    l2 = sps.kld_loss_from_features(z2, p2, lab, tables, cfg)
    l2.backward()
    _close_loss(l2, l1.detach())
    _close_grad(z2.grad, z1.grad.cpu()), _close_grad(p2.grad, p1.grad.cpu())
    void = (labels == 0).unsqueeze(1).expand_as(z)
    assert float(z2.grad.cpu()[void].abs().max()) == 0.0


def test_unchanged_trainer_step_in_fused_mode(built_lib):
    """The lines of the reference's training step that touch the head (module_multiscale.py:209-242, shipped config:
    entropy and norm weights 0): outputs = ppnet.forward(image, return_activations=False); cross entropy on
    outputs[0]; kld_loss_func(prototype_distances=outputs[1], target_labels=target).  With set_deferred_distances
    the same lines run without a (B,P,H,W) tensor and give the same loss and gradients as the CPU classes."""
    import torch.nn.functional as F
    from standin_model import Stage1Standin
    from test_gpu_dropin import _patched
    sub = _patched(Stage1Standin)
    sps.install_on_classes(sub)
    sps.install_kld(_KLDLossStandin)
    old_math = sps.default_math()
    sps.set_default_math(sps.MATH_F16X3)
    sps.set_deferred_distances(True)
    try:
        P, K, S = 228, 19, 4
        torch.manual_seed(11)
        ref = Stage1Standin(P, K)
        got = sub(P, K)
        got.load_state_dict(ref.state_dict())
        got = got.to(DEV).train()
        ref.train()
        x = torch.randn(2, 256, 14, 9)
        target = torch.stack([ho.make_labels(300 + i, 14, 9, K, block=4) for i in range(2)])

        def step(model, kld_mod, x, target):
            outputs = model.forward(x, return_activations=False)
            output, prototype_distances = outputs[0], outputs[1]
            ce = F.cross_entropy(output.reshape(-1, K), (target.reshape(-1) - 1).clamp_min(0))
            kld = kld_mod(prototype_distances=prototype_distances, target_labels=target)
            (ce + 0.25 * kld).backward()
            return ce.detach(), kld.detach(), prototype_distances

        kld_ref = _KLDLossStandin(ref.prototype_class_identity, S, ref.scale_num_prototypes)
        xr = x.clone().requires_grad_(True)
        ce_r, kld_r, _ = step(ref, kld_ref, xr, target)
        xg = x.to(DEV).requires_grad_(True)
        ce_g, kld_g, handle = step(got, kld_ref, xg, target.to(DEV))
        assert isinstance(handle, sps.DeferredDistances) and handle._dense is None      # never materialised
        assert tuple(handle.shape) == (2, P, 14, 9)
        _close_loss(kld_g, kld_r)
        assert abs(float(ce_g) - float(ce_r)) <= 1e-3 * max(abs(float(ce_r)), 1.0)
        _close_grad(xg.grad, xr.grad, tol=2e-3)
        _close_grad(got.prototype_vectors.grad, ref.prototype_vectors.grad, tol=2e-3)
        _close_grad(got.last_layer.weight.grad, ref.last_layer.weight.grad, tol=2e-3)
        # any other use of the handle still behaves like the tensor
        got.eval()
        with torch.no_grad():
            dense = got.forward(x.to(DEV))[1]
        assert isinstance(dense, torch.Tensor)
        assert torch.allclose(handle.detach().cpu(), dense.cpu(), rtol=1e-3, atol=1e-3)
    finally:
        sps.set_deferred_distances(False)
        sps.set_default_math(old_math)
        sps.uninstall()


# ---- KLDLossGroup (segmentation/model/loss.py:467-545): the stage-2 loss over the group activations -----------------------
@pytest.mark.parametrize("name", ["kld_group_em", "kld_group_city_ragged"])
def test_kld_group_golden(built_lib, golden_dir, name):
    """Value and gradient against the fixtures the unmodified reference produced (class without prototypes, class absent
    from the labels, void-only image), through the same kernels as KLDLoss with the group activations as rows."""
    g = _load(golden_dir, name)
    J = int(g["J"])
    gident = torch.from_numpy(g["group_identity"])
    tables = sps.KldTables(gident, [(0, gident.shape[0])], DEV)
    acts = torch.from_numpy(g["acts"]).to(DEV).requires_grad_(True)
    labels = torch.from_numpy(g["labels"]).to(DEV)
    loss = sps.kld_group_loss(list(torch.split(acts, J, dim=-1)), labels, tables)
    assert abs(float(loss) - float(g["loss"])) <= 1e-5 * max(float(g["loss"]), 1e-3)
    (da,) = torch.autograd.grad(loss * float(g["upstream"]), acts)
    assert_grad(da, g["dacts"], 1e-4, "kld group dacts")


def test_kld_group_dropin_on_a_standin(built_lib, golden_dir):
    """install_kld_group binds the fused loss over a class with the reference's three attributes (loss.py:470-476);
    CPU inputs keep the class's own loop."""
    g = _load(golden_dir, "kld_group_city_ragged")
    J = int(g["J"])

    class KLDLossGroupStandin(torch.nn.Module):
        def __init__(self, prototype_class_identity, group_class_identity, num_groups):
            super().__init__()
            self.prototype_class_identity, self.group_class_identity, self.num_groups = prototype_class_identity, group_class_identity, num_groups

        def forward(self, list_group_activation, target_labels):
            return ho.kld_group_loss(torch.cat(list_group_activation, -1), target_labels, self.group_class_identity)

    sps.install_kld_group(KLDLossGroupStandin)
    mod = KLDLossGroupStandin(torch.from_numpy(g["proto_identity"]), torch.from_numpy(g["group_identity"]), J)
    acts = torch.from_numpy(g["acts"])
    cpu = mod(list(torch.split(acts, J, dim=-1)), torch.from_numpy(g["labels"]))
    gpu = mod(list(torch.split(acts.to(DEV), J, dim=-1)), torch.from_numpy(g["labels"]).to(DEV))
    assert gpu.is_cuda and abs(float(gpu) - float(cpu)) <= 1e-5 * max(float(cpu), 1e-3)
    assert abs(float(gpu) - float(g["loss"])) <= 1e-5 * max(float(g["loss"]), 1e-3)

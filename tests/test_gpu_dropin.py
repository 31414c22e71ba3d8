"""Drop-in behaviour on the GPU: the patched methods keep the reference's signatures, return
conventions and autograd behaviour (stand-in classes with the reference's attribute names, see
tests/standin_model.py), plus edge cases and full-size properties."""
import copy
import os

import pytest
import torch

import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
from standin_model import Stage1Standin, Stage2Standin
from test_gpu_parity import DEV, assert_grad, assert_rel

pytestmark = pytest.mark.gpu
TOL = 5e-5
IMPORT_TIME_DEFAULT = sps.default_math()   # before the autouse fixture below switches to the exact mode


@pytest.fixture(autouse=True)
def _exact_math():
    """The drop-in behaviour tests run the exact-operand mode so they can hold the tight bound; the default
    (tensor-core) mode has its own test below."""
    old = sps.default_math()
    sps.set_default_math(sps.MATH_FP32)
    yield
    sps.set_default_math(old)


def _patched(cls):
    """A patched subclass so the un-patched class keeps serving as the CPU oracle."""
    sub = type(cls.__name__ + "Patched", (cls,), {})
    # copy the methods onto the subclass so install_on_classes wraps them there
    for name in ("_scale_l2_convolution", "_l2_convolution", "prototype_distances", "push_forward",
                 "forward_from_conv_features", "compute_group", "run_last_layer"):
        if hasattr(cls, name):
            setattr(sub, name, getattr(cls, name))
    return sub


@pytest.fixture(scope="module")
def patched_classes(built_lib):
    s1, s2 = _patched(Stage1Standin), _patched(Stage2Standin)
    sps.install_on_classes(s1, s2)
    return s1, s2


def _pair(cls_cpu, cls_gpu, *args, **kw):
    torch.manual_seed(7)
    ref = cls_cpu(*args, **kw)
    got = cls_gpu(*args, **kw)
    got.load_state_dict(ref.state_dict())
    return ref, got.to(DEV)


@pytest.mark.parametrize("flags", [(False, False), (True, False), (True, True), (False, True)])
def test_stage1_return_conventions(patched_classes, flags):
    ref, got = _pair(Stage1Standin, patched_classes[0], 24, 2)
    x = torch.randn(2, 256, 7, 5)
    ra, rd = flags
    o_ref = ref.forward(x, return_activations=ra, return_distances=rd)
    o_got = got.forward(x.to(DEV), return_activations=ra, return_distances=rd)
    assert len(o_ref) == len(o_got)
    for a, b in zip(o_got, o_ref):
        assert a.is_cuda and a.dtype == torch.float32
        assert_rel(a, b, TOL, "output")


def test_stage1_list_input_and_training_step(patched_classes):
    """MSC returns a LIST of maps of different sizes in train mode (utils.py:108-109); the head maps
    over it (model_multiscale.py:335-336, 358-359).  Then one optimiser-free training step: CE-like
    loss on logits + a loss straight on the distances (as KLDLoss does, module_multiscale.py:242)."""
    ref, got = _pair(Stage1Standin, patched_classes[0], 228, 19)
    xs = [torch.randn(1, 256, 9, 9), torch.randn(1, 256, 5, 7), torch.randn(1, 256, 3, 3)]
    o_ref = ref.forward_from_conv_features([ref.add_on_layers(x) for x in xs])
    o_got = got.forward_from_conv_features([got.add_on_layers(x.to(DEV)) for x in xs])
    assert isinstance(o_got, list) and len(o_got) == 3
    for (lg, dg), (lr, dr) in zip(o_got, o_ref):
        assert_rel(lg, lr, TOL, "logits"), assert_rel(dg, dr, TOL, "dist")

    def loss_of(m, x):
        logits, d = m.forward(x)
        return torch.logsumexp(logits, -1).mean() + 1e-3 * d.clamp(max=5).mean()

    loss_of(ref, xs[0]).backward()
    loss_of(got, xs[0].to(DEV)).backward()
    assert_grad(got.prototype_vectors.grad, ref.prototype_vectors.grad, 1e-3, "dprotos")
    assert_grad(got.last_layer.weight.grad, ref.last_layer.weight.grad, 1e-3, "dlast")


def test_stage1_ragged_after_prune(patched_classes):
    ref, got = _pair(Stage1Standin, patched_classes[0], 228, 19)
    prune = [0, 1, 60, 113, 114, 200, 227]
    ref.prune_prototypes(prune), got.prune_prototypes(prune)
    x = torch.randn(1, 256, 6, 6)
    (lr, dr), (lg, dg) = ref.forward(x), got.forward(x.to(DEV))
    assert dg.shape[1] == 221
    assert_rel(lg, lr, TOL, "logits"), assert_rel(dg, dr, TOL, "dist")
    assert_rel(got.prototype_distances(x.to(DEV)), ref.prototype_distances(x), TOL, "prototype_distances")
    c, d = got.push_forward(x.to(DEV))
    assert_rel(c, torch.sigmoid(x), 1e-6, "conv_output"), assert_rel(d, dr, TOL, "push_forward")


def test_stage2_dropin(patched_classes):
    ref, got = _pair(Stage2Standin, patched_classes[1], 228, 19)
    x = torch.randn(2, 256, 5, 4)
    lr, dr, ar = ref.forward(x, return_activations=True, return_distances=True)
    lg, dg, ag = got.forward(x.to(DEV), return_activations=True, return_distances=True)
    assert_rel(lg, lr, 4 * TOL, "logits"), assert_rel(dg, dr, TOL, "dist"), assert_rel(ag, ar, TOL, "act")
    gr, gg = ref.compute_group(ar), got.compute_group(ag)
    assert len(gr) == len(gg) == 19 and gg[0].shape == (40, 3)
    for a, b in zip(gg, gr):
        assert_rel(a, b, 4 * TOL, "group")
    (lr.square().mean() + torch.cat(gr, -1).mean()).backward()
    (lg.square().mean() + torch.cat(gg, -1).mean()).backward()
    for a, b in zip(got.group_projection, ref.group_projection):
        assert_grad(a.weight.grad, b.weight.grad, 1e-3, "dG")
    assert_grad(got.last_layer_group.weight.grad, ref.last_layer_group.weight.grad, 1e-3, "dWg")
    assert_grad(got.prototype_vectors.grad, ref.prototype_vectors.grad, 1e-3, "dprotos")


def test_unsupported_config_keeps_original(patched_classes):
    ref, got = _pair(Stage1Standin, patched_classes[0], 24, 2)
    got.prototype_activation_function = lambda d: 1.0 / (1.0 + d)   # callable activation: not covered
    assert not sps.model_api.is_supported(got)


def test_cpu_input_raises(patched_classes):
    _, got = _pair(Stage1Standin, patched_classes[0], 24, 2)
    with pytest.raises(RuntimeError, match="CUDA only"):
        got.forward_from_conv_features(torch.rand(1, 256, 3, 3))


def test_state_dict_unchanged_by_install(patched_classes):
    ref, got = _pair(Stage1Standin, patched_classes[0], 24, 2)
    got.forward(torch.randn(1, 256, 3, 3, device=DEV))
    assert sorted(ref.state_dict()) == sorted(got.state_dict())
    assert set(vars(got)) == set(vars(ref))      # no new attributes (whole-module pickles stay loadable)


@pytest.mark.parametrize("shape", [(0, 4, 4), (1, 1, 1), (1, 1, 63), (3, 1, 65), (1, 13, 5)])
def test_edge_shapes(built_lib, shape):
    B, H, W = shape
    P, K = 24, 2
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K)
    z = torch.sigmoid(torch.randn(B, 256, H, W))
    p = torch.rand(P, 64)
    w = ho.make_last_layer(ho.make_identity(P, K, 4))
    lg, dg, ag = sps.head_forward_raw(z.to(DEV), p.to(DEV), w.to(DEV), cfg, True, True)
    lr, dr, ar = ho.head_stage1(z, p.view(P, 64, 1, 1), ho.uniform_ranges(P, 4), w)
    assert_rel(lg.view(lr.shape), lr, TOL, "logits"), assert_rel(dg, dr, TOL, "dist"), assert_rel(ag, ar, TOL, "act")


def test_pushed_prototype_gives_zero_distance(built_lib):
    """After push a prototype equals a feature vector exactly: relu clamps d at 0 and the log
    similarity peaks at log(1/eps) (SURVEY 8c edge case)."""
    P, K = 24, 2
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K)
    z = torch.sigmoid(torch.randn(1, 256, 4, 4))
    p = torch.rand(P, 64)
    p[7] = z[0, 64:128, 2, 3]      # prototype 7 belongs to scale 1
    _, dg, ag = sps.head_forward_raw(z.to(DEV), p.to(DEV), None, cfg, True, True)
    assert dg[0, 7, 2, 3].item() <= 2e-5
    assert ag[2 * 4 + 3, 7].item() > 7.0


def test_full_size_properties(built_lib):
    """BASELINE config C2 (B=16, 65x65, P=760, K=19): size-independent checks instead of a CPU replay.
    (1) logits == act @ W^T recomputed by torch from the returned activations; (2) act == f(dist)
    after the layout permute; (3) fused mode (no maps written) gives the same logits; (4) a random
    subset of positions replayed through the oracle."""
    B, H, W, P, K = 16, 65, 65, 760, 19
    g = torch.Generator().manual_seed(ho.SEED)
    z = torch.sigmoid(torch.randn(B, 256, H, W, generator=g)).to(DEV)
    p = torch.rand(P, 64, generator=g).to(DEV)
    ident = ho.make_identity(P, K, 4)
    w = ho.make_last_layer(ident).to(DEV)
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, 4), K)
    lg, dg, ag = sps.head_forward_raw(z, p, w, cfg, True, True)
    lf, _, _ = sps.head_forward_raw(z, p, w, cfg, False, False)
    assert torch.equal(lg, lf)
    assert_rel(lg, ag @ w.t(), 2e-4, "logits vs act@W")
    a2 = torch.log((dg + 1) / (dg + 1e-4)).permute(0, 2, 3, 1).reshape(-1, P)
    assert_rel(ag, a2, 1e-5, "act vs f(dist)")
    sel = torch.randint(0, B * H * W, (257,), generator=g)
    zc = z.cpu().permute(0, 2, 3, 1).reshape(-1, 256)[sel].t().reshape(1, 256, 1, 257)
    lr, dr, ar = ho.head_stage1(zc, p.cpu().view(P, 64, 1, 1), ho.uniform_ranges(P, 4), w.cpu())
    assert_rel(lg[sel.to(DEV)], lr.view(257, K), TOL, "logits subset")
    assert_rel(ag[sel.to(DEV)], ar, TOL, "act subset")


def test_default_math_is_tensor_core_and_training_step_matches(patched_classes):
    """install() defaults to the tcgen05 kernels (3-term fp16 products).  One training-style step through the
    patched model: logits-only loss -> fused forward + tensor-core backward; the parameter gradients match the
    un-patched CPU model within the north-star tolerance."""
    assert IMPORT_TIME_DEFAULT == sps.MATH_F16X3
    sps.set_default_math(sps.MATH_F16X3)
    ref, got = _pair(Stage1Standin, patched_classes[0], 228, 19)
    x = torch.randn(2, 256, 9, 9)
    lr, dr = ref.forward(x)
    lg, dg = got.forward(x.to(DEV))
    assert_rel(lg, lr, 1e-3, "logits"), assert_rel(dg, dr, 1e-3, "dist")
    target = torch.randint(0, 19, (2, 9, 9))
    torch.nn.functional.cross_entropy(lr.permute(0, 3, 1, 2), target).backward()
    torch.nn.functional.cross_entropy(lg.permute(0, 3, 1, 2), target.to(DEV)).backward()
    assert_grad(got.prototype_vectors.grad, ref.prototype_vectors.grad, 1e-3, "dprotos")
    assert_grad(got.last_layer.weight.grad, ref.last_layer.weight.grad, 1e-3, "dlast")


# ---- single-scale ProtoSeg baseline (segmentation/model/model.py:74) ------------------------------
@pytest.fixture(scope="module")
def patched_single(built_lib):
    from standin_model import SingleScaleStandin
    sub = _patched(SingleScaleStandin)
    sps.install_on_classes(single_cls=sub)
    return SingleScaleStandin, sub


@pytest.mark.parametrize("return_activations", [False, True])
def test_single_scale_forward_and_step(patched_single, return_activations):
    """PPNet.forward_from_conv_features (model.py:306-330): (logits, distances) or (logits, activations), and the
    gradients of one CE-like step into the features, the prototypes and the last layer."""
    base, sub = patched_single
    ref, got = _pair(base, sub, 57, 19)
    x = torch.randn(2, 64, 9, 6)
    xr, xg = x.clone().requires_grad_(True), x.to(DEV).requires_grad_(True)
    o_ref = ref.forward(xr, return_activations=return_activations)
    o_got = got.forward(xg, return_activations=return_activations)
    assert len(o_ref) == len(o_got) == 2
    for a, b in zip(o_got, o_ref):
        assert a.is_cuda and a.shape == b.shape
        assert_rel(a, b, TOL, "output")
    w = torch.randn(o_ref[0].shape) / o_ref[0].numel()
    (o_ref[0] * w).sum().backward()
    (o_got[0] * w.to(DEV)).sum().backward()
    for a, b, name in [(xg.grad, xr.grad, "dx"), (got.prototype_vectors.grad, ref.prototype_vectors.grad, "dprotos"),
                       (got.last_layer.weight.grad, ref.last_layer.weight.grad, "dW")]:
        scale = b.abs().max().clamp_min(1e-30)
        assert_rel(a.cpu() / scale, b / scale, 20 * TOL, name)


def test_single_scale_list_push_forward_and_image_level_branch(patched_single):
    base, sub = patched_single
    ref, got = _pair(base, sub, 40, 5)
    maps = [torch.randn(1, 64, 5, 4), torch.randn(1, 64, 3, 7)]
    for (lr, dr), (lg, dg) in zip(ref.forward_from_conv_features([torch.sigmoid(m) for m in maps]),
                                  got.forward_from_conv_features([torch.sigmoid(m).to(DEV) for m in maps])):
        assert_rel(lg, lr, TOL, "list logits"), assert_rel(dg, dr, TOL, "list distances")
    cr, dr = ref.push_forward(maps[0])
    cg, dg = got.push_forward(maps[0].to(DEV))
    assert_rel(cg, cr, TOL, "conv_output"), assert_rel(dg, dr, TOL, "push distances")
    assert_rel(got.prototype_distances(maps[1].to(DEV)), ref.prototype_distances(maps[1]), TOL, "prototype_distances")
    # image-level (original ProtoPNet) head: not covered by the kernels' fused path, keeps the class's own code,
    # which still gets its distances from the patched _l2_convolution
    ref2, got2 = _pair(base, sub, 40, 5, patch_classification=False)
    o_ref = ref2.forward(maps[0], return_distances=True)
    o_got = got2.forward(maps[0].to(DEV), return_distances=True)
    assert len(o_got) == 3
    for a, b in zip(o_got, o_ref):
        assert_rel(a, b, TOL, "image-level output")


# ---- push drop-ins (install.make_push_functions): compute_distances / min_across_dataset / global_min ---------------------
class _FakePushDataset:
    """What the three functions read from a PatchClassificationDataset (push_multiscale_optimization.py:116-131)."""
    convert_targets = None
    image_margin_size = 0

    def __init__(self, n, ann_dir):
        self.img_ids = [f"img{i:03d}" for i in range(n)]
        self.annotations_dir = ann_dir

    def __len__(self):
        return len(self.img_ids)


def _push_fixture_env(golden_dir, name, tmp_path, patched_cls):
    import types
    import numpy as np
    from test_gpu_parity import _load
    g = _load(golden_dir, name)
    n = g["x"].shape[0]
    ds = _FakePushDataset(n, str(tmp_path))
    for i, iid in enumerate(ds.img_ids):
        np.save(os.path.join(str(tmp_path), iid + ".npy"), g["labels"][i].astype(np.int64))
    maps = {iid: torch.from_numpy(g["x"][i]) for i, iid in enumerate(ds.img_ids)}
    # the same stand-ins oracle/gen_golden.py::push_case used with the reference's own functions: a pass-through
    # to_normalized_tensor (the "image" is the pre-sigmoid map) and a same-size label resize
    pm = types.SimpleNamespace(to_normalized_tensor=lambda t: t,
                               resize_label=lambda t, size: torch.from_numpy(np.asarray(t)).long())
    P, K = g["protos"].shape[0], int(g["K"])
    torch.manual_seed(0)
    model = patched_cls(P, K).to(DEV)
    model.prototype_vectors.data.copy_(torch.from_numpy(g["protos"]))
    from scaleprotoseg_b200.install import make_push_functions
    fns = make_push_functions(pm, load_image=lambda dataset, iid: maps[iid])
    return g, ds, pm, model, fns


@pytest.mark.parametrize("name", ["push_small", "push_city228"])
def test_push_dropins_golden(patched_classes, golden_dir, tmp_path, name):
    """The installed replacements of compute_distances / min_across_dataset / global_min
    (segmentation/push_multiscale_optimization.py:35-190) driven the way push_prototypes_multiscale drives them (:286-293),
    against the fixtures the reference's own functions produced: per-image indices and values, winning image per prototype
    (first image wins a tie, void-only image, class absent from the dataset) and the projected prototypes."""
    import numpy as np
    sps.set_default_math(sps.PUSH_MATH)       # what install() runs the scan in
    g, ds, pm, model, (compute_distances, min_across_dataset, global_min) = _push_fixture_env(golden_dir, name, tmp_path, patched_classes[0])
    K = int(g["K"])
    n = len(ds)
    for i in range(n):
        idx, val = compute_distances(model, ds, torch.from_numpy(g["x"][i]), g["labels"][i].astype(np.int64), num_classes=K,
                                     void_class=0, device=DEV)
        assert idx.shape == (1, g["idx"].shape[1]) and idx.dtype == torch.int64 and val.dtype == torch.float32
        ref_val = torch.from_numpy(g["val"][i])
        absent = ref_val >= 1e10
        assert torch.equal(val.cpu()[0] >= 1e10, absent)
        assert_rel(val.cpu()[0][~absent], ref_val[~absent], 1e-3, "compute_distances values")
    proto_min_dist, list_idx = min_across_dataset(ds, model, K, void_class=0, device=DEV)
    assert proto_min_dist.dtype == torch.int64 and len(list_idx) == n and list_idx[0].shape == (1, g["idx"].shape[1])
    assert torch.equal(proto_min_dist.cpu(), torch.from_numpy(g["win"]))
    got_idx = torch.cat(list_idx).cpu()
    ref_idx = torch.from_numpy(g["idx"])
    win = torch.from_numpy(g["win"])
    cols = torch.arange(win.numel())
    live = torch.from_numpy(g["val"])[win, cols] < 1e10
    assert torch.equal(got_idx[win, cols][live], ref_idx[win, cols][live])      # the winning positions
    patches = global_min(proto_min_dist, list_idx, ds, model, device=DEV)
    assert len(patches) == win.numel() and patches[0].shape == (64, 1, 1) and patches[0].dtype == np.float32
    new = np.reshape(patches, g["new_protos"].shape)
    assert np.allclose(new[live.numpy()], g["new_protos"][live.numpy()], rtol=0, atol=1e-6)

"""The CPU oracle (oracle/head_oracle.py) against vectors produced by the UNMODIFIED reference
(oracle/gen_golden.py, run in the build container).  fp32 on both sides, same op order: tight."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import head_oracle as ho

TOL = dict(rtol=2e-5, atol=2e-6)


def _load(golden_dir, name):
    return {k: v for k, v in np.load(os.path.join(golden_dir, name + ".npz")).items()}


def _ranges(g):
    return [tuple(int(v) for v in r) for r in g["ranges"]]


@pytest.mark.parametrize("name", ["stage1_em", "stage1_em_ragged", "stage1_linear", "stage1_city228"])
def test_stage1_forward_backward(golden_dir, name):
    g = _load(golden_dir, name)
    z = torch.from_numpy(g["z"]).requires_grad_(True)
    p = torch.from_numpy(g["protos"]).requires_grad_(True)
    w = torch.from_numpy(g["last_w"]).requires_grad_(True)
    logits, d, a = ho.head_stage1(z, p, _ranges(g), w, kind=str(g["act_kind"]))
    np.testing.assert_allclose(d.detach().numpy(), g["dist"], **TOL)
    np.testing.assert_allclose(a.detach().numpy(), g["act"], **TOL)
    np.testing.assert_allclose(logits.detach().numpy(), g["logits"], rtol=2e-5, atol=2e-5)
    loss = (logits * torch.from_numpy(g["dlogits"])).sum() + (d * torch.from_numpy(g["ddist"])).sum() + \
        (a * torch.from_numpy(g["dact"])).sum()
    dz, dp, dw = torch.autograd.grad(loss, [z, p, w])
    np.testing.assert_allclose(dz.numpy(), g["dz"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(dp.numpy(), g["dprotos"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(dw.numpy(), g["dlast_w"], rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("name", ["stage2_em", "stage2_city228_ragged"])
def test_stage2_forward_backward(golden_dir, name):
    g = _load(golden_dir, name)
    z = torch.from_numpy(g["z"]).requires_grad_(True)
    p = torch.from_numpy(g["protos"]).requires_grad_(True)
    wg = torch.from_numpy(g["last_group_w"]).requires_grad_(True)
    gws = [torch.from_numpy(g[f"group_w_{k}"]).requires_grad_(True) for k in range(int(g["n_group_mods"]))]
    ident = torch.from_numpy(g["identity"])
    logits, d, a, groups = ho.head_stage2(z, p, _ranges(g), ident, gws, wg)
    np.testing.assert_allclose(d.detach().numpy(), g["dist"], **TOL)
    np.testing.assert_allclose(groups.detach().numpy(), g["groups"], rtol=5e-5, atol=1e-6)
    np.testing.assert_allclose(logits.detach().numpy(), g["logits"], rtol=5e-5, atol=5e-5)
    loss = (logits * torch.from_numpy(g["dlogits"])).sum() + (d * torch.from_numpy(g["ddist"])).sum() + \
        (a * torch.from_numpy(g["dact"])).sum() + (groups * torch.from_numpy(g["dgroups"])).sum()
    grads = torch.autograd.grad(loss, [z, p, wg] + gws)
    np.testing.assert_allclose(grads[0].numpy(), g["dz"], rtol=2e-4, atol=1e-6)
    np.testing.assert_allclose(grads[1].numpy(), g["dprotos"], rtol=2e-4, atol=1e-6)
    np.testing.assert_allclose(grads[2].numpy(), g["dlast_group_w"], rtol=2e-4, atol=1e-6)
    for k in range(len(gws)):
        np.testing.assert_allclose(grads[3 + k].numpy(), g[f"dgroup_w_{k}"], rtol=2e-4, atol=1e-6)


@pytest.mark.parametrize("name", ["push_small", "push_city228"])
def test_push(golden_dir, name):
    g = _load(golden_dir, name)
    K = int(g["K"])
    x = torch.from_numpy(g["x"])
    z_maps = [torch.sigmoid(x[i:i + 1]) for i in range(x.shape[0])]
    labels = [torch.from_numpy(g["labels"][i].astype(np.int64)) for i in range(x.shape[0])]
    win, idx, val, newp = ho.push_full(z_maps, labels, torch.from_numpy(g["protos"]), _ranges(g),
                                       torch.from_numpy(g["identity"]), K)
    assert np.array_equal(torch.cat(idx).numpy(), g["idx"])      # bit-exact indices
    assert np.array_equal(torch.cat(val).numpy(), g["val"])      # bit-exact fp32 minima
    assert np.array_equal(win.numpy(), g["win"])
    assert np.array_equal(newp, g["new_protos"])


def test_golden_files_present(golden_dir):
    assert len(glob.glob(os.path.join(golden_dir, "*.npz"))) >= 8


KLD_CASES = ["kld_em", "kld_small_ragged", "kld_city228"]


@pytest.mark.parametrize("name", KLD_CASES)
def test_kld_loss_and_gradient(golden_dir, name):
    """ho.kld_loss against KLDLoss.forward of the reference (segmentation/model/loss.py:66-146): value, and the
    gradient autograd takes through the reference's masked_select / log_softmax / kl_div chain."""
    g = _load(golden_dir, name)
    dist = torch.from_numpy(g["dist"]).requires_grad_(True)
    loss = ho.kld_loss(dist, torch.from_numpy(g["labels"]), torch.from_numpy(g["identity"]), _ranges(g))
    np.testing.assert_allclose(loss.detach().numpy(), g["loss"], rtol=1e-5, atol=1e-7)
    (ddist,) = torch.autograd.grad(loss * float(g["upstream"]), dist)
    np.testing.assert_allclose(ddist.numpy(), g["ddist"], rtol=1e-4, atol=1e-7)


@pytest.mark.parametrize("name", ["kld_group_em", "kld_group_city_ragged"])
def test_kld_group_loss_and_gradient(golden_dir, name):
    """KLDLossGroup.forward (segmentation/model/loss.py:467-545) run by the unmodified reference on random group
    activations (class without prototypes, class absent from the labels, void-only image) vs the oracle's restatement."""
    g = _load(golden_dir, name)
    acts = torch.from_numpy(g["acts"]).requires_grad_(True)
    loss = ho.kld_group_loss(acts, torch.from_numpy(g["labels"]), torch.from_numpy(g["group_identity"]))
    np.testing.assert_allclose(loss.detach().numpy(), g["loss"], rtol=1e-5, atol=1e-7)
    (da,) = torch.autograd.grad(loss * float(g["upstream"]), acts)
    np.testing.assert_allclose(da.numpy(), g["dacts"], rtol=1e-4, atol=1e-7 * float(np.abs(g["dacts"]).max()))


def test_topk_ranking_from_the_reference(golden_dir):
    """find_k_nearest_patches_to_prototypes (find_nearest.py:71-352) run by the unmodified reference on a small on-disk
    dataset: the ranked image list per prototype (ascending distance of the image's closest non-void patch)."""
    g = _load(golden_dir, "topk_small")
    ranges = _ranges(g)
    protos = torch.from_numpy(g["protos"])
    maps = [ho.scale_l2_convolution(torch.sigmoid(torch.from_numpy(x)).unsqueeze(0), protos, ranges) for x in g["x"]]
    val, img, pix = ho.topk_nearest(maps, [torch.from_numpy(l.astype(np.int64)) for l in g["labels"]], int(g["k"]))
    assert np.array_equal(img.numpy(), g["rank_img"])
    assert (val[:, 1:] >= val[:, :-1]).all() and (val < 1e6).all()


def test_kld_no_contribution_is_zero():
    """Void-only labels, or one pixel per class: no pair contributes and the loss is 0 (loss.py:142-146)."""
    ident, ranges = ho.make_identity(24, 2, 4), ho.uniform_ranges(24, 4)
    dist = torch.rand(1, 24, 3, 3)
    assert float(ho.kld_loss(dist, torch.zeros(1, 3, 3, dtype=torch.int64), ident, ranges)) == 0.0
    one_each = torch.zeros(1, 3, 3, dtype=torch.int64)
    one_each[0, 0, 0], one_each[0, 2, 2] = 1, 2
    assert float(ho.kld_loss(dist, one_each, ident, ranges)) == 0.0

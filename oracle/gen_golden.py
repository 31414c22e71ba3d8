"""Generate tests/golden/*.npz by running the UNMODIFIED reference (imported from /root/reference).

Run in the build container only (the GPU box has no /root/reference):
    python oracle/gen_golden.py
The three absent packages (gin, pytorch_lightning, matplotlib) are satisfied by tests/_stubs.
TEST INFRASTRUCTURE ONLY.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def import_reference():
    for k in ["DATA_PATH_CITY", "DATA_PATH_PASCAL", "DATA_PATH_ADE", "DATA_PATH_EM", "DATA_PATH_COCO", "LOG_DIR"]:
        os.environ.setdefault(k, "/tmp/sps_ref_env")
    os.makedirs("/tmp/sps_ref_env", exist_ok=True)
    for p in (REF, os.path.join(ROOT, "tests", "_stubs")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from segmentation.model import model_multiscale as mm
    from segmentation.model import model_multiscale_group as mg
    import segmentation.push_multiscale_optimization as pm
    return mm, mg, pm


import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402

sys.path.insert(0, ROOT)
from oracle import head_oracle as ho  # noqa: E402


class DeepLabStub(nn.Module):
    """Only there so the reference ctor finds two Conv2d under a DEEPLAB* name
    (model_multiscale.py:162-165); replaced by Identity right after construction."""

    def __init__(self, c):
        super().__init__()
        self.a = nn.Conv2d(3, c, 1)
        self.b = nn.Conv2d(c, c, 1)


def build_stage1(mm, P, K, S=4, D=64, act="log"):
    m = mm.PPNetMultiScale(features=DeepLabStub(S * D), img_size=65, prototype_shape=(P, D, 1, 1),
                           proto_layer_rf_info=[], num_classes=K, init_weights=True,
                           prototype_activation_function=act, add_on_layers_type="deeplab_simple",
                           patch_classification=True, num_scales=S)
    m.features = nn.Identity()
    return m


def build_stage2(mg, P, K, S=4, D=64, num_groups=3):
    m = mg.PPNetMultiScale(features=DeepLabStub(S * D), img_size=65, prototype_shape=(P, D, 1, 1),
                           proto_layer_rf_info=[], num_classes=K, init_weights=True,
                           prototype_activation_function="log", add_on_layers_type="deeplab_simple",
                           patch_classification=True, num_scales=S, num_groups=num_groups)
    m.features = nn.Identity()
    return m


def ranges_of(m):
    return np.array([m.scale_num_prototypes[s] for s in range(m.num_scales)], dtype=np.int64)


def stage1_case(mm, name, B, H, W, P, K, prune=(), act="log", seed=0):
    g = torch.Generator().manual_seed(ho.SEED + seed)
    m = build_stage1(mm, P, K, act=act)
    if prune:
        m.prune_prototypes(list(prune))
    x = torch.randn(B, 256, H, W, generator=g)  # PRE-sigmoid; add_on_layers = Sigmoid
    z = torch.sigmoid(x).detach().requires_grad_(True)
    m.add_on_layers = nn.Sequential()  # feed post-sigmoid z so dz is w.r.t. the head input
    logits, d, a = m.forward_from_conv_features(z, return_activations=True, return_distances=True)
    Pn = m.num_prototypes
    N = B * H * W
    dlogits = torch.randn(logits.shape, generator=g) / N
    ddist = torch.randn(d.shape, generator=g) * 1e-3
    dact = torch.randn(a.shape, generator=g) * 1e-3
    loss = (logits * dlogits).sum() + (d * ddist).sum() + (a * dact).sum()
    dz, dp, dW = torch.autograd.grad(loss, [z, m.prototype_vectors, m.last_layer.weight])
    np.savez_compressed(
        os.path.join(ROOT, "tests", "golden", name + ".npz"),
        z=z.detach().numpy(), protos=m.prototype_vectors.detach().numpy(), ranges=ranges_of(m),
        identity=m.prototype_class_identity.numpy(), last_w=m.last_layer.weight.detach().numpy(),
        logits=logits.detach().numpy(), dist=d.detach().numpy(), act=a.detach().numpy(),
        dlogits=dlogits.numpy(), ddist=ddist.numpy(), dact=dact.numpy(),
        dz=dz.numpy(), dprotos=dp.numpy(), dlast_w=dW.numpy(), act_kind=np.array(act))
    print(name, "P", Pn, "ranges", ranges_of(m).tolist())


def stage2_case(mm, mg, name, B, H, W, P, K, prune=(), drop_class=None, seed=0):
    g = torch.Generator().manual_seed(ho.SEED + 100 + seed)
    m = build_stage2(mg, P, K)
    if prune or drop_class is not None:
        # Same transplant as finetune_wandb_group.py:74-87: stage-1 state with a ragged layout.
        s1 = build_stage1(mm, P, K)
        pr = set(prune)
        if drop_class is not None:
            pr |= set(torch.nonzero(s1.prototype_class_identity[:, drop_class]).flatten().tolist())
        s1.prune_prototypes(sorted(pr))
        m.prototype_vectors = nn.Parameter(s1.prototype_vectors.data.clone())
        m.ones = nn.Parameter(s1.ones.data.clone(), requires_grad=False)
        m.scale_num_prototypes = dict(s1.scale_num_prototypes)
        m.prototype_class_identity = s1.prototype_class_identity.clone()
        m._initialize_groups()
        m._initialize_weights()
    # non-trivial simplex weights
    for gp in m.group_projection:
        from segmentation.utils import projection_simplex_sort
        gp.weight.data = projection_simplex_sort(torch.rand(gp.weight.shape, generator=g))
    m.add_on_layers = nn.Sequential()
    z = torch.sigmoid(torch.randn(B, 256, H, W, generator=g)).requires_grad_(True)
    logits, d, a = m.forward_from_conv_features(z, return_activations=True, return_distances=True)
    groups = torch.cat(m.compute_group(a), dim=-1)
    N = B * H * W
    dlogits = torch.randn(logits.shape, generator=g) / N
    ddist = torch.randn(d.shape, generator=g) * 1e-3
    dact = torch.randn(a.shape, generator=g) * 1e-3
    dgroups = torch.randn(groups.shape, generator=g) * 1e-3
    loss = (logits * dlogits).sum() + (d * ddist).sum() + (a * dact).sum() + (groups * dgroups).sum()
    params = [z, m.prototype_vectors, m.last_layer_group.weight] + [gp.weight for gp in m.group_projection]
    grads = torch.autograd.grad(loss, params)
    out = dict(
        z=z.detach().numpy(), protos=m.prototype_vectors.detach().numpy(), ranges=ranges_of(m),
        identity=m.prototype_class_identity.numpy(), last_group_w=m.last_layer_group.weight.detach().numpy(),
        logits=logits.detach().numpy(), dist=d.detach().numpy(), act=a.detach().numpy(), groups=groups.detach().numpy(),
        dlogits=dlogits.numpy(), ddist=ddist.numpy(), dact=dact.numpy(), dgroups=dgroups.numpy(),
        dz=grads[0].numpy(), dprotos=grads[1].numpy(), dlast_group_w=grads[2].numpy(),
        n_group_mods=np.array(len(m.group_projection)))
    for k, gp in enumerate(m.group_projection):
        out[f"group_w_{k}"] = gp.weight.detach().numpy()
        out[f"dgroup_w_{k}"] = grads[3 + k].numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), **out)
    print(name, "P", m.num_prototypes, "mods", len(m.group_projection))


class _FakeDataset:
    convert_targets = None


def push_case(mm, pm, name, n_img, h, w, P, K, absent_class, dup=False, seed=0):
    """Runs the reference's own compute_distances (push_multiscale_optimization.py:35-91) per map,
    with `to_normalized_tensor` rebound to a pass-through so a (C,h,w) pre-sigmoid map stands in
    for the PIL image, then :135-137 and the gather lines :182-188."""
    m = build_stage1(mm, P, K)
    m.eval()
    pm.to_normalized_tensor = lambda t: t
    xs, labels, list_idx, list_val = [], [], [], []
    for i in range(n_img):
        g = torch.Generator().manual_seed(ho.SEED + 104729 * (i + 1) + seed)
        x = torch.randn(256, h, w, generator=g)
        if dup and i == 1:
            x = xs[0].clone()  # identical map -> exact cross-image ties, first image must win
        lab = ho.make_labels(i + seed, h, w, K, absent_class=absent_class)
        if i == n_img - 1:
            lab = torch.zeros_like(lab)  # void-only image
        idx, val = pm.compute_distances(m, _FakeDataset(), x, lab.numpy(), num_classes=K, void_class=0, device="cpu")
        xs.append(x), labels.append(lab), list_idx.append(idx), list_val.append(val)
    win = torch.concat(list_val, dim=0).argmin(dim=0)
    n_per_scale = P // m.num_scales
    patches = []
    for p in range(P):
        cf = m.conv_features(xs[int(win[p])].unsqueeze(0))
        _, C, H, W = cf.shape
        cf = cf.view(m.num_scales, C // m.num_scales, H, W)
        flat = list_idx[int(win[p])][:, p].item()
        i, j = flat // cf.shape[-1], flat % cf.shape[-1]
        patches.append(cf[p // n_per_scale, :, i:i + 1, j:j + 1].detach().numpy())
    new_protos = np.reshape(patches, (P, 64, 1, 1))
    np.savez_compressed(
        os.path.join(ROOT, "tests", "golden", name + ".npz"),
        x=torch.stack(xs).numpy(), labels=torch.stack(labels).numpy().astype(np.int16),
        protos=m.prototype_vectors.detach().numpy(), ranges=ranges_of(m), identity=m.prototype_class_identity.numpy(),
        idx=torch.cat(list_idx).numpy(), val=torch.cat(list_val).numpy(), win=win.numpy(), new_protos=new_protos,
        K=np.array(K))
    print(name, "absent->1e10:", int((torch.cat(list_val).min(0).values >= 1e10).sum()))


def kld_case(name, B, H, W, P, K, S=4, prune=(), absent_class=None, void_only=False, seed=0):
    """Reference KLDLoss (segmentation/model/loss.py:53-146) on a random distance map; value and gradient."""
    from segmentation.model.loss import KLDLoss
    g = torch.Generator().manual_seed(ho.SEED + 31 * seed)
    ident = ho.make_identity(P, K, S)
    ranges = ho.uniform_ranges(P, S)
    if prune:
        keep = [p for p in range(P) if p not in set(prune)]
        ident = ident[keep]
        new, lo = [], 0
        for a, b in ranges:
            n = sum(1 for p in keep if a <= p < b)
            new.append((lo, lo + n))
            lo += n
        ranges, P = new, len(keep)
    dist = (torch.rand(B, P, H, W, generator=g) * 6.0).requires_grad_(True)
    labels = torch.stack([ho.make_labels(seed * 100 + i, H, W, K, absent_class=absent_class, block=3) for i in range(B)])
    if void_only:
        labels[0] = 0
    ref = KLDLoss(ident, len(ranges), {s: tuple(r) for s, r in enumerate(ranges)})
    loss = ref(dist, labels)
    if loss.requires_grad:
        (ddist,) = torch.autograd.grad(loss * 3.0, dist)
    else:
        ddist = torch.zeros_like(dist)
    mine = ho.kld_loss(dist.detach(), labels, ident, ranges)
    lv = float(loss.detach())
    assert abs(float(mine) - lv) <= 1e-5 * max(1.0, abs(lv)), (name, float(mine), lv)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), dist=dist.detach().numpy(),
                        labels=labels.numpy(), identity=ident.numpy(), ranges=np.array(ranges, dtype=np.int64),
                        loss=np.float32(loss.detach().item()), ddist=ddist.numpy(), upstream=np.float32(3.0))
    print(name, "loss", lv)


def kld_group_case(name, B, H, W, K, J=3, per_class=3, drop_class=None, absent_class=None, void_only=False, seed=0):
    """Reference KLDLossGroup (segmentation/model/loss.py:467-545) on random group activations; value and gradient."""
    from segmentation.model.loss import KLDLossGroup
    g = torch.Generator().manual_seed(ho.SEED + 57 * (seed + 1))
    classes = [c for c in range(K) if c != drop_class]           # a class without prototypes has no projection (:486-487)
    pident = torch.zeros(per_class * len(classes), K)
    gident = torch.zeros(J * len(classes), K)
    for k, c in enumerate(classes):
        pident[k * per_class:(k + 1) * per_class, c] = 1
        gident[k * J:(k + 1) * J, c] = 1                          # model_multiscale_group.py:262-267
    N = B * H * W
    acts = torch.exp(torch.randn(N, J * len(classes), generator=g)).requires_grad_(True)   # compute_group ends in exp
    labels = torch.stack([ho.make_labels(seed * 100 + i, H, W, K, absent_class=absent_class, block=3) for i in range(B)])
    if void_only:
        labels[0] = 0
    ref = KLDLossGroup(pident, gident, J)
    loss = ref(list(torch.split(acts, J, dim=-1)), labels)
    if loss.requires_grad:
        (dacts,) = torch.autograd.grad(loss * 2.0, acts)
    else:
        dacts = torch.zeros_like(acts)
    mine = ho.kld_group_loss(acts.detach(), labels, gident)
    lv = float(loss.detach())
    assert abs(float(mine) - lv) <= 1e-5 * max(1.0, abs(lv)), (name, float(mine), lv)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), acts=acts.detach().numpy(), labels=labels.numpy(),
                        proto_identity=pident.numpy(), group_identity=gident.numpy(), J=np.array(J),
                        loss=np.float32(lv), dacts=dacts.numpy(), upstream=np.float32(2.0))
    print(name, "loss", lv)


def topk_case(mm, name, n_img, h, w, P, K, k, seed=0):
    """Runs the reference's own find_k_nearest_patches_to_prototypes (find_nearest.py:71-352) on a small dataset ON DISK
    (PNG images + .npy annotations in a temp directory).  to_normalized_tensor is rebound so that the PNG, whose first
    pixel carries the image number, stands for a pre-sigmoid feature map; Tensor.cuda is a no-op here (no GPU).  The
    function returns, per prototype, the annotation maps of the k nearest images in ascending distance: matching them
    against the (distinct) annotations gives the ranked image list, which is what pins the oracle."""
    import tempfile
    import types
    from PIL import Image
    import find_nearest as fn
    m = build_stage1(mm, P, K)
    m.eval()
    xs, labels = [], []
    tmp = tempfile.mkdtemp(prefix="sps_topk_")
    ids = [f"im{i}" for i in range(n_img)]
    for i in range(n_img):
        g = torch.Generator().manual_seed(ho.SEED + 7001 * (i + 1) + seed)
        xs.append(torch.randn(256, h, w, generator=g))
        lab = ho.make_labels(i + seed, h, w, K, block=3)
        lab[0, 0] = 0                                            # every image holds a void pixel (the +10e6 mask, :138)
        labels.append(lab)
        px = np.zeros((h, w, 3), dtype=np.uint8)
        px[0, 0, 0] = i
        Image.fromarray(px).save(os.path.join(tmp, ids[i] + ".png"))
        np.save(os.path.join(tmp, ids[i] + ".npy"), lab.numpy())
    ds = types.SimpleNamespace(img_ids=ids, get_img_path=lambda iid: os.path.join(tmp, iid + ".png"), image_margin_size=0,
                               annotations_dir=tmp, convert_targets=None, __len__=lambda: n_img)

    class _DS:
        img_ids, image_margin_size, annotations_dir, convert_targets = ids, 0, tmp, None

        def get_img_path(self, iid):
            return os.path.join(tmp, iid + ".png")

        def __len__(self):
            return n_img

    old_norm, old_cuda = fn.to_normalized_tensor, torch.Tensor.cuda
    fn.to_normalized_tensor = lambda img: xs[img.getpixel((0, 0))[0]]
    torch.Tensor.cuda = lambda self, *a, **kw: self
    try:
        out = fn.find_k_nearest_patches_to_prototypes(_DS(), types.SimpleNamespace(module=m, eval=lambda: None), k=k,
                                                      full_save=False, log=lambda *a: None)
    finally:
        fn.to_normalized_tensor, torch.Tensor.cuda = old_norm, old_cuda
    shifted = [lab.numpy() - 1 for lab in labels]                  # the function keeps search_y - 1 (:127)
    rank = np.full((P, k), -1, dtype=np.int64)
    for p in range(P):
        for r in range(k):
            hits = [i for i in range(n_img) if np.array_equal(out[p][r], shifted[i])]
            assert len(hits) == 1, (p, r, hits)
            rank[p, r] = hits[0]
    ranges = ranges_of(m)
    mine = ho.topk_nearest([ho.scale_l2_convolution(torch.sigmoid(x).unsqueeze(0), m.prototype_vectors.detach(), [tuple(r) for r in ranges])
                            for x in xs], labels, k)
    assert np.array_equal(mine[1].numpy(), rank), name
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), x=torch.stack(xs).numpy(),
                        labels=torch.stack(labels).numpy().astype(np.int16), protos=m.prototype_vectors.detach().numpy(),
                        ranges=ranges, identity=m.prototype_class_identity.numpy(), rank_img=rank, K=np.array(K), k=np.array(k))
    print(name, "ranked images of prototype 0:", rank[0].tolist())


def main():
    mm, mg, pm = import_reference()
    torch.manual_seed(ho.SEED)
    os.makedirs(os.path.join(ROOT, "tests", "golden"), exist_ok=True)
    # EM-config shape (configs/scaleproto_em.gin: 24 prototypes, 2 classes)
    stage1_case(mm, "stage1_em", B=2, H=9, W=7, P=24, K=2)
    stage1_case(mm, "stage1_em_ragged", B=1, H=5, W=6, P=24, K=2, prune=(1, 7, 8, 20), seed=1)
    stage1_case(mm, "stage1_linear", B=1, H=4, W=5, P=24, K=2, act="linear", seed=2)
    # Cityscapes shipped shape, tiny map
    stage1_case(mm, "stage1_city228", B=1, H=6, W=5, P=228, K=19, seed=3)
    stage2_case(mm, mg, "stage2_em", B=2, H=5, W=4, P=24, K=2)
    stage2_case(mm, mg, "stage2_city228_ragged", B=1, H=4, W=5, P=228, K=19, prune=(0, 5, 100, 227), drop_class=4, seed=1)
    push_case(mm, pm, "push_small", n_img=5, h=9, w=11, P=40, K=5, absent_class=3, dup=True)
    push_case(mm, pm, "push_city228", n_img=3, h=17, w=9, P=228, K=19, absent_class=7, seed=5)
    kld_case("kld_em", B=2, H=9, W=7, P=24, K=2)
    kld_case("kld_small_ragged", B=2, H=7, W=8, P=40, K=5, prune=(0, 1, 11, 39), absent_class=3, void_only=True, seed=1)
    kld_case("kld_city228", B=1, H=12, W=10, P=228, K=19, seed=2)
    kld_group_case("kld_group_em", B=2, H=9, W=7, K=2)
    kld_group_case("kld_group_city_ragged", B=2, H=8, W=9, K=19, drop_class=4, absent_class=3, void_only=True, seed=1)
    topk_case(mm, "topk_small", n_img=7, h=9, w=11, P=40, K=5, k=3)


if __name__ == "__main__":
    main()

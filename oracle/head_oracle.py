"""CPU oracle for the ScaleProtoSeg prototype head and push scan.  TEST INFRASTRUCTURE ONLY.

This file restates, in plain torch-CPU / numpy, the arithmetic of the reference's hot path so that
the CUDA kernels can be checked on a box where /root/reference does not exist.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may import it.
The product package (`scaleprotoseg_b200/`) never does.

Pinning: the reference ships no tests and no golden vectors, so this restatement is
pinned against the reference's own code imported from /root/reference (oracle/gen_golden.py writes
tests/golden/*.npz; tests/test_oracle_golden.py replays them).

Reference lines followed (relative to /root/reference):
  l2_convolution            segmentation/model/model_multiscale.py:256-281
  scale_l2_convolution      segmentation/model/model_multiscale.py:283-317
  distance_2_similarity     segmentation/model/model_multiscale.py:324-330
  head_stage1               segmentation/model/model_multiscale.py:340-388 (+ run_last_layer :243)
  compute_group/head_stage2 segmentation/model/model_multiscale_group.py:283-308, 404-452
  push_image_min            segmentation/push_multiscale_optimization.py:68-91
  push_dataset_argmin       segmentation/push_multiscale_optimization.py:135-137
  push_gather               segmentation/push_multiscale_optimization.py:161-188
  topk_nearest              find_nearest.py:89-175 (pinned by topk_small.npz: the reference's own function run on a small
                            on-disk dataset, oracle/gen_golden.py::topk_case)
  kld_loss                  segmentation/model/loss.py:53-146 (KLDLoss.forward)
  kld_group_loss            segmentation/model/loss.py:467-545 (KLDLossGroup.forward; pinned by kld_group_*.npz)
"""
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

EPSILON = 1e-4  # model_multiscale.py:106


def l2_convolution(x: torch.Tensor, protos: torch.Tensor) -> torch.Tensor:
    """x (B,D,H,W), protos (Ps,D,1,1) -> relu(|x|^2 - 2 x.p + |p|^2) (B,Ps,H,W).

    Same op sequence as model_multiscale.py:267-279 (the all-ones conv computes |x|^2 broadcast
    to every prototype channel)."""
    ones = torch.ones_like(protos)
    x2_patch_sum = F.conv2d(x * x, ones)
    p2 = torch.sum(protos * protos, dim=(1, 2, 3)).view(-1, 1, 1)
    xp = F.conv2d(x, protos)
    return F.relu(x2_patch_sum + (-2 * xp + p2))


def scale_l2_convolution(z: torch.Tensor, protos: torch.Tensor, scale_ranges: Sequence[Tuple[int, int]]) -> torch.Tensor:
    """z (B,C,H,W) with C = S*D; protos (P,D,1,1); scale_ranges[s] = (lo, hi) -> (B,P,H,W).

    model_multiscale.py:295 views channels as (S, D); scale s uses channel block s and prototype
    rows [lo_s, hi_s); outputs are concatenated in scale order (:317)."""
    B, C, H, W = z.shape
    S = len(scale_ranges)
    D = C // S
    zs = z.view(B, S, D, H, W)
    outs = []
    for s in range(S):
        lo, hi = scale_ranges[s]
        outs.append(l2_convolution(zs[:, s], protos[lo:hi]))
    return torch.cat(outs, dim=1)


def distance_2_similarity(d: torch.Tensor, kind: str = "log", eps: float = EPSILON) -> torch.Tensor:
    if kind == "log":
        return torch.log((d + 1) / (d + eps))
    if kind == "linear":
        return -d
    raise ValueError(kind)


def head_stage1(z, protos, scale_ranges, last_w, kind: str = "log"):
    """-> logits (B,H,W,K), distances (B,P,H,W), activations (B*H*W,P)."""
    d = scale_l2_convolution(z, protos, scale_ranges)
    B, P, H, W = d.shape
    act = distance_2_similarity(d.permute(0, 2, 3, 1).contiguous().reshape(-1, P), kind)
    logits = F.linear(act, last_w).reshape(B, H, W, last_w.shape[0])
    return logits, d, act


def class_members(identity: torch.Tensor) -> List[np.ndarray]:
    """Per class WITH prototypes: the prototype rows of that class, ascending
    (model_multiscale_group.py:294-298)."""
    out = []
    for c in range(identity.shape[1]):
        idx = torch.nonzero(identity[:, c]).flatten().cpu().numpy()
        if len(idx) > 0:
            out.append(idx)
    return out


def compute_group(act: torch.Tensor, identity: torch.Tensor, group_ws: Sequence[torch.Tensor]) -> List[torch.Tensor]:
    """group_ws[k] is (num_groups, n_k) for the k-th class that has prototypes."""
    outs = []
    for k, members in enumerate(class_members(identity)):
        outs.append(torch.exp(F.linear(act[:, members], group_ws[k])))
    return outs


def head_stage2(z, protos, scale_ranges, identity, group_ws, last_group_w, kind: str = "log"):
    """-> logits (B,H,W,K), distances, activations, groups (N, sum num_groups)."""
    d = scale_l2_convolution(z, protos, scale_ranges)
    B, P, H, W = d.shape
    act = distance_2_similarity(d.permute(0, 2, 3, 1).contiguous().reshape(-1, P), kind)
    groups = torch.cat(compute_group(act, identity, group_ws), dim=-1)
    logits = F.linear(groups, last_group_w).reshape(B, H, W, last_group_w.shape[0])
    return logits, d, act, groups


# --------------------------------------------------------------------------------------------
# Push / nearest-patch projection (tensor-level restatement; the enclosing reference functions
# take PIL images and a dataset object, the arithmetic lines are followed one for one).
# --------------------------------------------------------------------------------------------
MAX_DIST = 1e10  # push_multiscale_optimization.py:41


def push_image_min(d: torch.Tensor, label: torch.Tensor, identity: torch.Tensor, num_classes: int,
                   void_class: Optional[int] = 0) -> Tuple[torch.Tensor, torch.Tensor]:
    """d (1,P,h,w) fp32 distances of ONE image, label (h,w) int64 already resized to (h,w).

    Follows push_multiscale_optimization.py:73-91: one-hot over K(+1) classes, drop the void
    column, `@ identity.T`-transposed, mask = 1e10*(1-match), min/argmin over the flattened map.
    Returns (indices (1,P) int64, values (1,P) fp32)."""
    class_identity = identity.T.clone().to(torch.float32)  # (K,P)
    t = label.unsqueeze(0)
    one_hot = F.one_hot(t, num_classes=num_classes if void_class is None else num_classes + 1).to(torch.float32)
    if void_class is not None:
        one_hot = torch.cat([one_hot[..., :void_class], one_hot[..., void_class + 1:]], dim=-1)
    m = torch.matmul(one_hot, class_identity)  # (1,h,w,P)
    m = MAX_DIST * (1 - m.permute(0, 3, 1, 2))
    masked = (d + m).flatten(-2, -1)
    r = masked.min(dim=-1)
    return r.indices, r.values


def push_dataset_argmin(list_dist: List[torch.Tensor]) -> torch.Tensor:
    """list of (1,P) per-image minima -> (P,) index of the first image holding the minimum
    (push_multiscale_optimization.py:135-137)."""
    return torch.cat(list_dist, dim=0).argmin(dim=0)


def push_gather(z_of_image, proto_min_img: torch.Tensor, list_idx: List[torch.Tensor], num_scales: int,
                num_prototypes: int) -> List[np.ndarray]:
    """z_of_image(i) -> (1,C,h,w) post-sigmoid feature map of image i.

    push_multiscale_optimization.py:161-188: scale of prototype p is p // (P/S) (uniform ranges
    assumed by the reference), the flat index is unravelled with the map WIDTH, and the D-vector
    of that scale is taken."""
    n_per_scale = num_prototypes // num_scales
    out = []
    for p in range(num_prototypes):
        s = p // n_per_scale
        img = int(proto_min_img[p])
        z = z_of_image(img)
        _, C, H, W = z.shape
        zs = z.view(num_scales, C // num_scales, H, W)
        flat = int(list_idx[img][:, p].item())
        i, j = flat // W, flat % W
        out.append(zs[s, :, i:i + 1, j:j + 1].detach().cpu().numpy())
    return out


def push_full(z_maps: Sequence[torch.Tensor], labels: Sequence[torch.Tensor], protos: torch.Tensor,
              scale_ranges, identity: torch.Tensor, num_classes: int):
    """Whole push over a list of (1,C,h,w) post-sigmoid maps and (h,w) labels.
    Returns (proto_min_img (P,), list_idx, list_val, new_prototypes (P,D,1,1) float32)."""
    list_idx, list_val = [], []
    for z, lab in zip(z_maps, labels):
        d = scale_l2_convolution(z, protos, scale_ranges)
        idx, val = push_image_min(d, lab, identity, num_classes, void_class=0)
        list_idx.append(idx)
        list_val.append(val)
    win = push_dataset_argmin(list_val)
    P = protos.shape[0]
    patches = push_gather(lambda i: z_maps[i], win, list_idx, len(scale_ranges), P)
    new_protos = np.reshape(patches, tuple(protos.shape)).astype(np.float32)
    return win, list_idx, list_val, new_protos


# --------------------------------------------------------------------------------------------
# Synthetic inputs.  Counter-based so every rank / GPU count sees the same data.
# --------------------------------------------------------------------------------------------
SEED = 20220227  # configs/scaleproto_cityscapes.gin:15


def kld_loss(dist: torch.Tensor, target: torch.Tensor, identity: torch.Tensor,
             scale_ranges: Sequence[Tuple[int, int]]) -> torch.Tensor:
    """KLDLoss.forward (segmentation/model/loss.py:66-146) restated with one matrix per (image, class, scale).

    dist (B,P,H,W) distances; target (B,H,W) integer labels, 0 = void, c+1 = class c (loss.py:77);
    identity (P,K) prototype_class_identity.  For every image, every class present in it (loss.py:88-91)
    and every scale (loss.py:103), the class's prototypes of that scale give one distribution each over the
    class's pixels: log_softmax of the DISTANCES along the pixels (loss.py:112-115).  Every unordered pair
    with at least two pixels contributes kld = (KL(a||b) + KL(b||a)) / 2 (loss.py:121-140); the loss is the
    mean of exp(-kld) over all contributions (loss.py:142-146), 0 when there are none."""
    B, P = dist.shape[0], dist.shape[1]
    K = identity.shape[1]
    lab = target.reshape(B, -1).long() - 1
    d = dist.permute(0, 2, 3, 1).reshape(B, -1, P)
    terms = []
    for b in range(B):
        for c in sorted(set(lab[b].tolist())):
            if c < 0 or c >= K:
                continue
            members = torch.nonzero(identity[:, c]).flatten()
            mask = lab[b] == c
            if members.numel() == 0 or int(mask.sum()) < 2:
                continue
            for lo, hi in scale_ranges:
                ms = members[(members >= lo) & (members < hi)]
                if ms.numel() < 2:
                    continue
                logp = torch.log_softmax(d[b][mask][:, ms], dim=0)            # (n, m)
                prob = logp.exp()
                cross = prob.t() @ logp                                        # cross[a, b] = sum_x p_a log p_b
                kl = cross.diagonal().unsqueeze(1) - cross                     # kl[a, b] = KL(a || b)
                sym = 0.5 * (kl + kl.t())
                iu = torch.triu_indices(ms.numel(), ms.numel(), offset=1)
                terms.append(sym[iu[0], iu[1]])
    if not terms:
        return torch.zeros((), dtype=dist.dtype)
    return torch.exp(-torch.cat(terms)).mean()


def kld_group_loss(acts: torch.Tensor, target: torch.Tensor, group_identity: torch.Tensor) -> torch.Tensor:
    """KLDLossGroup.forward (segmentation/model/loss.py:478-545): acts (N, J*Kc) = the concatenated group activations
    (compute_group's list, class-slot order), target (B,H,W) with 0 = void, group_identity (J*Kc, K).  It is KLDLoss with
    the group activations in the role of the distances and one 'scale' owning every column."""
    B = target.shape[0]
    G = acts.shape[1]
    as_map = acts.reshape(B, -1, G).permute(0, 2, 1).reshape(B, G, -1, 1)
    return kld_loss(as_map, target.reshape(B, -1, 1), group_identity, [(0, G)])


def topk_nearest(dist_maps: Sequence[torch.Tensor], labels: Sequence[torch.Tensor], k: int):
    """find_k_nearest_patches_to_prototypes (find_nearest.py:89-175) on resident maps: per image the void pixels get
    +10e6 (:138), the closest patch per prototype is the minimum over the map (:146), and a heap per prototype keeps the
    k images with the smallest such distance (:160-175).  dist_maps[i] (1,P,h,w), labels[i] (h,w) with 0 = void.
    Returns (dist (P,k) ascending, image (P,k), flat position (P,k))."""
    mins, args = [], []
    for d, lab in zip(dist_maps, labels):
        masked = d[0] + 10e6 * (lab == 0).to(d.dtype).unsqueeze(0)
        v, a = masked.flatten(1).min(dim=1)
        mins.append(v), args.append(a)
    mins, args = torch.stack(mins), torch.stack(args)                 # (n, P)
    val, img = torch.topk(mins, k, dim=0, largest=False)              # (k, P)
    return val.t().contiguous(), img.t().contiguous(), torch.gather(args, 0, img).t().contiguous()


def uniform_ranges(P: int, S: int):
    n = P // S
    return [(s * n, (s + 1) * n) for s in range(S)]


def make_identity(P: int, K: int, S: int) -> torch.Tensor:
    """model_multiscale.py:129-141."""
    ident = torch.zeros(P, K)
    per_scale = P // S
    per_cs = P // K // S
    for i in range(S):
        for j in range(K):
            ident[i * per_scale + j * per_cs: i * per_scale + (j + 1) * per_cs, j] = 1
    return ident


def make_last_layer(identity: torch.Tensor, incorrect: float = -0.5) -> torch.Tensor:
    """model_multiscale.py:449-464 -> (K,P)."""
    pos = identity.t()
    return 1.0 * pos + incorrect * (1 - pos)


def make_labels(img_id: int, h: int, w: int, K: int, absent_class: Optional[int] = None, block: int = 8) -> torch.Tensor:
    """Piecewise-constant block labels in {0..K}, 0 = void (~5 %), from seed+img_id."""
    g = torch.Generator().manual_seed(SEED + 7919 * (img_id + 1))
    hb, wb = (h + block - 1) // block, (w + block - 1) // block
    cls = torch.randint(1, K + 1, (hb, wb), generator=g)
    void = torch.rand(hb, wb, generator=g) < 0.05
    cls[void] = 0
    if absent_class is not None:
        cls[cls == absent_class] = (absent_class % K) + 1
    lab = cls.repeat_interleave(block, 0).repeat_interleave(block, 1)[:h, :w]
    return lab.contiguous().to(torch.int64)


def make_feature_map(img_id: int, C: int, h: int, w: int) -> torch.Tensor:
    g = torch.Generator().manual_seed(SEED + 104729 * (img_id + 1))
    return torch.sigmoid(torch.randn(1, C, h, w, generator=g))

"""Host side of the fused KLD loss between same-class prototypes.

Drop-in for `KLDLoss.forward(prototype_distances, target_labels)` (segmentation/model/loss.py:66-146, called from
segmentation/model/module_multiscale.py:242).  The reference walks image x class x scale x prototype-pair in Python
(one host sync per class, about six small kernels per pair); here one autograd Function calls `sps_kld_fwd` /
`sps_kld_bwd` of libscaleproto_sm100.so.  CUDA tensors only; there is no CPU path.
"""
import ctypes as C
from typing import Dict, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import SpsHeadDesc, SpsKldDesc
from .head import _check_cuda_f32, _ptr, _stream


class KldTables:
    """(class, scale) groups and their prototype pairs, derived once from prototype_class_identity (P x K) and
    scale_num_prototypes — what loss.py:92-109 recomputes per image and class."""

    def __init__(self, identity: torch.Tensor, scale_ranges: Sequence[Tuple[int, int]], device):
        ident = identity.detach().to("cpu")
        P, K = ident.shape
        ranges = [(int(a), int(b)) for a, b in scale_ranges]
        S = len(ranges)
        if (ident != 0).sum(dim=1).max().item() > 1 if P else False:
            raise RuntimeError("prototype_class_identity must give every prototype at most one class")
        pclass = torch.where((ident != 0).any(dim=1), (ident != 0).float().argmax(dim=1), torch.full((P,), -1)).to(torch.int32) \
            if P else torch.zeros(0, dtype=torch.int32)
        pgroup = torch.full((P,), -1, dtype=torch.int32)
        plocal = torch.zeros((P,), dtype=torch.int32)
        moff, poff, members, pa, pb = [0], [0], [], [], []
        for c in range(K):
            cls = torch.nonzero(ident[:, c]).flatten().tolist()
            for s, (lo, hi) in enumerate(ranges):
                ms = [p for p in cls if lo <= p < hi]          # loss.py:105-110
                g = c * S + s
                for r, p in enumerate(ms):
                    pgroup[p], plocal[p] = g, r
                members += ms
                for i in range(len(ms)):                       # loss.py:121-127: j, then k > j
                    for j in range(i + 1, len(ms)):
                        pa.append(ms[i]), pb.append(ms[j])
                moff.append(len(members))
                poff.append(len(pa))
        # prototypes outside every scale range or without a class never take part; park them in an extra group of size 1
        self.P, self.K, self.S = P, K, S
        self.G, self.Q = K * S, len(pa)
        i32 = dict(dtype=torch.int32, device=device)
        self.proto_class = pclass.to(device)
        self.mmax = max([b - a for a, b in zip(moff[:-1], moff[1:])] + [1])
        pclass = torch.where(pgroup < 0, torch.full_like(pclass, -1), pclass)   # outside every scale range: never used
        pgroup = torch.where(pgroup < 0, torch.full_like(pgroup, self.G), pgroup)
        moff.append(moff[-1]), poff.append(poff[-1])           # the parking group: no members listed, no pairs
        self.proto_class = pclass.to(device)
        self.proto_group, self.proto_local = pgroup.to(device), plocal.to(device)
        self.group_moff, self.group_poff = torch.tensor(moff, **i32), torch.tensor(poff, **i32)
        self.members = torch.tensor(members + [0] * max(0, 1 - len(members)), **i32)
        self.pair_a = torch.tensor(pa + [0] * max(0, 1 - len(pa)), **i32)
        self.pair_b = torch.tensor(pb + [0] * max(0, 1 - len(pb)), **i32)

    def desc(self) -> SpsKldDesc:
        k = SpsKldDesc()
        k.G, k.Q, k.mmax, k.reserved = self.G, self.Q, self.mmax, 0
        k.proto_class, k.proto_group, k.proto_local = (self.proto_class.data_ptr(), self.proto_group.data_ptr(),
                                                       self.proto_local.data_ptr())
        k.group_moff, k.group_poff, k.members = (self.group_moff.data_ptr(), self.group_poff.data_ptr(),
                                                 self.members.data_ptr())
        k.pair_a, k.pair_b = self.pair_a.data_ptr(), self.pair_b.data_ptr()
        return k


def _desc(B: int, P: int, H: int, W: int, K: int) -> SpsHeadDesc:
    d = SpsHeadDesc()
    d.B, d.H, d.W, d.S, d.D, d.P, d.K = B, H, W, 1, 1, P, K
    d.scale_lo[0] = 0
    for i in range(1, _lib.SPS_MAX_SCALES + 1):
        d.scale_lo[i] = P
    d.act_kind, d.eps, d.math, d.flags = _lib.ACT_LOG, 1e-4, _lib.MATH_FP32, 0
    return d


def labels_int16(target_labels: torch.Tensor, num_classes: int) -> torch.Tensor:
    """loss.py:77 subtracts 1 and skips what falls outside [0, K) (:90-91): 0 = void stays 0, values above K never
    match.  Clamped before the 16-bit cast so that e.g. 255 / 65535 cannot wrap into a class id."""
    t = target_labels.to(torch.int64)
    return torch.where((t >= 1) & (t <= num_classes), t, torch.zeros_like(t)).to(torch.int16).contiguous()


class KldLossFn(torch.autograd.Function):
    """(distances (B,P,H,W), labels int16 (B,H,W)) -> scalar mean of exp(-kld) over the contributing pairs."""

    @staticmethod
    def forward(ctx, dist, labels, tables: KldTables):
        lib = _lib.load()
        _check_cuda_f32("prototype_distances", dist)
        if dist.dim() != 4 or dist.shape[1] != tables.P:
            raise RuntimeError(f"prototype_distances must be (B,{tables.P},H,W), got {tuple(dist.shape)}")
        if labels.dtype != torch.int16 or not labels.is_cuda or labels.numel() != dist.shape[0] * dist.shape[2] * dist.shape[3]:
            raise RuntimeError("labels must be CUDA int16 of shape (B,H,W) matching the distance map")
        dist, labels = dist.contiguous(), labels.contiguous()
        B, P, H, W = dist.shape
        desc, kd = _desc(B, P, H, W, tables.K), tables.desc()
        nbytes = int(lib.sps_kld_workspace_bytes(C.byref(desc), C.byref(kd)))
        ws = torch.empty(max(nbytes, 16), device=dist.device, dtype=torch.uint8)
        out = torch.empty(2, device=dist.device, dtype=torch.float32)
        with torch.cuda.device(dist.device):
            _lib.check(lib.sps_kld_fwd(C.byref(desc), C.byref(kd), _ptr(dist), _ptr(labels), _ptr(ws), nbytes, _ptr(out),
                                       _stream()), "sps_kld_fwd")
        ctx.tables, ctx.nbytes = tables, nbytes
        ctx.save_for_backward(dist, ws)
        ctx.mark_non_differentiable(labels)
        return out[0]

    @staticmethod
    def backward(ctx, dloss):
        lib = _lib.load()
        dist, ws = ctx.saved_tensors
        if not ctx.needs_input_grad[0]:
            return None, None, None
        B, P, H, W = dist.shape
        desc, kd = _desc(B, P, H, W, ctx.tables.K), ctx.tables.desc()
        g = dloss.to(torch.float32).reshape(1).contiguous()
        ddist = torch.empty_like(dist, memory_format=torch.contiguous_format)
        with torch.cuda.device(dist.device):
            _lib.check(lib.sps_kld_bwd(C.byref(desc), C.byref(kd), _ptr(dist), _ptr(ws), ctx.nbytes, _ptr(g), _ptr(ddist),
                                       _stream()), "sps_kld_bwd")
        return ddist, None, None


def kld_loss(dist: torch.Tensor, target_labels: torch.Tensor, tables: KldTables) -> torch.Tensor:
    """Same value as KLDLoss(...)(prototype_distances=dist, target_labels=target_labels) of the reference."""
    return KldLossFn.apply(dist, labels_int16(target_labels, tables.K).to(dist.device), tables)


class KldFeatureLossFn(torch.autograd.Function):
    """(z (B,S*D,H,W), prototype_vectors (P,D,1,1), labels int16) -> the same scalar as KldLossFn on the head's distance
    map, without that map: sps_kld_feat_fwd / sps_kld_feat_bwd ("feature mode").  The gradient flows into z and the
    prototypes directly, so the head itself can run in fused mode (no (B,P,H,W) tensor in HBM) during training."""

    @staticmethod
    def forward(ctx, z, protos, labels, tables: KldTables, cfg):
        from .head import make_desc, _shape_of
        lib = _lib.load()
        _check_cuda_f32("z", z), _check_cuda_f32("prototype_vectors", protos)
        z = z.contiguous()
        p2 = protos.reshape(protos.shape[0], -1).contiguous()
        B, H, W, S, D, P = _shape_of(z, p2, cfg)
        if P != tables.P or S != tables.S or cfg.num_classes != tables.K:
            raise RuntimeError("KLD tables do not match the head configuration")
        if labels.dtype != torch.int16 or not labels.is_cuda or labels.numel() != B * H * W:
            raise RuntimeError("labels must be CUDA int16 of shape (B,H,W) matching the feature map")
        labels = labels.contiguous()
        desc, kd = make_desc(cfg, B, H, W, D, P), tables.desc()
        nbytes = int(lib.sps_kld_feat_workspace_bytes(C.byref(desc), C.byref(kd)))
        ws = torch.empty(max(nbytes, 16), device=z.device, dtype=torch.uint8)
        out = torch.empty(2, device=z.device, dtype=torch.float32)
        with torch.cuda.device(z.device):
            _lib.check(lib.sps_kld_feat_fwd(C.byref(desc), C.byref(kd), _ptr(z), _ptr(p2), _ptr(labels), _ptr(ws), nbytes,
                                            _ptr(out), _stream()), "sps_kld_feat_fwd")
        ctx.tables, ctx.cfg, ctx.nbytes, ctx.proto_shape = tables, cfg, nbytes, protos.shape
        ctx.save_for_backward(z, p2, labels, ws)
        return out[0]

    @staticmethod
    def backward(ctx, dloss):
        from .head import make_desc
        lib = _lib.load()
        z, p2, labels, ws = ctx.saved_tensors
        need_dz, need_dp = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        if not (need_dz or need_dp):
            return None, None, None, None, None
        B, _, H, W = z.shape
        P, D = p2.shape
        desc, kd = make_desc(ctx.cfg, B, H, W, D, P), ctx.tables.desc()
        g = dloss.to(torch.float32).reshape(1).contiguous()
        dz = torch.empty_like(z) if need_dz else None
        dp = torch.empty_like(p2) if need_dp else None
        with torch.cuda.device(z.device):
            _lib.check(lib.sps_kld_feat_bwd(C.byref(desc), C.byref(kd), _ptr(z), _ptr(p2), _ptr(labels), _ptr(ws), ctx.nbytes,
                                            _ptr(g), _ptr(dz), _ptr(dp), _stream()), "sps_kld_feat_bwd")
        return dz, (dp.view(ctx.proto_shape) if dp is not None else None), None, None, None


def kld_loss_from_features(z: torch.Tensor, protos: torch.Tensor, target_labels: torch.Tensor, tables: KldTables,
                           cfg) -> torch.Tensor:
    """KLDLoss(...)(prototype_distances=head(z), target_labels) without materialising prototype_distances."""
    return KldFeatureLossFn.apply(z, protos, labels_int16(target_labels, tables.K).to(z.device), tables, cfg)


import weakref

# loss module -> (signature of what the tables were built from, tables).  Weak keys: an entry dies with its module, so a
# recycled id() / data_ptr() can never serve stale tables.
_TABLE_CACHE: "weakref.WeakKeyDictionary" = weakref.WeakKeyDictionary()


def tables_for(self, device) -> KldTables:
    """Tables of a KLDLoss module (attributes prototype_class_identity, num_scales, scale_num_prototypes; loss.py:56-64),
    rebuilt only when what they were derived from changes."""
    ranges = tuple(tuple(int(v) for v in self.scale_num_prototypes[s]) for s in range(self.num_scales))
    ident = self.prototype_class_identity
    sig = (ranges, tuple(ident.shape), int(ident.data_ptr()), int(ident._version), str(device))
    hit: Optional[Tuple[tuple, KldTables]] = _TABLE_CACHE.get(self)
    if hit is None or hit[0] != sig:
        hit = (sig, KldTables(ident, ranges, device))
        _TABLE_CACHE[self] = hit
    return hit[1]


def kld_supported(self, device) -> bool:
    """The kernels hold one (class, scale) group of at most 32 prototypes per warp; a wider group keeps the reference's loop."""
    return tables_for(self, device).mmax <= 32


def kld_forward(self, prototype_distances: torch.Tensor, target_labels: torch.Tensor) -> torch.Tensor:
    """Bound over KLDLoss.forward by install(): `self` is the reference's KLDLoss module (attributes
    prototype_class_identity, num_scales, scale_num_prototypes; loss.py:56-64)."""
    from .model_api import DeferredDistances
    if not prototype_distances.is_cuda:
        raise RuntimeError("scaleprotoseg_b200 KLD loss runs on CUDA only: there is no CPU path")
    tables = tables_for(self, prototype_distances.device)
    if isinstance(prototype_distances, DeferredDistances):     # fused-mode training: no (B,P,H,W) map exists
        d = prototype_distances
        return kld_loss_from_features(d.z, d.protos, target_labels, tables, d.cfg)
    return kld_loss(prototype_distances.float(), target_labels, tables)


def kld_group_loss(list_group_activation, target_labels: torch.Tensor, tables: KldTables) -> torch.Tensor:
    """Same value as KLDLossGroup(...)(list_group_activation, target_labels) of the reference (loss.py:467-545).

    The loss is KLDLoss with the group activations in the role of the distances: per image and class, log-softmax over the
    class's pixels of each of the class's `num_groups` group columns, symmetric KL of every column pair, mean of exp(-kld).
    So the same kernels run on the (B, J*Kc, HW) view of the concatenated group activations, with tables built from
    group_class_identity (one 'scale' owning every column: the groups of a class are its members, pairs j < k as :521-527).
    list_group_activation: Kc tensors (N, J) in class-slot order (compute_group's output), N = B*H*W."""
    acts = torch.cat(list(list_group_activation), dim=-1) if not torch.is_tensor(list_group_activation) else list_group_activation
    B = target_labels.shape[0]
    G = acts.shape[-1]
    if G != tables.P:
        raise RuntimeError(f"group activations have {G} columns, group_class_identity has {tables.P} rows")
    hw = acts.shape[0] // max(B, 1)
    # (N, G) -> (B, G, HW, 1): the kernels read a channel-major map; a 4-byte-per-element layout change done by torch
    # (plumbing; autograd carries the gradient back through it)
    as_map = acts.float().reshape(B, hw, G).permute(0, 2, 1).contiguous().view(B, G, hw, 1)
    labels = labels_int16(target_labels.reshape(B, hw, 1), tables.K).to(acts.device)
    return KldLossFn.apply(as_map, labels, tables)


def kld_group_forward(self, list_group_activation, target_labels: torch.Tensor) -> torch.Tensor:
    """Bound over KLDLossGroup.forward by install(): `self` is the reference's module (attributes prototype_class_identity,
    group_class_identity, num_groups; loss.py:470-476)."""
    first = list_group_activation[0]
    if not first.is_cuda:
        raise RuntimeError("scaleprotoseg_b200 KLD loss runs on CUDA only: there is no CPU path")
    gident = self.group_class_identity
    sig = (tuple(gident.shape), int(gident.data_ptr()), int(gident._version), str(first.device), int(self.num_groups))
    hit: Optional[Tuple[tuple, KldTables]] = _TABLE_CACHE.get(self)
    if hit is None or hit[0] != sig:
        hit = (sig, KldTables(gident, [(0, int(gident.shape[0]))], first.device))
        _TABLE_CACHE[self] = hit
    return kld_group_loss(list_group_activation, target_labels, hit[1])

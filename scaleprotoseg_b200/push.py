"""Host side of the push / nearest-patch projection.

Tensor-level API (used by the bench and the tests) plus drop-in versions of the three functions of
segmentation/push_multiscale_optimization.py that `push_prototypes_multiscale` looks up as module
globals (:286, :291): `compute_distances` (:35-91), `min_across_dataset` (:94-137), `global_min`
(:141-190).

Multi-GPU: maps are independent, so each rank scans a contiguous slice of the dataset and ONE
min all-reduce over P packed (distance, global map index) int64 keys combines the ranks; the owner
of each winning map contributes its D-vector through one sum all-reduce of a (P, D) buffer that
is zero everywhere else.
"""
import ctypes as C
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from .head import HeadConfig, _ptr, _stream, make_desc, prepare_z

KEY_INIT = 0x7FFFFFFFFFFFFFFF  # > any packed key; positive as int64 so an NCCL int64 min works
ABSENT_BITS = 0x501502F9       # float bits of 1e10f (push_multiscale_optimization.py:41)


def proto_class_of(identity: torch.Tensor) -> torch.Tensor:
    """(P,K) one-hot -> int32 class per prototype."""
    return identity.detach().argmax(dim=1).to(torch.int32)


def proto_scale_reference(P: int, S: int) -> torch.Tensor:
    """Scale whose channel block global_min reads for prototype p: p // (P // S)
    (push_multiscale_optimization.py:161,167 — uniform ranges assumed there)."""
    n = max(P // S, 1)
    return torch.clamp(torch.arange(P) // n, max=S - 1).to(torch.int32)


def scan_maps(z: torch.Tensor, labels: torch.Tensor, protos: torch.Tensor, proto_class: torch.Tensor,
              cfg: HeadConfig, keys_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """z (B,C,h,w) post-sigmoid maps (fp32 NCHW, or fp16 / bf16 channels-last: TMA-fed), labels (B,h,w) int16
    (0 = void, c+1 = class c).  -> keys (B,P) int64: (float_bits(min distance) << 32) | flat index."""
    lib = _lib.load()
    z, cfg = prepare_z(z, cfg)
    p2 = protos.reshape(protos.shape[0], -1).contiguous()
    B, Cc, H, W = z.shape
    P, D = p2.shape
    if labels.dtype != torch.int16 or tuple(labels.shape) != (B, H, W) or not labels.is_cuda:
        raise RuntimeError("labels must be a CUDA int16 tensor of shape (B,h,w)")
    labels = labels.contiguous()
    proto_class = proto_class.to(device=z.device, dtype=torch.int32).contiguous()
    keys = keys_out if keys_out is not None else torch.empty((B, P), device=z.device, dtype=torch.int64)
    desc = make_desc(cfg, B, H, W, D, P, z)
    nbytes = int(lib.sps_head_fwd_workspace_bytes(C.byref(desc)))
    ws = torch.empty(nbytes, device=z.device, dtype=torch.uint8) if nbytes else None
    with torch.cuda.device(z.device):
        _lib.check(lib.sps_push_scan(C.byref(desc), _ptr(z), _ptr(p2), _ptr(labels), _ptr(proto_class), _ptr(keys),
                                     _ptr(ws), nbytes, _stream()), "sps_push_scan")
    return keys


def reduce_keys(keys: torch.Tensor, map_base: int, best: Optional[torch.Tensor] = None) -> torch.Tensor:
    """keys (n_maps,P) -> best (P,) int64 merged by min: (float_bits(d) << 32) | global map index."""
    lib = _lib.load()
    n_maps, P = keys.shape
    if best is None:
        best = torch.full((P,), KEY_INIT, device=keys.device, dtype=torch.int64)
    with torch.cuda.device(keys.device):
        _lib.check(lib.sps_push_reduce(_ptr(keys), n_maps, P, int(map_base), _ptr(best), _stream()),
                   "sps_push_reduce")
    return best


def gather_winners(z: torch.Tensor, keys: torch.Tensor, best: torch.Tensor, map_base: int, proto_scale: torch.Tensor,
                   cfg: HeadConfig, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Writes out[p] = z[win_map(p), scale(p)*D:(scale(p)+1)*D, i, j] for the winners held in z."""
    lib = _lib.load()
    z, cfg = prepare_z(z, cfg)
    B, Cc, H, W = z.shape
    P = keys.shape[1]
    S = len(cfg.scale_lo) - 1
    D = Cc // S
    if out is None:
        out = torch.zeros((P, D), device=z.device, dtype=torch.float32)
    proto_scale = proto_scale.to(device=z.device, dtype=torch.int32).contiguous()
    desc = make_desc(cfg, B, H, W, D, P, z)
    with torch.cuda.device(z.device):
        _lib.check(lib.sps_push_gather(C.byref(desc), _ptr(z), _ptr(keys), _ptr(best), B, int(map_base),
                                       _ptr(proto_scale), _ptr(out), _stream()), "sps_push_gather")
    return out


def decode_keys(keys: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """int64 keys -> (low word as int64, high word reinterpreted as fp32)."""
    low = keys & 0xFFFFFFFF
    val = (keys >> 32).to(torch.int32).view(torch.float32)
    return low, val


def push_project(z: torch.Tensor, labels: torch.Tensor, protos: torch.Tensor, identity: torch.Tensor,
                 cfg: HeadConfig, map_base: int = 0, group=None, batch: int = 32):
    """Whole projection over the maps THIS rank holds (all of the same size): scan in batches,
    dataset arg-min, (optional) cross-rank exchange, gather.

    Returns (win_map (P,) int64 global map index, win_pix (P,) int64, win_dist (P,) fp32,
             new_prototypes (P,D) fp32, keys (n_local,P) int64)."""
    n, P = z.shape[0], protos.shape[0]
    S = len(cfg.scale_lo) - 1
    pc = proto_class_of(identity).to(z.device)
    ps = proto_scale_reference(P, S).to(z.device)
    keys = torch.empty((n, P), device=z.device, dtype=torch.int64)
    for i in range(0, n, batch):
        scan_maps(z[i:i + batch], labels[i:i + batch], protos, pc, cfg, keys_out=keys[i:i + batch])
    best = reduce_keys(keys, map_base)
    if group is not None:
        import torch.distributed as dist
        dist.all_reduce(best, op=dist.ReduceOp.MIN, group=group)
    out = gather_winners(z, keys, best, map_base, ps, cfg)
    win_map, win_dist = decode_keys(best)
    local = win_map - map_base
    mine = (local >= 0) & (local < n)
    pix = torch.zeros_like(win_map)
    if n > 0:
        idx = torch.clamp(local, 0, max(n - 1, 0))
        pix = torch.where(mine, keys[idx, torch.arange(P, device=z.device)] & 0xFFFFFFFF, pix)
    if group is not None:
        import torch.distributed as dist
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(pix, op=dist.ReduceOp.SUM, group=group)
    return win_map, pix, win_dist, out, keys


def nonvoid_labels(labels: torch.Tensor) -> torch.Tensor:
    """Labels for a scan that masks only the void class, as find_nearest.py:138 does (distances + 10e6 at void pixels):
    1 at every labelled pixel, 0 at void; use with proto_class = zeros."""
    return (labels != 0).to(torch.int16)


def topk_keys(keys: torch.Tensor, k: int, map_base: int = 0, top: Optional[Tuple[torch.Tensor, torch.Tensor]] = None):
    """keys (n_maps,P) -> (top_key (P,k) int64 ascending, top_pix (P,k) int32), merged into `top` when given."""
    lib = _lib.load()
    n_maps, P = keys.shape
    if top is None:
        top = (torch.full((P, k), -1, device=keys.device, dtype=torch.int64),     # 0xFF..FF as uint64
               torch.zeros((P, k), device=keys.device, dtype=torch.int32))
    with torch.cuda.device(keys.device):
        _lib.check(lib.sps_push_topk(_ptr(keys.contiguous()), n_maps, P, int(map_base), int(k), _ptr(top[0]), _ptr(top[1]),
                                     _stream()), "sps_push_topk")
    return top


def topk_nearest(z: torch.Tensor, labels: torch.Tensor, protos: torch.Tensor, cfg: HeadConfig, k: int,
                 proto_class: Optional[torch.Tensor] = None, map_base: int = 0, group=None, batch: int = 32):
    """The k maps whose closest patch is nearest to each prototype (find_nearest.py:71-175 on resident maps).

    labels (n,h,w) int16: 0 = void, c+1 = class c.  proto_class None = the reference's void-only mask; a (P,) class
    vector = class-pure lists.  Returns (dist (P,k) fp32 ascending, map (P,k) int64 global index, pix (P,k) int64 flat
    position); entries with dist >= 1e10 mean fewer than k maps had an admissible pixel.  With `group`, every rank
    ends up with the merged lists (all-gather of the per-rank lists, merged by the same kernel)."""
    n, P = z.shape[0], protos.shape[0]
    if proto_class is None:
        labels, proto_class = nonvoid_labels(labels), torch.zeros(P, dtype=torch.int32)
    pc = proto_class.to(z.device)
    keys = torch.empty((n, P), device=z.device, dtype=torch.int64)
    for i in range(0, n, batch):
        scan_maps(z[i:i + batch], labels[i:i + batch], protos, pc, cfg, keys_out=keys[i:i + batch])
    top_key, top_pix = topk_keys(keys, k, map_base)
    if group is not None:
        import torch.distributed as dist
        world = dist.get_world_size(group)
        all_key = [torch.empty_like(top_key) for _ in range(world)]
        all_pix = [torch.empty_like(top_pix) for _ in range(world)]
        dist.all_gather(all_key, top_key, group=group)
        dist.all_gather(all_pix, top_pix, group=group)
        top_key, top_pix = merge_topk(all_key, all_pix)
    dist_bits = (top_key >> 32).to(torch.int32).view(torch.float32)
    return dist_bits, top_key & 0xFFFFFFFF, top_pix.to(torch.int64)


def merge_topk(all_key, all_pix):
    """Merge per-rank (P,k) lists: ascending packed keys are totally ordered, so a sort of the concatenation and the
    first k columns is the merged list (a few KB; plumbing, not a kernel)."""
    k = all_key[0].shape[1]
    cat_key, cat_pix = torch.cat(all_key, dim=1), torch.cat(all_pix, dim=1)
    order = torch.argsort(cat_key.view(torch.int64) ^ (-0x8000000000000000), dim=1, stable=True)[:, :k]   # unsigned order
    return torch.gather(cat_key, 1, order), torch.gather(cat_pix, 1, order)

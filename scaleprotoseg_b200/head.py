"""Host side of the prototype head: tensor checks, descriptors and the two autograd Functions that
call libscaleproto_sm100.so.  PyTorch is plumbing here (device memory, streams, autograd graph);
the arithmetic is in csrc/.

Mirrors, argument for argument, what the reference computes in
segmentation/model/model_multiscale.py:283-388 (stage 1) and
segmentation/model/model_multiscale_group.py:283-308 (stage 2).
"""
import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import (ACT_LINEAR, ACT_LOG, FLAG_ZPACK, FLAG_ZRELAYED, MATH_BF16, MATH_BF16X3, MATH_F16, MATH_F16X3, MATH_FP32,
                   SpsGroupDesc, SpsHeadDesc, Z_BF16, Z_F16, Z_F32, Z_NCHW, Z_NHWC)

# what install() uses: tcgen05 kernels with 3-term fp16 products (fp32-class distances, |z|,|p| < 6e4 — the
# reference's add-on layers end in a Sigmoid, segmentation/model/model_multiscale.py:206-208)
_DEFAULT_MATH = MATH_F16X3


def set_default_math(math: int):
    global _DEFAULT_MATH
    _DEFAULT_MATH = int(math)


def default_math() -> int:
    return _DEFAULT_MATH


@dataclass(frozen=True)
class HeadConfig:
    scale_lo: Tuple[int, ...]   # S+1 entries, scale s owns prototypes [scale_lo[s], scale_lo[s+1])
    num_classes: int
    act_kind: int = ACT_LOG
    eps: float = 1e-4
    math: int = MATH_FP32
    flags: int = 0

    @staticmethod
    def from_ranges(ranges: Sequence[Tuple[int, int]], num_classes: int, act: str = "log", eps: float = 1e-4,
                    math: Optional[int] = None) -> "HeadConfig":
        lo = [int(ranges[0][0])]
        for a, b in ranges:
            if int(a) != lo[-1]:
                raise ValueError(f"scale ranges must be contiguous, got {list(ranges)}")
            lo.append(int(b))
        kind = {"log": ACT_LOG, "linear": ACT_LINEAR}[act]
        return HeadConfig(tuple(lo), int(num_classes), kind, float(eps), _DEFAULT_MATH if math is None else int(math))


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _check_cuda_f32(name: str, t: torch.Tensor):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: the prototype head has no CPU path")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")


_SPLIT = (MATH_F16X3, MATH_BF16X3)


def prepare_z(z: torch.Tensor, cfg: HeadConfig):
    """Feature maps as the kernels take them: fp32 NCHW (the reference's own layout), or fp16 / bf16 channels-last
    (what the stock backbone emits under channels_last + autocast; fed to the tensor cores by TMA).
    -> (z in a layout the library accepts, HeadConfig whose math mode matches the element type)."""
    if not z.is_cuda:
        raise RuntimeError("z must be a CUDA tensor: the prototype head has no CPU path")
    if z.dim() != 4:
        raise RuntimeError(f"z must be (B,C,H,W), got {tuple(z.shape)}")
    if z.dtype == torch.float32:
        return z.contiguous(), cfg
    if z.dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError(f"z must be float32, float16 or bfloat16, got {z.dtype}")
    if cfg.math == MATH_FP32:   # exact-operand mode asked for: widen (a 16-bit value is exact in fp32)
        return z.float().contiguous(), cfg
    split = cfg.math in _SPLIT
    math = (MATH_F16X3 if split else MATH_F16) if z.dtype == torch.float16 else (MATH_BF16X3 if split else MATH_BF16)
    if math != cfg.math:
        cfg = HeadConfig(cfg.scale_lo, cfg.num_classes, cfg.act_kind, cfg.eps, math, cfg.flags)
    return z.contiguous(memory_format=torch.channels_last), cfg


def make_desc(cfg: HeadConfig, B: int, H: int, W: int, D: int, P: int, z: Optional[torch.Tensor] = None,
              extra_flags: int = 0) -> SpsHeadDesc:
    S = len(cfg.scale_lo) - 1
    d = SpsHeadDesc()
    d.B, d.H, d.W, d.S, d.D, d.P, d.K = B, H, W, S, D, P, cfg.num_classes
    for i in range(_lib.SPS_MAX_SCALES + 1):
        d.scale_lo[i] = cfg.scale_lo[i] if i <= S else P
    d.act_kind, d.eps, d.math, d.flags = cfg.act_kind, cfg.eps, cfg.math, cfg.flags | extra_flags
    d.z_dtype, d.z_layout = Z_F32, Z_NCHW
    if z is not None and z.dtype != torch.float32:
        d.z_dtype, d.z_layout = (Z_F16 if z.dtype == torch.float16 else Z_BF16), Z_NHWC
    return d


def _shape_of(z: torch.Tensor, protos: torch.Tensor, cfg: HeadConfig):
    if z.dim() != 4:
        raise RuntimeError(f"z must be (B,C,H,W), got {tuple(z.shape)}")
    B, Cc, H, W = z.shape
    S = len(cfg.scale_lo) - 1
    P, D = protos.shape[0], protos.shape[1]
    if protos.numel() != P * D:
        raise RuntimeError("only 1x1 prototypes are supported")
    if Cc != S * D:
        raise RuntimeError(f"z has {Cc} channels, expected num_scales*D = {S}*{D}")
    if cfg.scale_lo[-1] != P:
        raise RuntimeError(f"scale ranges end at {cfg.scale_lo[-1]} but there are {P} prototypes")
    return B, H, W, S, D, P


def wants_relay(z: torch.Tensor, cfg: HeadConfig) -> bool:
    """fp32 NCHW map + FLAG_ZPACK + tensor-core math + D = 64: the map is re-laid ONCE per step (sps_z_relay) as 16-bit
    channels-last hi / lo arrays and both the forward and the backward run their TMA-fed kernels on it."""
    S = len(cfg.scale_lo) - 1
    # asked for, or a wide last layer (more than 32 classes: forward and backward both run TMA-fed kernels)
    asked = bool(cfg.flags & FLAG_ZPACK) or cfg.num_classes > 32
    return (z.dtype == torch.float32 and asked and cfg.math != MATH_FP32 and z.shape[1] == 64 * S and z.numel() > 0)


def z_relay(z: torch.Tensor, protos: torch.Tensor, cfg: HeadConfig) -> torch.Tensor:
    lib = _lib.load()
    z = z.contiguous()
    B, H, W, S, D, P = _shape_of(z, protos, cfg)
    desc = make_desc(cfg, B, H, W, D, P)
    nbytes = int(lib.sps_z_relay_bytes(C.byref(desc)))
    out = torch.empty(max(nbytes, 16), device=z.device, dtype=torch.uint8)
    with torch.cuda.device(z.device):
        _lib.check(lib.sps_z_relay(C.byref(desc), _ptr(z), _ptr(out), nbytes, _stream()), "sps_z_relay")
    return out


def head_forward_raw(z, protos, last_w, cfg: HeadConfig, want_dist: bool, want_act: bool, relay=None):
    """No autograd.  -> (logits (N,K) | None, dist (B,P,H,W) | None, act (N,P) | None).
    relay: a buffer written by z_relay for this z (then z itself is only used for its shape)."""
    lib = _lib.load()
    _check_cuda_f32("prototype_vectors", protos)
    z, cfg = prepare_z(z, cfg)
    protos = protos.contiguous()
    B, H, W, S, D, P = _shape_of(z, protos, cfg)
    N = B * H * W
    logits = dist = act = None
    if last_w is not None:
        _check_cuda_f32("last_layer.weight", last_w)
        last_w = last_w.contiguous()
        if tuple(last_w.shape) != (cfg.num_classes, P):
            raise RuntimeError(f"last layer weight {tuple(last_w.shape)} != (K={cfg.num_classes}, P={P})")
        logits = torch.empty((N, cfg.num_classes), device=z.device, dtype=torch.float32)
    if want_dist:
        dist = torch.empty((B, P, H, W), device=z.device, dtype=torch.float32)
    if want_act:
        act = torch.empty((N, P), device=z.device, dtype=torch.float32)
    desc = make_desc(cfg, B, H, W, D, P, z, FLAG_ZRELAYED if relay is not None else 0)
    nbytes = int(lib.sps_head_fwd_workspace_bytes(C.byref(desc)))
    ws = torch.empty(nbytes, device=z.device, dtype=torch.uint8) if nbytes else None
    with torch.cuda.device(z.device):
        _lib.check(lib.sps_head_fwd(C.byref(desc), _ptr(relay if relay is not None else z), _ptr(protos), _ptr(last_w),
                                    _ptr(logits), _ptr(dist), _ptr(act), _ptr(ws), nbytes, _stream()), "sps_head_fwd")
    return logits, dist, act


def head_backward_raw(z, protos, last_w, cfg: HeadConfig, dlogits, ddist, dact, need_dz, need_dp, need_dw, relay=None):
    lib = _lib.load()
    z_in = z
    z, cfg = prepare_z(z, cfg)
    protos = protos.contiguous()
    B, H, W, S, D, P = _shape_of(z, protos, cfg)
    desc = make_desc(cfg, B, H, W, D, P, z, FLAG_ZRELAYED if relay is not None else 0)
    if last_w is not None:
        last_w = last_w.contiguous()
    # dz mirrors z (element type and layout); preserve_format keeps channels-last for the 16-bit maps
    dz = torch.empty_like(z, memory_format=torch.preserve_format) if need_dz else None
    dp = torch.empty_like(protos, memory_format=torch.contiguous_format) if need_dp else None
    dw = torch.empty_like(last_w, memory_format=torch.contiguous_format) if (need_dw and last_w is not None) else None
    if dlogits is not None:
        dlogits = dlogits.contiguous().view(-1, cfg.num_classes)
        if dlogits.dtype != torch.float32:
            dlogits = dlogits.float()
    if ddist is not None:
        ddist = ddist.contiguous()
    if dact is not None:
        dact = dact.contiguous()
    nbytes = int(lib.sps_workspace_bytes(C.byref(desc)))
    ws = torch.empty(max(nbytes, 16), device=z.device, dtype=torch.uint8)
    with torch.cuda.device(z.device):
        _lib.check(lib.sps_head_bwd(C.byref(desc), _ptr(relay if relay is not None else z), _ptr(protos), _ptr(last_w),
                                    _ptr(dlogits), _ptr(ddist), _ptr(dact), _ptr(dz), _ptr(dp), _ptr(dw), _ptr(ws), nbytes,
                                    _stream()), "sps_head_bwd")
    if dz is not None and dz.dtype != z_in.dtype:     # a 16-bit map widened for the exact fp32 mode
        dz = dz.to(z_in.dtype)
    return dz, dp, dw


class PrototypeHeadFn(torch.autograd.Function):
    """(z, prototype_vectors, last_layer.weight | None) -> (logits | None, distances | None, activations | None)."""

    @staticmethod
    def forward(ctx, z, protos, last_w, cfg: HeadConfig, want_dist: bool, want_act: bool):
        p2 = protos.reshape(protos.shape[0], -1)
        relay = z_relay(z, p2, cfg) if wants_relay(z, cfg) else None
        logits, dist, act = head_forward_raw(z, p2, last_w, cfg, want_dist, want_act, relay=relay)
        ctx.cfg = cfg
        ctx.proto_shape = protos.shape
        ctx.has_w = last_w is not None
        ctx.relay = relay          # one re-lay per step: the backward's TMA-fed kernel reads the same buffer
        ctx.save_for_backward(z, p2, last_w if last_w is not None else torch.empty(0, device=z.device))
        ctx.set_materialize_grads(False)
        return logits, dist, act

    @staticmethod
    def backward(ctx, dlogits, ddist, dact):
        z, p2, last_w = ctx.saved_tensors
        if not ctx.has_w:
            last_w = None
        need_dz, need_dp, need_dw = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        if dlogits is None and ddist is None and dact is None:
            return None, None, None, None, None, None
        relay = ctx.relay if (dact is None and dlogits is not None and ctx.has_w) else None
        dz, dp, dw = head_backward_raw(z, p2, last_w, ctx.cfg, dlogits, ddist, dact, need_dz, need_dp,
                                       need_dw and ctx.has_w, relay=relay)
        ctx.relay = None
        if dp is not None:
            dp = dp.view(ctx.proto_shape)
        return dz, dp, dw, None, None, None


# ------------------------------------------------------------------------------------------------
# Stage 2: group projections
# ------------------------------------------------------------------------------------------------
@dataclass
class GroupTables:
    """Index tables derived once from prototype_class_identity (P x K one-hot, CPU tensor in the
    reference: model_multiscale_group.py:294-298 recomputes them per class per call with a host sync)."""
    Kc: int
    J: int
    class_ids: List[int]            # class id of slot k'
    sizes: List[int]                # n_k'
    proto_slot: torch.Tensor        # int32 [P] device
    proto_col: torch.Tensor         # int32 [P]
    class_off: torch.Tensor         # int32 [Kc+1]
    members: torch.Tensor           # int32 [P]
    members_host: List[torch.Tensor]


def build_group_tables(identity: torch.Tensor, num_groups: int, device) -> GroupTables:
    ident = identity.detach().to("cpu")
    P, K = ident.shape
    class_ids, sizes, members_host = [], [], []
    slot = torch.full((P,), -1, dtype=torch.int32)
    col = torch.zeros((P,), dtype=torch.int32)
    off = [0]
    for c in range(K):
        idx = torch.nonzero(ident[:, c]).flatten()
        if idx.numel() == 0:
            continue
        k = len(class_ids)
        class_ids.append(c)
        sizes.append(int(idx.numel()))
        members_host.append(idx)
        slot[idx] = k
        col[idx] = torch.arange(idx.numel(), dtype=torch.int32)
        off.append(off[-1] + int(idx.numel()))
    members = torch.cat(members_host).to(torch.int32) if members_host else torch.zeros(0, dtype=torch.int32)
    if members.numel() != P:
        raise RuntimeError("prototype_class_identity must assign every prototype to exactly one class")
    return GroupTables(len(class_ids), int(num_groups), class_ids, sizes, slot.to(device), col.to(device),
                       torch.tensor(off, dtype=torch.int32, device=device), members.to(device), members_host)


def _group_desc(t: GroupTables, G: torch.Tensor, Wg: torch.Tensor) -> SpsGroupDesc:
    g = SpsGroupDesc()
    g.Kc, g.J = t.Kc, t.J
    g.proto_slot, g.proto_col = t.proto_slot.data_ptr(), t.proto_col.data_ptr()
    g.class_off, g.members = t.class_off.data_ptr(), t.members.data_ptr()
    g.G, g.Wg = G.data_ptr(), Wg.data_ptr()
    return g


class GroupHeadFn(torch.autograd.Function):
    """(act (N,P), G packed (J*P,), Wg (K, J*Kc)) -> (groups (N, J*Kc), logits (N,K))."""

    @staticmethod
    def forward(ctx, act, G, Wg, cfg: HeadConfig, tables: GroupTables, hw: Tuple[int, int, int]):
        lib = _lib.load()
        _check_cuda_f32("prototype_activations", act), _check_cuda_f32("group weights", G)
        act, G, Wg = act.contiguous(), G.contiguous(), Wg.contiguous()
        N, P = act.shape
        Gt = tables.J * tables.Kc
        desc = make_desc(cfg, hw[0], hw[1], hw[2], 1, P)
        desc.S, desc.D = 1, 1
        desc.scale_lo[0], desc.scale_lo[1] = 0, P
        groups = torch.empty((N, Gt), device=act.device, dtype=torch.float32)
        logits = torch.empty((N, cfg.num_classes), device=act.device, dtype=torch.float32)
        gd = _group_desc(tables, G, Wg)
        with torch.cuda.device(act.device):
            _lib.check(lib.sps_group_fwd(C.byref(desc), C.byref(gd), _ptr(act), _ptr(groups), _ptr(logits), _stream()),
                       "sps_group_fwd")
        ctx.cfg, ctx.tables, ctx.hw = cfg, tables, hw
        ctx.save_for_backward(act, G, Wg, groups)     # groups: 1/13 of the activation map; saves its re-read in the backward
        ctx.set_materialize_grads(False)
        return groups, logits

    @staticmethod
    def backward(ctx, dgroups, dlogits):
        lib = _lib.load()
        act, G, Wg, groups = ctx.saved_tensors
        if dgroups is None and dlogits is None:
            return None, None, None, None, None, None
        cfg, tables, hw = ctx.cfg, ctx.tables, ctx.hw
        N, P = act.shape
        desc = make_desc(cfg, hw[0], hw[1], hw[2], 1, P)
        desc.S, desc.D = 1, 1
        desc.scale_lo[0], desc.scale_lo[1] = 0, P
        gd = _group_desc(tables, G, Wg)
        need_da, need_dG, need_dWg = ctx.needs_input_grad[:3]
        dact = torch.empty_like(act, memory_format=torch.contiguous_format) if need_da else None
        dG = torch.empty_like(G, memory_format=torch.contiguous_format) if need_dG else None
        dWg = torch.empty_like(Wg, memory_format=torch.contiguous_format) if need_dWg else None
        if dgroups is not None:
            dgroups = dgroups.contiguous()
        if dlogits is not None:
            dlogits = dlogits.contiguous()
        nbytes = int(lib.sps_group_workspace_bytes(C.byref(desc), C.byref(gd)))
        ws = torch.empty(max(nbytes, 16), device=act.device, dtype=torch.uint8)
        with torch.cuda.device(act.device):
            _lib.check(lib.sps_group_bwd(C.byref(desc), C.byref(gd), _ptr(act), _ptr(groups), _ptr(dlogits), _ptr(dgroups),
                                         _ptr(dact), _ptr(dG), _ptr(dWg), _ptr(ws), nbytes, _stream()),
                       "sps_group_bwd")
        if dWg is not None and dlogits is None:
            dWg.zero_()
        return dact, dG, dWg, None, None, None

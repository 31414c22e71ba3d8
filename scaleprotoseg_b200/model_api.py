"""Drop-in replacements for the head methods of the reference's two `PPNetMultiScale` classes.

Every function here takes the reference module as `self` and keeps the reference signature and return
convention (segmentation/model/model_multiscale.py:243-398 and
segmentation/model/model_multiscale_group.py:283-460).  `install.install()` binds them onto the
reference classes; nothing is added to the module's state, so the whole-module pickles the reference
writes (module_multiscale.py:348-353) stay loadable with or without this package.
"""
import weakref
from typing import Any, List, Union

import torch

from .head import (GroupHeadFn, GroupTables, HeadConfig, PrototypeHeadFn, build_group_tables, default_math)

_group_cache: "weakref.WeakKeyDictionary" = weakref.WeakKeyDictionary()


def head_config(m, math=None) -> HeadConfig:
    if hasattr(m, "scale_num_prototypes"):
        ranges = [tuple(m.scale_num_prototypes[s]) for s in range(m.num_scales)]
    else:   # single-scale ProtoSeg baseline (segmentation/model/model.py:74): one scale owning every prototype
        ranges = [(0, int(m.prototype_vectors.shape[0]))]
    return HeadConfig.from_ranges(ranges, m.num_classes, m.prototype_activation_function, m.epsilon,
                                  default_math() if math is None else math)


MAX_CLASSES, MAX_DEPTH, MAX_CHUNKS, MAX_GROUPS_PER_CLASS = 192, 64, 160, 8     # limits of csrc/ (api.cu, tc_common.cuh, group.cu)


def is_supported(m) -> bool:
    """Configurations the kernels cover; anything else keeps running the reference's own code (never an error in the middle
    of training): activation kind, 1x1 prototypes, no scale head, and the kernels' shape limits - at most 192 classes, a
    prototype depth that is a multiple of 8 up to 64, at most 160 chunks of 64 prototypes over the scales, and at most 8
    groups per class for the stage-2 head."""
    act = m.prototype_activation_function
    shape = tuple(m.prototype_vectors.shape)
    if not (getattr(m, "scale_head", None) is None and isinstance(act, str) and act in ("log", "linear")
            and len(shape) == 4 and shape[2] == 1 and shape[3] == 1):
        return False
    if shape[1] > MAX_DEPTH or shape[1] % 8 != 0 or int(m.num_classes) > MAX_CLASSES or int(m.num_classes) < 1:
        return False
    if hasattr(m, "scale_num_prototypes"):
        ranges = [tuple(m.scale_num_prototypes[s]) for s in range(m.num_scales)]
    else:
        ranges = [(0, shape[0])]
    if sum((int(b) - int(a) + 63) // 64 for a, b in ranges) > MAX_CHUNKS:
        return False
    if hasattr(m, "num_groups") and not (1 <= int(m.num_groups) <= MAX_GROUPS_PER_CLASS):
        return False
    return True


_DEFER_DISTANCES = False


def set_deferred_distances(on: bool):
    """Opt in: while training, the stage-1 `forward_from_conv_features(x)` (the call of module_multiscale.py:209, whose
    distances go nowhere but into KLDLoss, :242) returns a DeferredDistances handle instead of the (B,P,H,W) map, the
    head runs in fused mode, and the patched KLDLoss.forward computes the loss straight from the features."""
    global _DEFER_DISTANCES
    _DEFER_DISTANCES = bool(on)


class DeferredDistances:
    """Stands in for the dense distance map of one head call.  Carries what is needed to compute any part of it
    (features, prototypes, head configuration).  The patched KLDLoss.forward consumes it without materialising it.
    Named attributes and methods of a tensor (`.shape`, `.mean()`, `.permute(...)`, ...), indexing, `len()` and
    `torch.*(handle)` calls materialise the map through the head kernels and delegate, with autograd intact.  It is NOT a
    Tensor subclass: `isinstance(handle, torch.Tensor)` is False and arithmetic operators are not forwarded - call
    `.materialize()` first where a real tensor is required.  Opt-in (set_deferred_distances), for the reference trainer
    whose only use of the map is KLDLoss."""

    def __init__(self, z: torch.Tensor, protos: torch.Tensor, cfg: HeadConfig):
        self.z, self.protos, self.cfg = z, protos, cfg
        self._dense = None

    @property
    def shape(self):
        B, _, H, W = self.z.shape
        return torch.Size((B, self.protos.shape[0], H, W))

    @property
    def is_cuda(self):
        return self.z.is_cuda

    @property
    def device(self):
        return self.z.device

    def materialize(self) -> torch.Tensor:
        if self._dense is None:
            _, self._dense, _ = PrototypeHeadFn.apply(self.z, self.protos, None, self.cfg, True, False)
        return self._dense

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return getattr(self.materialize(), name)

    def __getitem__(self, idx):
        return self.materialize()[idx]

    def __len__(self):
        return int(self.z.shape[0])

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        unwrap = lambda a: a.materialize() if isinstance(a, DeferredDistances) else a      # noqa: E731
        return func(*[unwrap(a) for a in args], **{k: unwrap(v) for k, v in (kwargs or {}).items()})


def _require_cuda(x: torch.Tensor):
    if not x.is_cuda:
        raise RuntimeError("scaleprotoseg_b200 is installed: the prototype head runs on CUDA only "
                           "(got a CPU tensor; there is no CPU fallback)")


def group_tables(m, device) -> GroupTables:
    ident = m.prototype_class_identity
    ent = _group_cache.get(m)
    if ent is not None:
        tables, snap = ent
        if (snap.shape == ident.shape and tables.proto_slot.device == torch.device(device)
                and tables.J == m.num_groups and torch.equal(snap, ident.detach().cpu())):
            return tables
    tables = build_group_tables(ident, m.num_groups, device)
    _group_cache[m] = (tables, ident.detach().cpu().clone())
    return tables


# ---- methods shared by both classes -------------------------------------------------------------
def _scale_l2_convolution(self, x: torch.Tensor) -> torch.Tensor:
    _require_cuda(x)
    _, dist, _ = PrototypeHeadFn.apply(x.float(), self.prototype_vectors, None, head_config(self), True, False)
    return dist


def prototype_distances(self, x: torch.Tensor) -> torch.Tensor:
    return self._scale_l2_convolution(self.conv_features(x))


def push_forward(self, x: torch.Tensor) -> Any:
    conv_output = self.conv_features(x)
    if isinstance(conv_output, list):
        return [(c, self._scale_l2_convolution(c)) for c in conv_output]
    return conv_output, self._scale_l2_convolution(conv_output)


def _returns(logits, distances, activations, return_activations, return_distances):
    if return_activations and not return_distances:
        return logits, activations
    if return_activations and return_distances:
        return logits, distances, activations
    return logits, distances


# ---- stage 1 ------------------------------------------------------------------------------------
def forward_from_conv_features_stage1(self, conv_features: Union[torch.Tensor, List[torch.Tensor]],
                                      return_activations: bool = False, return_distances: bool = False) -> Any:
    if isinstance(conv_features, list):
        return [self.forward_from_conv_features(c) for c in conv_features]
    if not (hasattr(self, "patch_classification") and self.patch_classification):
        raise Exception("Original Prototype Network Implementation")
    _require_cuda(conv_features)
    B, _, H, W = conv_features.shape
    want_dist = return_distances or not return_activations
    cfg = head_config(self)
    if (_DEFER_DISTANCES and self.training and torch.is_grad_enabled() and not return_activations
            and not return_distances):
        z = conv_features.float()
        logits, _, _ = PrototypeHeadFn.apply(z, self.prototype_vectors, self.last_layer.weight, cfg, False, False)
        return logits.reshape(B, H, W, -1), DeferredDistances(z, self.prototype_vectors, cfg)
    logits, dist, act = PrototypeHeadFn.apply(conv_features.float(), self.prototype_vectors, self.last_layer.weight,
                                              cfg, want_dist, bool(return_activations))
    logits = logits.reshape(B, H, W, -1)
    return _returns(logits, dist, act, return_activations, return_distances)


# ---- stage 2 ------------------------------------------------------------------------------------
# module -> (key, weakref to the activation tensor, (groups, logits)).  The stage-2 trainer runs the group head TWICE per
# step on the same activations: once inside forward_from_conv_features for the logits and once through compute_group for
# the KLD / entropy losses (module_multiscale_group_train.py:224 -> :306, then :242).  The second call gets the first
# call's outputs (one autograd node, one backward) instead of recomputing them.
_apply_cache: "weakref.WeakKeyDictionary" = weakref.WeakKeyDictionary()


def _group_apply(self, prototype_activations: torch.Tensor, hw=None):
    _require_cuda(prototype_activations)
    tables = group_tables(self, prototype_activations.device)
    if len(self.group_projection) != tables.Kc:
        raise RuntimeError("group_projection does not match prototype_class_identity "
                           f"({len(self.group_projection)} modules, {tables.Kc} classes with prototypes)")
    key = (id(prototype_activations), prototype_activations._version, tuple(gp.weight._version for gp in self.group_projection),
           tuple(gp.weight.data_ptr() for gp in self.group_projection), self.last_layer_group.weight._version,
           self.last_layer_group.weight.data_ptr(), torch.is_grad_enabled())
    ent = _apply_cache.get(self)
    if ent is not None and ent[0] == key and ent[1]() is prototype_activations:
        return ent[2], tables
    G = torch.cat([gp.weight.reshape(-1) for gp in self.group_projection])
    N = prototype_activations.shape[0]
    hw = hw if hw is not None else (1, 1, N)
    res = GroupHeadFn.apply(prototype_activations.float(), G, self.last_layer_group.weight, head_config(self), tables, hw)
    _apply_cache[self] = (key, weakref.ref(prototype_activations), res)
    return res, tables


def compute_group(self, prototype_activations: torch.Tensor) -> List[torch.Tensor]:
    (groups, _), tables = _group_apply(self, prototype_activations)
    return list(torch.split(groups, tables.J, dim=-1))


def run_last_layer_group(self, prototype_activations: torch.Tensor) -> torch.Tensor:
    (_, logits), _ = _group_apply(self, prototype_activations)
    return logits


def forward_from_conv_features_stage2(self, conv_features: Union[torch.Tensor, List[torch.Tensor]],
                                      return_activations: bool = False, return_distances: bool = False) -> Any:
    if isinstance(conv_features, list):
        return [self.forward_from_conv_features(c) for c in conv_features]
    if not (hasattr(self, "patch_classification") and self.patch_classification):
        raise Exception("Original Prototype Network Implementation")
    _require_cuda(conv_features)
    B, _, H, W = conv_features.shape
    want_dist = return_distances or not return_activations
    _, dist, act = PrototypeHeadFn.apply(conv_features.float(), self.prototype_vectors, None, head_config(self),
                                         want_dist, True)
    (_, logits), _ = _group_apply(self, act, (B, H, W))
    logits = logits.reshape(B, H, W, -1)
    return _returns(logits, dist, act, return_activations, return_distances)


# ---- single-scale ProtoSeg baseline (segmentation/model/model.py, class PPNet) --------------------
def _l2_convolution_single(self, x: torch.Tensor) -> torch.Tensor:
    """model.py:250-268: the S = 1 case of the same kernels (C = D)."""
    _require_cuda(x)
    _, dist, _ = PrototypeHeadFn.apply(x.float(), self.prototype_vectors, None, head_config(self), True, False)
    return dist


def prototype_distances_single(self, x: torch.Tensor) -> torch.Tensor:
    return self._l2_convolution(self.conv_features(x))          # model.py:270-276


def push_forward_single(self, x: torch.Tensor) -> Any:
    conv_output = self.conv_features(x)                          # model.py:345-354
    if isinstance(conv_output, list):
        return [(c, self._l2_convolution(c)) for c in conv_output]
    return conv_output, self._l2_convolution(conv_output)


def forward_from_conv_features_single(self, conv_features: Union[torch.Tensor, List[torch.Tensor]],
                                      return_activations: bool = False, return_distances: bool = False) -> Any:
    """model.py:306-343.  The patch-classification branch runs fused; the image-level branch (global min pooling, the
    original ProtoPNet head) keeps the reference's code."""
    if isinstance(conv_features, list):
        return [self.forward_from_conv_features(c) for c in conv_features]
    if not (hasattr(self, "patch_classification") and self.patch_classification):
        original = type(self).forward_from_conv_features.__sps_original__
        return original(self, conv_features, return_activations=return_activations, return_distances=return_distances)
    _require_cuda(conv_features)
    B, _, H, W = conv_features.shape
    logits, dist, act = PrototypeHeadFn.apply(conv_features.float(), self.prototype_vectors, self.last_layer.weight,
                                              head_config(self), not return_activations, bool(return_activations))
    logits = logits.reshape(B, H, W, -1)
    if return_activations:
        return logits, act                                       # model.py:327-328
    return logits, dist

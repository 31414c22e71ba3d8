"""`install()` — bind the B200 head and push onto the reference's own classes and module globals.

    import scaleprotoseg_b200; scaleprotoseg_b200.install()
    # then run segmentation.train_wandb_multiscale / finetune_wandb_group / eval_valid_multiscale unchanged

What is patched (reference file:line of what each replaces):
  segmentation.model.model_multiscale.PPNetMultiScale
      _scale_l2_convolution :283, prototype_distances :319, forward_from_conv_features :340, push_forward :390
  segmentation.model.model_multiscale_group.PPNetMultiScale
      _scale_l2_convolution :347, compute_group :283, run_last_layer :305, forward_from_conv_features :404,
      push_forward :454
  segmentation.model.model.PPNet (single-scale ProtoSeg baseline = one scale owning every prototype)
      _l2_convolution :250, prototype_distances :270, forward_from_conv_features :306 (patch branch), push_forward :345
  segmentation.push_multiscale_optimization (module globals, looked up at call time :286,:291)
      compute_distances :35, min_across_dataset :94, global_min :141
  segmentation.push_optimization (the single-scale PPNet's push driver, :276,:281)
      compute_distances :35, min_across_dataset :89, global_min :136
  segmentation.model.loss.KLDLoss
      forward :66 (the image x class x scale x pair loop; CPU inputs keep the reference's loop)
  segmentation.model.loss.KLDLossGroup
      forward :478 (the same loop over the stage-2 group activations)
Configurations the kernels do not cover (scale_head set, callable activation, non-1x1 prototypes)
keep the reference's original method.  CPU tensors raise: there is no CPU fallback.
"""
import functools
import os
from typing import List, Optional, Tuple

import numpy as np
import torch

from . import _lib, model_api
from . import kld as _kld
from . import push as _push

_installed = {}


def _guard(new_fn, old_fn):
    @functools.wraps(old_fn)
    def method(self, *args, **kwargs):
        if not model_api.is_supported(self):
            return old_fn(self, *args, **kwargs)
        return new_fn(self, *args, **kwargs)

    method.__sps_original__ = old_fn
    return method


def _patch(cls, name, new_fn):
    old = getattr(cls, name)
    if hasattr(old, "__sps_original__"):
        return
    setattr(cls, name, _guard(new_fn, old))
    _installed[(cls, name)] = old


def install_on_classes(stage1_cls=None, stage2_cls=None, single_cls=None):
    """Patch the given classes (any class with the reference's attribute names works)."""
    _lib.load()  # fail now, loudly, if the CUDA library has not been built
    if single_cls is not None:
        _patch(single_cls, "_l2_convolution", model_api._l2_convolution_single)
        _patch(single_cls, "prototype_distances", model_api.prototype_distances_single)
        _patch(single_cls, "push_forward", model_api.push_forward_single)
        _patch(single_cls, "forward_from_conv_features", model_api.forward_from_conv_features_single)
    if stage1_cls is not None:
        _patch(stage1_cls, "_scale_l2_convolution", model_api._scale_l2_convolution)
        _patch(stage1_cls, "prototype_distances", model_api.prototype_distances)
        _patch(stage1_cls, "push_forward", model_api.push_forward)
        _patch(stage1_cls, "forward_from_conv_features", model_api.forward_from_conv_features_stage1)
    if stage2_cls is not None:
        _patch(stage2_cls, "_scale_l2_convolution", model_api._scale_l2_convolution)
        _patch(stage2_cls, "push_forward", model_api.push_forward)
        _patch(stage2_cls, "compute_group", model_api.compute_group)
        _patch(stage2_cls, "run_last_layer", model_api.run_last_layer_group)
        _patch(stage2_cls, "forward_from_conv_features", model_api.forward_from_conv_features_stage2)


def uninstall():
    for (cls, name), old in list(_installed.items()):
        setattr(cls, name, old)
    _installed.clear()
    _push_uninstall()


# ---- push drop-ins --------------------------------------------------------------------------------
_push_saved = {}


def _labels_for_scan(target_tensor: torch.Tensor, void_class: Optional[int]) -> torch.Tensor:
    """Reference semantics (push_multiscale_optimization.py:73-82): one-hot over K (+1 with a void
    class) and the void column dropped.  The kernel wants 0 = never matches, c+1 = class column c."""
    t = target_tensor.to(torch.int64)
    if void_class is None:
        lab = t + 1
    else:
        lab = torch.where(t == void_class, torch.zeros_like(t), torch.where(t < void_class, t + 1, t))
    return lab.to(torch.int16)


def _dist_info():
    """(rank, world, group-less) of the default process group; (0, 1) when torch.distributed is not initialised."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous slice of the dataset a rank scans: contiguous so that 'first image wins' (argmin(dim=0),
    push_multiscale_optimization.py:137) stays 'smallest global index wins' after the min all-reduce of packed keys."""
    per = (n + world - 1) // world
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


SCAN_BATCH = 32     # same-size feature maps scanned per sps_push_scan launch (1.1 GB of full-resolution Cityscapes maps; the scan
                    # kernel has a fixed start-up of a few microseconds per launch, 7 % of a 16-map launch)


def make_push_functions(pm, load_image=None):
    """Builds the three replacements against the reference module `pm`
    (segmentation.push_multiscale_optimization, or segmentation.push_optimization for the single-scale PPNet) so they use
    ITS helpers (to_normalized_tensor, resize_label, PIL loading) for everything that is not the scan itself.
    load_image(dataset, img_id) -> image: overrides the PIL loader (tests drive the functions with in-memory maps).

    When torch.distributed is initialised with more than one rank, `min_across_dataset` scans only this rank's contiguous
    slice of `dataset.img_ids` and the ranks meet in ONE min all-reduce of P packed (distance, image) keys plus one
    all-gather of the per-image index lists; `global_min` gathers the winners on the rank that owns each winning image and
    one sum all-reduce of the (P, D) block hands them to everybody.  push_prototypes_multiscale itself runs unchanged
    (every rank ends with identical prototypes)."""

    def _load(dataset, img_id):
        if load_image is not None:
            return load_image(dataset, img_id)
        from PIL import Image
        with open(dataset.get_img_path(img_id), "rb") as f:
            img = Image.open(f).convert("RGB")
        m = dataset.image_margin_size
        return img.crop((m, m, img.width - m, img.height - m))

    def _features_and_labels(ppnet, dataset, img, target, void_class, device):
        img_tensor = pm.to_normalized_tensor(img).unsqueeze(0).to(device)
        z = ppnet.conv_features(img_tensor)      # a CPU map is refused by the scan itself (no CPU path)
        if dataset.convert_targets is not None:
            target = dataset.convert_targets(target)
        target_tensor = pm.resize_label(target, (z.shape[3], z.shape[2])).unsqueeze(0)
        return z.float(), _labels_for_scan(target_tensor, void_class).to(z.device)

    @torch.no_grad()
    def compute_distances(ppnet, dataset, img, target, num_classes, max_dist=1e10, device="cpu", void_class=None):
        ppnet.to(device)
        ppnet.eval()
        z, labels = _features_and_labels(ppnet, dataset, img, target, void_class, device)
        cfg = model_api.head_config(ppnet)
        pc = _push.proto_class_of(ppnet.prototype_class_identity)
        keys = _push.scan_maps(z, labels, ppnet.prototype_vectors.detach(), pc, cfg)
        idx, val = _push.decode_keys(keys)
        return idx, val

    @torch.no_grad()
    def min_across_dataset(dataset, ppnet, num_classes, void_class=None, device="cpu"):
        import torch.distributed as dist
        rank, world = _dist_info()
        ids = list(dataset.img_ids)
        n = len(ids)
        lo, hi = shard_bounds(n, rank, world)
        ppnet.to(device)
        ppnet.eval()
        cfg = model_api.head_config(ppnet)
        protos = ppnet.prototype_vectors.detach()
        P = protos.shape[0]
        pc = _push.proto_class_of(ppnet.prototype_class_identity)
        dev = protos.device
        keys = torch.empty((hi - lo, P), device=dev, dtype=torch.int64)
        pend_z, pend_l, pend0 = [], [], lo

        def flush():
            nonlocal pend_z, pend_l, pend0
            if pend_z:
                _push.scan_maps(torch.cat(pend_z), torch.cat(pend_l), protos, pc, cfg,
                                keys_out=keys[pend0 - lo:pend0 - lo + len(pend_z)])
                pend0 += len(pend_z)
                pend_z, pend_l = [], []

        for i in range(lo, hi):
            img = _load(dataset, ids[i])
            gt_ann = np.load(os.path.join(dataset.annotations_dir, ids[i] + ".npy"))
            z, labels = _features_and_labels(ppnet, dataset, img, gt_ann, void_class, device)
            if pend_z and (z.shape != pend_z[0].shape or len(pend_z) >= SCAN_BATCH):
                flush()
            pend_z.append(z), pend_l.append(labels)
        flush()
        # dataset arg-min: (float_bits(d) << 32) | global image index, merged by min (first image wins a tie)
        best = _push.reduce_keys(keys, lo) if hi > lo else torch.full((P,), _push.KEY_INIT, device=dev, dtype=torch.int64)
        idx_local = (keys & 0xFFFFFFFF)
        if world > 1:
            dist.all_reduce(best, op=dist.ReduceOp.MIN)
            per = (n + world - 1) // world
            pad = torch.zeros((per, P), device=dev, dtype=torch.int64)
            pad[:hi - lo] = idx_local
            parts = [torch.empty_like(pad) for _ in range(world)]
            dist.all_gather(parts, pad)
            idx_all = torch.cat(parts)[:n]
        else:
            idx_all = idx_local
        list_idx = [idx_all[i:i + 1] for i in range(n)]
        return (best & 0xFFFFFFFF), list_idx

    @torch.no_grad()
    def global_min(proto_min_dist, list_min_patch, dataset, ppnet, device="cpu"):
        """One backbone forward per DISTINCT winning image (on the rank that scanned it) instead of one per prototype;
        the winners' vectors are copied by sps_push_gather."""
        import torch.distributed as dist
        rank, world = _dist_info()
        P = ppnet.num_prototypes
        S = int(getattr(ppnet, "num_scales", 1))
        ppnet.to(device)
        cfg = model_api.head_config(ppnet)
        dev = ppnet.prototype_vectors.device
        D = ppnet.prototype_vectors.shape[1]
        ps = _push.proto_scale_reference(P, S).to(dev)       # p // (P // S), as :161,:167
        best = proto_min_dist.to(device=dev, dtype=torch.int64).contiguous()
        out = torch.zeros((P, D), device=dev, dtype=torch.float32)
        lo, hi = shard_bounds(len(dataset.img_ids), rank, world)
        for img_index in sorted(set(int(v) for v in proto_min_dist.tolist())):
            if not (lo <= img_index < hi):
                continue
            img = _load(dataset, dataset.img_ids[img_index])
            cf = ppnet.conv_features(pm.to_normalized_tensor(img).unsqueeze(0).to(device)).float()
            krow = list_min_patch[img_index].to(device=dev, dtype=torch.int64).reshape(1, P).contiguous()
            _push.gather_winners(cf, krow, best, img_index, ps, cfg, out=out)
        if world > 1:
            dist.all_reduce(out, op=dist.ReduceOp.SUM)
        host = out.cpu().numpy()
        return [host[p].reshape(D, 1, 1).copy() for p in range(P)]

    return compute_distances, min_across_dataset, global_min


def _push_uninstall():
    for (mod, name), old in list(_push_saved.items()):
        setattr(mod, name, old)
    _push_saved.clear()


def install_push(pm, load_image=None):
    """Rebind compute_distances / min_across_dataset / global_min of a push driver module
    (segmentation.push_multiscale_optimization :35,:94,:141 or segmentation.push_optimization :35,:89,:136)."""
    fns = make_push_functions(pm, load_image)
    for name, fn in zip(("compute_distances", "min_across_dataset", "global_min"), fns):
        if (pm, name) not in _push_saved:
            _push_saved[(pm, name)] = getattr(pm, name)
        setattr(pm, name, fn)


def install_kld(kld_cls):
    """Bind the fused KLD loss over `kld_cls.forward` (the reference's KLDLoss or any module with its three
    attributes).  Distances on the CPU keep the original loop."""
    _lib.load()
    old = getattr(kld_cls, "forward")
    if hasattr(old, "__sps_original__"):
        return

    @functools.wraps(old)
    def forward(self, prototype_distances, target_labels):
        if not prototype_distances.is_cuda or not _kld.kld_supported(self, prototype_distances.device):
            if isinstance(prototype_distances, model_api.DeferredDistances):
                prototype_distances = prototype_distances.materialize()
            return old(self, prototype_distances, target_labels)
        return _kld.kld_forward(self, prototype_distances, target_labels)

    forward.__sps_original__ = old
    setattr(kld_cls, "forward", forward)
    _installed[(kld_cls, "forward")] = old


def install_kld_group(kld_cls):
    """Bind the fused loss over `KLDLossGroup.forward` (loss.py:478-545: the image x class x group-pair loop of stage-2
    training, module_multiscale_group_train.py:260).  CPU inputs keep the original loop."""
    _lib.load()
    old = getattr(kld_cls, "forward")
    if hasattr(old, "__sps_original__"):
        return

    @functools.wraps(old)
    def forward(self, list_group_activation, target_labels):
        if len(list_group_activation) == 0 or not list_group_activation[0].is_cuda or int(self.num_groups) > 32:
            return old(self, list_group_activation, target_labels)
        return _kld.kld_group_forward(self, list_group_activation, target_labels)

    forward.__sps_original__ = old
    setattr(kld_cls, "forward", forward)
    _installed[(kld_cls, "forward")] = old


def install():
    """Patch the reference (must be importable: `segmentation` on sys.path)."""
    from segmentation.model import model_multiscale as mm
    from segmentation.model import model_multiscale_group as mg

    from segmentation.model import model as m1

    install_on_classes(mm.PPNetMultiScale, mg.PPNetMultiScale, m1.PPNet)
    import segmentation.push_multiscale_optimization as pm

    install_push(pm)
    import segmentation.push_optimization as pm1     # single-scale ProtoSeg push driver (PPNet, num_scales = 1)

    install_push(pm1)
    from segmentation.model import loss as ml

    install_kld(ml.KLDLoss)
    install_kld_group(ml.KLDLossGroup)

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
  segmentation.model.loss.KLDLoss
      forward :66 (the image x class x scale x pair loop; CPU inputs keep the reference's loop)
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


def make_push_functions(pm):
    """Builds the three replacements against the reference module `pm`
    (segmentation.push_multiscale_optimization) so they use ITS helpers (to_normalized_tensor,
    resize_label, PIL loading) for everything that is not the scan itself."""
    from PIL import Image

    @torch.no_grad()
    def compute_distances(ppnet, dataset, img, target, num_classes, max_dist=1e10, device="cpu", void_class=None):
        ppnet.to(device)
        ppnet.eval()
        img_tensor = pm.to_normalized_tensor(img).unsqueeze(0).to(device)
        z = ppnet.conv_features(img_tensor)
        if not z.is_cuda:
            raise RuntimeError("scaleprotoseg_b200 push runs on CUDA only (device=%r)" % (device,))
        if dataset.convert_targets is not None:
            target = dataset.convert_targets(target)
        target_tensor = pm.resize_label(target, (z.shape[3], z.shape[2])).unsqueeze(0)
        labels = _labels_for_scan(target_tensor, void_class).to(z.device)
        cfg = model_api.head_config(ppnet)
        pc = _push.proto_class_of(ppnet.prototype_class_identity)
        keys = _push.scan_maps(z.float(), labels, ppnet.prototype_vectors.detach(), pc, cfg)
        idx, val = _push.decode_keys(keys)
        return idx, val

    def _load(dataset, img_id):
        with open(dataset.get_img_path(img_id), "rb") as f:
            img = Image.open(f).convert("RGB")
        m = dataset.image_margin_size
        return img.crop((m, m, img.width - m, img.height - m))

    def min_across_dataset(dataset, ppnet, num_classes, void_class=None, device="cpu"):
        list_idx, list_dist = [], []
        for img_id in dataset.img_ids:
            img = _load(dataset, img_id)
            gt_ann = np.load(os.path.join(dataset.annotations_dir, img_id + ".npy"))
            idx, dist = pm.compute_distances(ppnet, dataset, img, gt_ann, num_classes=num_classes,
                                             void_class=void_class, device=device)
            list_dist.append(dist)
            list_idx.append(idx)
        tot = torch.concat(list_dist, dim=0)
        # first map holding the minimum: same packed-key order as the kernel uses across shards
        bits = tot.view(torch.int32).to(torch.int64) << 32
        keyed = bits | torch.arange(tot.shape[0], device=tot.device, dtype=torch.int64).unsqueeze(1)
        return (keyed.min(dim=0).values & 0xFFFFFFFF), list_idx

    @torch.no_grad()
    def global_min(proto_min_dist, list_min_patch, dataset, ppnet, device="cpu"):
        """One backbone forward per DISTINCT winning image instead of one per prototype."""
        P, S = ppnet.num_prototypes, ppnet.num_scales
        n_per_scale = P // S
        ppnet.to(device)
        out: List[Optional[np.ndarray]] = [None] * P
        wins = [int(v) for v in proto_min_dist.tolist()]
        for img_index in sorted(set(wins)):
            img = _load(dataset, dataset.img_ids[img_index])
            cf = ppnet.conv_features(pm.to_normalized_tensor(img).unsqueeze(0).to(device))
            _, Cc, H, W = cf.shape
            cf = cf.view(S, Cc // S, H, W)
            for p in [q for q in range(P) if wins[q] == img_index]:
                flat = int(list_min_patch[img_index][:, p].item())
                i, j = flat // W, flat % W
                out[p] = cf[p // n_per_scale, :, i:i + 1, j:j + 1].detach().cpu().numpy()
        return out

    return compute_distances, min_across_dataset, global_min


def _push_uninstall():
    for (mod, name), old in list(_push_saved.items()):
        setattr(mod, name, old)
    _push_saved.clear()


def install_push(pm):
    fns = make_push_functions(pm)
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
        if not prototype_distances.is_cuda:
            return old(self, prototype_distances, target_labels)
        return _kld.kld_forward(self, prototype_distances, target_labels)

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
    from segmentation.model import loss as ml

    install_kld(ml.KLDLoss)

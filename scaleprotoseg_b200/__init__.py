"""scaleprotoseg_b200 — B200-native (sm_100a) prototype head and push for ScaleProtoSeg.

Only the hot path lives here: csrc/ (CUDA kernels + the C ABI of include/scaleproto_b200.h), the
ctypes binding, the autograd wrappers and the method patches that drop the kernels into the
reference's own classes.  Backbone, losses, trainers and data loading stay the reference's.
"""
from ._lib import (ACT_LINEAR, ACT_LOG, EXPORTED_SYMBOLS, FLAG_DETERMINISTIC, LIB_PATH, MATH_BF16, MATH_BF16X3, MATH_F16,
                   MATH_F16X3, MATH_FP32)
from .head import (GroupHeadFn, HeadConfig, PrototypeHeadFn, build_group_tables, default_math, head_backward_raw,
                   head_forward_raw, set_default_math)
from .install import install, install_kld, install_kld_group, install_on_classes, install_push, uninstall
from .kld import KldFeatureLossFn, KldLossFn, KldTables, kld_group_loss, kld_loss, kld_loss_from_features
from . import evalmap, kld, model_api, push
from .evalmap import upsample_argmax, upsample_argmin
from .model_api import DeferredDistances, set_deferred_distances

PUSH_MATH = MATH_F16X3    # math mode the push scan runs in (tcgen05, 3-term fp16 products)
HAS_TC_PUSH = True
HAS_TC_BACKWARD = True   # sps_head_bwd runs the tcgen05 backward for the fused-mode step in the 16-bit math modes

__all__ = ["upsample_argmax", "upsample_argmin", "evalmap", "DeferredDistances", "set_deferred_distances", "install", "install_kld", "install_kld_group", "kld_group_loss", "KldLossFn", "KldFeatureLossFn", "KldTables", "kld_loss", "kld_loss_from_features", "kld", "install_on_classes", "install_push", "uninstall", "HeadConfig", "PrototypeHeadFn",
           "GroupHeadFn", "head_forward_raw", "head_backward_raw", "build_group_tables", "set_default_math",
           "default_math", "push", "model_api", "MATH_FP32", "MATH_BF16", "MATH_BF16X3", "MATH_F16", "MATH_F16X3", "ACT_LOG", "ACT_LINEAR",
           "LIB_PATH", "EXPORTED_SYMBOLS", "FLAG_DETERMINISTIC"]

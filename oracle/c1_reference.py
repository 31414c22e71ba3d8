"""BASELINE.json configs[0] (C1): the UNMODIFIED reference's whole-model CPU forward on one synthetic 1x3x513x513 image,
random-init DeepLabv2-ResNet101 multi-scale backbone + ScaleProtoSeg Cityscapes head (P = 760, K = 19), split into
backbone (`conv_features`) and prototype head (`forward_from_conv_features`, segmentation/model/model_multiscale.py:332-388).

Runs only where /root/reference exists (the build container).  TEST / MEASUREMENT INFRASTRUCTURE ONLY.
    python oracle/c1_reference.py  ->  one JSON line (committed as profiles/r02_c1_reference_cpu.json)
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.gen_golden import import_reference  # noqa: E402

import torch  # noqa: E402


def main():
    mm, _, _ = import_reference()
    from segmentation.model import deeplab_features as df
    torch.manual_seed(20220227)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    feat = df.deeplabv2_resnet101_features_multiscale(pretrained=False, deeplab_n_features=64, scales=[])
    m = mm.PPNetMultiScale(features=feat, img_size=513, prototype_shape=(760, 64, 1, 1), proto_layer_rf_info=[], num_classes=19,
                           init_weights=True, prototype_activation_function="log", add_on_layers_type="deeplab_simple",
                           patch_classification=True, num_scales=4)
    m.eval()
    x = torch.randn(1, 3, 513, 513)
    tb, th = [], []
    with torch.no_grad():
        m.forward_from_conv_features(m.conv_features(x))
        for _ in range(5):
            t0 = time.perf_counter()
            z = m.conv_features(x)
            t1 = time.perf_counter()
            m.forward_from_conv_features(z)
            t2 = time.perf_counter()
            tb.append((t1 - t0) * 1e3), th.append((t2 - t1) * 1e3)
    b, h = min(tb), min(th)
    print(json.dumps({"workload": "C1: reference PPNetMultiScale forward, 1x3x513x513, random-init DeepLabv2-R101 multi-scale, P=760, K=19, CPU",
                      "cores": cores, "backbone_ms": b, "head_ms": h, "head_share": h / (b + h), "feature_map": list(z.shape),
                      "note": "best of 5 after 1 warm-up; measured in the build container (the GPU box has no /root/reference)"}))


if __name__ == "__main__":
    main()

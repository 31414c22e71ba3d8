"""No-GPU checks of the drop-in boundary: the library builds, loads and exports every symbol the
header declares; argument validation works without touching a device."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "scaleproto_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sps_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_exported(built_lib):
    syms = _declared_symbols()
    assert len(syms) >= 10
    raw = C.CDLL(os.path.join(ROOT, "scaleprotoseg_b200", "libscaleproto_sm100.so"))
    for s in syms:
        assert hasattr(raw, s), f"{s} declared in the header but not exported"


def test_binding_covers_header(built_lib):
    from scaleprotoseg_b200 import EXPORTED_SYMBOLS
    assert sorted(EXPORTED_SYMBOLS) == _declared_symbols()


def test_version_and_validation(built_lib):
    from scaleprotoseg_b200._lib import SpsHeadDesc
    assert built_lib.sps_version() == 2
    d = SpsHeadDesc()
    d.B, d.H, d.W, d.S, d.D, d.P, d.K = 1, 4, 4, 4, 64, 24, 2
    for i in range(9):
        d.scale_lo[i] = min(i, 4) * 6
    d.scale_lo[4] = 23  # does not end at P
    rc = built_lib.sps_head_fwd(C.byref(d), None, None, None, None, None, None, None, 0, None)
    assert rc == 1 and b"scale_lo" in built_lib.sps_last_error()
    d.scale_lo[4] = 24
    for i in range(5, 9):
        d.scale_lo[i] = 24
    rc = built_lib.sps_head_fwd(C.byref(d), None, None, None, None, None, None, None, 0, None)
    assert rc == 1 and b"NULL" in built_lib.sps_last_error()
    assert built_lib.sps_workspace_bytes(C.byref(d)) > 0
    # feature-map element type / layout (z_dtype, z_layout): 16-bit maps are channels-last and need the matching math
    d.z_dtype, d.z_layout, d.math = 1, 0, 3                     # fp16 but NCHW
    rc = built_lib.sps_head_fwd(C.byref(d), None, None, None, None, None, None, None, 0, None)
    assert rc == 2 and b"channels-last" in built_lib.sps_last_error()
    d.z_layout, d.math = 1, 2                                   # fp16 channels-last with bf16 math
    rc = built_lib.sps_head_fwd(C.byref(d), None, None, None, None, None, None, None, 0, None)
    assert rc == 2 and b"does not match" in built_lib.sps_last_error()
    d.math = 4                                                  # fp16 channels-last, 3-term fp16 math: valid (NULL pointers)
    rc = built_lib.sps_head_fwd(C.byref(d), None, None, None, None, None, None, None, 0, None)
    assert rc == 1 and b"NULL" in built_lib.sps_last_error()


def test_struct_layout_matches_header():
    from scaleprotoseg_b200._lib import SpsGroupDesc, SpsHeadDesc
    assert C.sizeof(SpsHeadDesc) == 4 * (7 + 9 + 4 + 2)
    assert C.sizeof(SpsGroupDesc) == 8 + 6 * 8


def test_cpu_tensor_is_rejected(built_lib):
    import torch
    import scaleprotoseg_b200 as sps
    cfg = sps.HeadConfig.from_ranges([(0, 6), (6, 12), (12, 18), (18, 24)], 2)
    with pytest.raises(RuntimeError, match="no CPU path"):
        sps.head_forward_raw(torch.zeros(1, 256, 2, 2), torch.zeros(24, 64), None, cfg, True, False)

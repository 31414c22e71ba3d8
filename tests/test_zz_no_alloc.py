"""`sps_head_fwd_static` — the forward that allocates nothing and keeps its scratch in a static device array — against
`sps_head_fwd` (caller workspace), through the C ABI. No counterpart in the reference;
collected last so that a failure here hides no other test."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
_SETUP_DONE = [False]
_STATIC_BYTES = 8 << 20


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _case(lib, head, sps, ho, seed, B, H, W, P, K, math, fp16_maps, dense):
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(seed)
    S, D = 4, 64
    z = torch.sigmoid(torch.randn(B, S * D, H, W, generator=g)).to(dev)
    if fp16_maps:
        z = z.half().contiguous(memory_format=torch.channels_last)
    protos = torch.rand(P, D, generator=g).to(dev)
    last_w = ho.make_last_layer(ho.make_identity(P, K, S)).contiguous().to(dev)
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, S), K, math=math)
    desc = head.make_desc(cfg, B, H, W, D, P, z)
    N = B * H * W

    def outputs():
        return (torch.full((N, K), float("nan"), device=dev),
                torch.full((B, P, H, W), float("nan"), device=dev) if dense else None,
                torch.full((N, P), float("nan"), device=dev) if dense else None)

    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # existing entry point, caller workspace
    nbytes = int(lib.sps_head_fwd_workspace_bytes(C.byref(desc)))
    ws = torch.empty(max(nbytes, 1), device=dev, dtype=torch.uint8)
    ref = outputs()
    rc = lib.sps_head_fwd(C.byref(desc), _ptr(z), _ptr(protos), _ptr(last_w), _ptr(ref[0]), _ptr(ref[1]), _ptr(ref[2]),
                          _ptr(ws), nbytes, stream)
    assert rc == 0, lib.sps_last_error()
    # new entry point: outputs allocated BEFORE the call, nothing may be allocated by it
    got = outputs()
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free_before, _ = torch.cuda.mem_get_info()
    rc = lib.sps_head_fwd_static(C.byref(desc), _ptr(z), _ptr(protos), _ptr(last_w), _ptr(got[0]), _ptr(got[1]), _ptr(got[2]),
                                 stream)
    assert rc == 0, lib.sps_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    # no driver-level allocation either — except that the FIRST use in a process makes the runtime load the module that
    # holds the static array (lazy loading), which is when its 8 MiB appear on the device: set-up, once per device
    if _SETUP_DONE[0]:
        assert torch.cuda.mem_get_info()[0] >= free_before
    else:
        assert free_before - torch.cuda.mem_get_info()[0] <= _STATIC_BYTES + (4 << 20)
    _SETUP_DONE[0] = True
    for a, b, name in zip(got, ref, ("logits", "dist", "act")):
        if a is None:
            continue
        assert torch.isfinite(a).all(), name
        # same kernels, same operands, same order of operations: the two entry points agree bit for bit; the tolerance
        # is the one the tensor-core parity tests use for this mode (tests/test_gpu_parity.py), kept as the stated bar
        tol = 2e-3 if math in (sps.MATH_F16, sps.MATH_BF16) else 2e-5
        scale = max(float(b.abs().max()), 1.0)
        assert float((a - b).abs().max()) <= tol * scale, name
        assert torch.equal(a, b), name


def test_static_forward_matches_workspace_forward_two_calls_in_a_row():
    import scaleprotoseg_b200 as sps
    from scaleprotoseg_b200 import _lib, head
    from oracle import head_oracle as ho
    lib = _lib.load()
    # call 1: the larger shape in 3-term mode fills the static array; call 2: a smaller, ragged shape in another mode
    # and another map format finds call 1's bytes in it and must not be affected by them; call 3 repeats call 1
    _case(lib, head, sps, ho, 11, 2, 33, 29, 760, 19, sps.MATH_F16X3, False, True)
    _case(lib, head, sps, ho, 12, 1, 17, 13, 228, 19, sps.MATH_F16, True, False)
    _case(lib, head, sps, ho, 13, 2, 33, 29, 760, 19, sps.MATH_F16X3, False, False)
    _case(lib, head, sps, ho, 14, 1, 9, 7, 228, 19, sps.MATH_FP32, False, True)   # needs no scratch at all


def test_static_forward_rejects_shapes_beyond_the_static_array():
    import scaleprotoseg_b200 as sps
    from scaleprotoseg_b200 import _lib, head
    from oracle import head_oracle as ho
    lib = _lib.load()
    dev = torch.device("cuda:0")
    P, K, S, D = 6840, 171, 4, 64
    # the COCO-Stuff shape on fp32 maps: more than 32 classes need the 16-bit re-lay of the map inside the scratch (145 MB)
    B, H, W = 32, 65, 65
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, S), K, math=sps.MATH_F16X3)
    desc = head.make_desc(cfg, B, H, W, D, P)
    need = int(lib.sps_head_fwd_workspace_bytes(C.byref(desc)))
    assert need > _lib.STATIC_SCRATCH_BYTES
    dummy = torch.zeros(1024, device=dev)               # rejected before any pointer is touched
    rc = lib.sps_head_fwd_static(C.byref(desc), _ptr(dummy), _ptr(dummy), _ptr(dummy), _ptr(dummy), None, None, None)
    assert rc == 3   # SPS_ERR_WORKSPACE
    msg = lib.sps_last_error().decode()
    assert str(need) in msg and str(_lib.STATIC_SCRATCH_BYTES) in msg, msg
    # the same prototype set fits with 16-bit channels-last maps
    z16 = torch.zeros(1, dtype=torch.float16)
    desc = head.make_desc(cfg, 1, 9, 9, D, P, z16)
    need16 = int(lib.sps_head_fwd_workspace_bytes(C.byref(desc)))
    assert 0 < need16 <= _lib.STATIC_SCRATCH_BYTES

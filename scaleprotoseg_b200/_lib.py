"""ctypes binding of libscaleproto_sm100.so (the C ABI declared in include/scaleproto_b200.h).

There is no CPU or PyTorch fallback: if the library is missing, importing the compute entry points
raises.  Build it with `python __graft_entry__.py` (or `__graft_entry__.build()`).
"""
import ctypes as C
import os

SPS_MAX_SCALES = 8
ACT_LOG, ACT_LINEAR = 0, 1
MATH_FP32, MATH_BF16X3, MATH_BF16, MATH_F16, MATH_F16X3 = 0, 1, 2, 3, 4
FLAG_DETERMINISTIC = 1
FLAG_ZPACK = 2       # fp32 NCHW input: 16-bit channels-last re-lay in a pre-pass, then the TMA-fed kernels
FLAG_ZRELAYED = 4    # the z pointer is a re-lay written earlier by sps_z_relay
Z_F32, Z_F16, Z_BF16 = 0, 1, 2
Z_NCHW, Z_NHWC = 0, 1

LIB_NAME = "libscaleproto_sm100.so"
LIB_PATH = os.environ.get("SCALEPROTO_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), LIB_NAME)


class SpsHeadDesc(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("H", C.c_int32), ("W", C.c_int32),
        ("S", C.c_int32), ("D", C.c_int32),
        ("P", C.c_int32), ("K", C.c_int32),
        ("scale_lo", C.c_int32 * (SPS_MAX_SCALES + 1)),
        ("act_kind", C.c_int32),
        ("eps", C.c_float),
        ("math", C.c_int32),
        ("flags", C.c_int32),
        ("z_dtype", C.c_int32),
        ("z_layout", C.c_int32),
    ]


class SpsGroupDesc(C.Structure):
    _fields_ = [
        ("Kc", C.c_int32), ("J", C.c_int32),
        ("proto_slot", C.c_void_p), ("proto_col", C.c_void_p), ("class_off", C.c_void_p),
        ("members", C.c_void_p), ("G", C.c_void_p), ("Wg", C.c_void_p),
    ]


class SpsKldDesc(C.Structure):
    _fields_ = [
        ("G", C.c_int32), ("Q", C.c_int32), ("mmax", C.c_int32), ("reserved", C.c_int32),
        ("proto_class", C.c_void_p), ("proto_group", C.c_void_p), ("proto_local", C.c_void_p),
        ("group_moff", C.c_void_p), ("group_poff", C.c_void_p), ("members", C.c_void_p),
        ("pair_a", C.c_void_p), ("pair_b", C.c_void_p),
    ]


_P = C.c_void_p
STATIC_SCRATCH_BYTES = 8 << 20   # SPS_STATIC_SCRATCH_BYTES

_SIGNATURES = {
    "sps_version": (C.c_int, []),
    "sps_last_error": (C.c_char_p, []),
    "sps_launch_count": (C.c_uint64, []),
    "sps_fp32_reroute_count": (C.c_uint64, []),
    "sps_z_relay_bytes": (C.c_size_t, [C.POINTER(SpsHeadDesc)]),
    "sps_z_relay": (C.c_int, [C.POINTER(SpsHeadDesc), _P, _P, C.c_size_t, _P]),
    "sps_workspace_bytes": (C.c_size_t, [C.POINTER(SpsHeadDesc)]),
    "sps_head_fwd_workspace_bytes": (C.c_size_t, [C.POINTER(SpsHeadDesc)]),
    "sps_head_fwd": (C.c_int, [C.POINTER(SpsHeadDesc), _P, _P, _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "sps_head_fwd_static": (C.c_int, [C.POINTER(SpsHeadDesc), _P, _P, _P, _P, _P, _P, _P]),
    "sps_head_bwd": (C.c_int, [C.POINTER(SpsHeadDesc), _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "sps_group_fwd": (C.c_int, [C.POINTER(SpsHeadDesc), C.POINTER(SpsGroupDesc), _P, _P, _P, _P]),
    "sps_group_workspace_bytes": (C.c_size_t, [C.POINTER(SpsHeadDesc), C.POINTER(SpsGroupDesc)]),
    "sps_group_bwd": (C.c_int, [C.POINTER(SpsHeadDesc), C.POINTER(SpsGroupDesc), _P, _P, _P, _P, _P, _P, _P, _P,
                                C.c_size_t, _P]),
    "sps_kld_workspace_bytes": (C.c_size_t, [C.POINTER(SpsHeadDesc), C.POINTER(SpsKldDesc)]),
    "sps_kld_fwd": (C.c_int, [C.POINTER(SpsHeadDesc), C.POINTER(SpsKldDesc), _P, _P, _P, C.c_size_t, _P, _P]),
    "sps_kld_bwd": (C.c_int, [C.POINTER(SpsHeadDesc), C.POINTER(SpsKldDesc), _P, _P, C.c_size_t, _P, _P, _P]),
    "sps_kld_feat_workspace_bytes": (C.c_size_t, [C.POINTER(SpsHeadDesc), C.POINTER(SpsKldDesc)]),
    "sps_kld_feat_fwd": (C.c_int, [C.POINTER(SpsHeadDesc), C.POINTER(SpsKldDesc), _P, _P, _P, _P, C.c_size_t, _P, _P]),
    "sps_kld_feat_bwd": (C.c_int, [C.POINTER(SpsHeadDesc), C.POINTER(SpsKldDesc), _P, _P, _P, _P, C.c_size_t, _P, _P, _P,
                                   _P]),
    "sps_upsample_argext": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                      _P, _P, _P]),
    "sps_push_scan": (C.c_int, [C.POINTER(SpsHeadDesc), _P, _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "sps_push_reduce": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_uint32, _P, _P]),
    "sps_push_topk": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_uint32, C.c_int32, _P, _P, _P]),
    "sps_push_gather": (C.c_int, [C.POINTER(SpsHeadDesc), _P, _P, _P, C.c_int32, C.c_uint32, _P, _P, _P]),
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib = None


def load():
    """dlopen the in-tree library and attach the prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
            "There is no CPU fallback for the prototype head.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().sps_last_error()
        raise RuntimeError(f"{what} failed (status {rc}): {msg.decode() if msg else ''}")

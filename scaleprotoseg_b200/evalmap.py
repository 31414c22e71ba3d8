"""Host side of the evaluation epilogue (segmentation/eval_valid_multiscale.py:229-234): bilinear upsampling to the
annotation size + argmax / argmin over channels, through `sps_upsample_argext`.  CUDA only."""
import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from .head import _check_cuda_f32, _ptr, _stream


def _argext(maps: torch.Tensor, size: Sequence[int], argmin: bool, want_value: bool):
    lib = _lib.load()
    _check_cuda_f32("maps", maps)
    squeeze = maps.dim() == 3
    if squeeze:
        maps = maps.unsqueeze(0)
    if maps.dim() != 4:
        raise RuntimeError(f"maps must be (B,C,h,w) or (C,h,w), got {tuple(maps.shape)}")
    maps = maps.contiguous()
    B, Cc, h, w = maps.shape
    H, W = int(size[0]), int(size[1])
    idx = torch.empty((B, H, W), device=maps.device, dtype=torch.int32)
    val = torch.empty((B, H, W), device=maps.device, dtype=torch.float32) if want_value else None
    with torch.cuda.device(maps.device):
        _lib.check(lib.sps_upsample_argext(_ptr(maps), B, Cc, h, w, H, W, int(argmin), _ptr(idx), _ptr(val), _stream()),
                   "sps_upsample_argext")
    idx = idx.to(torch.int64)       # torch.argmax / argmin return int64
    if squeeze:
        idx, val = idx[0], (val[0] if val is not None else None)
    return (idx, val) if want_value else idx


def upsample_argmax(logits: torch.Tensor, size: Sequence[int], return_value: bool = False):
    """torch.argmax(F.interpolate(logits, size, mode="bilinear", align_corners=False), dim=channel)
    (eval_valid_multiscale.py:229,233) for channels-first logits (B,K,h,w) or (K,h,w)."""
    return _argext(logits, size, False, return_value)


def upsample_argmin(distances: torch.Tensor, size: Sequence[int], return_value: bool = False):
    """torch.argmin(F.interpolate(distances, size, mode="bilinear", align_corners=False), dim=channel)
    (eval_valid_multiscale.py:230-231): the nearest prototype per full-resolution pixel."""
    return _argext(distances, size, True, return_value)

"""GPU parity of the evaluation epilogue (sps_upsample_argext) against the reference's own lines
(segmentation/eval_valid_multiscale.py:229-234): F.interpolate(..., mode="bilinear", align_corners=False) followed by
torch.argmax / torch.argmin over channels, evaluated on the CPU in fp32.

Bar: the winning VALUE within 1e-5 * max(|ref|, 1); the INDEX bit-exact wherever the runner-up is further than 2e-5 *
max(|ref|, 1) from the winner (closer than that, the value at the returned index must still be within the tolerance)."""
import pytest
import torch
import torch.nn.functional as F

import scaleprotoseg_b200 as sps

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _check(maps, size, argmin):
    up = F.interpolate(maps, size=size, mode="bilinear", align_corners=False)
    ref_idx = up.argmin(dim=1) if argmin else up.argmax(dim=1)
    ref_val = up.gather(1, ref_idx.unsqueeze(1)).squeeze(1)
    fn = sps.upsample_argmin if argmin else sps.upsample_argmax
    idx, val = fn(maps.to(DEV), size, return_value=True)
    assert idx.dtype == torch.int64 and tuple(idx.shape) == tuple(ref_idx.shape)
    idx, val = idx.cpu(), val.cpu()
    tol = 1e-5 * ref_val.abs().clamp_min(1.0)
    assert ((val - ref_val).abs() <= tol).all()
    at_idx = up.gather(1, idx.unsqueeze(1)).squeeze(1)
    assert ((at_idx - ref_val).abs() <= 2 * tol).all()
    if maps.shape[1] < 2:
        assert (idx == 0).all()
        return
    top2 = (up if argmin else -up).topk(2, dim=1, largest=False).values
    clear = (top2[:, 1] - top2[:, 0]) > 2 * tol
    assert torch.equal(idx[clear], ref_idx[clear])
    assert clear.float().mean() > 0.99


@pytest.mark.parametrize("argmin", [False, True])
@pytest.mark.parametrize("shape", [(2, 19, 9, 7, (65, 50)), (1, 228, 17, 33, (129, 257)), (1, 5, 4, 4, (4, 4)),
                                   (1, 3, 6, 5, (3, 2)), (3, 1, 2, 2, (9, 9))])
def test_upsample_argext_vs_torch(built_lib, shape, argmin):
    B, Cc, h, w, size = shape
    g = torch.Generator().manual_seed(1234 + Cc + h)
    _check(torch.randn(B, Cc, h, w, generator=g) * 3.0, size, argmin)


def test_eval_lines_at_cityscapes_size(built_lib):
    """One validation image as eval_valid_multiscale.py:219-234 handles it: logits (19,129,257) and distances
    (760,129,257) upsampled to the 1024x2048 annotation; the 6.4 GB upsampled distance map is never built here, so the
    reference side is evaluated in horizontal strips."""
    g = torch.Generator().manual_seed(99)
    K, P, h, w, H, W = 19, 760, 129, 257, 1024, 2048
    logits = torch.randn(1, K, h, w, generator=g)
    dist = torch.rand(1, P, h, w, generator=g) * 8.0
    pred = sps.upsample_argmax(logits.to(DEV)[0], (H, W)).cpu()
    near, nval = sps.upsample_argmin(dist.to(DEV), (H, W), return_value=True)
    near, nval = near.cpu()[0], nval.cpu()[0]
    assert tuple(pred.shape) == (H, W) and tuple(near.shape) == (H, W)
    up_l = F.interpolate(logits, size=(H, W), mode="bilinear", align_corners=False)[0]
    top2 = up_l.topk(2, dim=0).values
    clear = (top2[0] - top2[1]) > 1e-4
    assert torch.equal(pred[clear], up_l.argmax(0)[clear]) and clear.float().mean() > 0.99
    # distances: rows 0..63 and the last 64 rows of the full-resolution map (strips of the reference computation)
    up_d = F.interpolate(dist.to(DEV), size=(H, W), mode="bilinear", align_corners=False)[0]   # torch on the GPU, 6.4 GB
    ref_val, ref_idx = up_d.min(dim=0)
    ref_val, ref_idx = ref_val.cpu(), ref_idx.cpu()
    del up_d
    assert ((nval - ref_val).abs() <= 1e-4).all()
    assert (near == ref_idx).float().mean() > 0.999

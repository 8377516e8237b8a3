"""Drop-in for the functions of the reference's python/utils.py that lie on the coded path.

rgb_to_yuv (python/utils.py:4-33) is what python/encode_ply.py:108 applies to the point-cloud colours before RAHT; here it
is one CUDA kernel (csrc/color.cu).  CUDA tensors only: this package has no CPU path."""
import torch

from raht_b200 import _lib as L
from raht_b200._lib import lib, check, ptr, stream


def rgb_to_yuv(rgb: torch.Tensor) -> torch.Tensor:
    """RGB -> YUV (BT.709 full range with 128/255 chroma offsets), Nx3 in [0, 255] -> Nx3 float64 in [0, 255], clipped."""
    if rgb.ndim != 2 or rgb.shape[1] != 3:
        raise ValueError("Expected Nx3 tensor")                   # utils.py:10-11
    L.require_cuda(rgb, "rgb")
    if rgb.dtype not in (torch.float32, torch.float64):
        rgb = rgb.to(torch.float64)
    rgb = rgb.contiguous()
    out = torch.empty(rgb.shape[0], 3, dtype=torch.float64, device=rgb.device)
    check(lib.raht_rgb_to_yuv(ptr(rgb), L.dtype_tag(rgb.dtype), rgb.shape[0], ptr(out), stream()), "raht_rgb_to_yuv")
    return out


def sanity_check_vector(T: torch.Tensor, C: torch.Tensor, rtol=1e-5, atol=1e-8) -> bool:
    """max(T) == sqrt(N) * mean(C)  (python/utils.py:46-57): the DC check the reference drivers print."""
    assert T.dim() == 1 and C.dim() == 1 and T.size(0) == C.size(0), "T and C must be 1D with same length"
    lhs = T.max()
    rhs = torch.sqrt(torch.tensor(float(T.size(0)), dtype=C.dtype, device=C.device)) * C.mean()
    return bool(torch.allclose(lhs, rhs, rtol=rtol, atol=atol))

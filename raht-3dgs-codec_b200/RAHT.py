"""Drop-in for python/RAHT.py of Daaohame/RAHT-3DGS-codec.

RAHT2_optimized (reference python/RAHT.py:253-336) is the production transform; RAHT, RAHT2 and
RAHT_optimized (:6-80, :84-167, :171-249) are bound to the same fp64 semantics here (the reference's
non-"2" variants compute the rotation in fp32 by accident -- SURVEY App. C).
Returns (T (N, D) float64, w (N, 1) float64) like the reference.
"""
import torch

from raht_b200 import ops


@torch.no_grad()
def RAHT2_optimized(C, List, Flags, weights, one_based: bool = False):
    if not isinstance(C, torch.Tensor) or not C.is_cuda:
        raise RuntimeError("RAHT: C must be a CUDA tensor (this build has no CPU path)")
    tree = ops.tree_of(List, Flags, weights, one_based)
    squeeze = C.dim() == 1
    Cm = C.reshape(C.shape[0], -1)
    T = ops.forward(tree, Cm, torch.float64)
    w = tree.w.clone().reshape(-1, 1)
    return (T.reshape(-1) if squeeze else T), w


RAHT = RAHT2 = RAHT_optimized = RAHT2_optimized

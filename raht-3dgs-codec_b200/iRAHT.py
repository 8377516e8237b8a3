"""Drop-in for python/iRAHT.py of Daaohame/RAHT-3DGS-codec.

inverse_RAHT_optimized (reference python/iRAHT.py:40-114, production) and inverse_RAHT (:5-37);
the drivers bind it as raht_fn["iRAHT"] (encode_3dgs.py:23-27), hence the iRAHT alias.
"""
import torch

from raht_b200 import ops


@torch.no_grad()
def inverse_RAHT_optimized(T, List, Flags, weights, one_based: bool = False):
    if not isinstance(T, torch.Tensor) or not T.is_cuda:
        raise RuntimeError("iRAHT: T must be a CUDA tensor (this build has no CPU path)")
    tree = ops.tree_of(List, Flags, weights, one_based)
    Tm = T.reshape(T.shape[0], -1).to(torch.float64)
    return ops.inverse(tree, Tm).reshape(T.shape)


inverse_RAHT = iRAHT = inverse_RAHT_optimized

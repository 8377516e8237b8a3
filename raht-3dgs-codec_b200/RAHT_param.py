"""Drop-in for python/RAHT_param.py of Daaohame/RAHT-3DGS-codec: same names, arguments and returns.

RAHT_param            (reference python/RAHT_param.py:6-83)
RAHT_param_reorder    (:86-187)
RAHT_param_reorder_fast (:191-279)  -- the production entry point (encode_3dgs.py:23-27,149)

The returned List / Flags / weights are lazy sequences (raht_b200.ops.LevelSequence): indexing a
level materialises exactly the tensors the reference returns (int64 / bool / int64, 0-based unless
return_one_based), while RAHT()/iRAHT() recognise them and run the fused plan instead.

Divergences: rows that are not Morton-sorted and unique raise ValueError (the
reference silently produces garbage); the out-of-bounds check of RAHT_param is applied in the
_fast variant too; for N <= 4 clouds whose tree is lower than two bit levels the reference's
order_RAGFT is not a permutation ([0,0] for two siblings) -- here it always is.
"""
import torch

from raht_b200 import ops
from raht_b200 import _lib as L


def _as_min3(minV):
    if isinstance(minV, torch.Tensor):
        v = minV.detach().to("cpu", torch.float64).reshape(-1).tolist()
    elif isinstance(minV, (int, float)):
        v = [float(minV)] * 3
    else:
        v = [float(x) for x in minV]
    if len(v) == 1:
        v = v * 3
    if len(v) != 3:
        raise ValueError("minV must have 3 components")
    return v


def _validated_tree(V, minV, width, depth):
    if not isinstance(V, torch.Tensor) or not V.is_cuda:
        raise RuntimeError("RAHT_param: V must be a CUDA tensor (this build has no CPU path)")
    if V.dim() != 2 or V.shape[1] != 3:
        raise ValueError("V must be (N, 3)")
    if not (1 <= 3 * depth <= 63):
        raise ValueError("RAHT_param: 3*depth must be <= 63")
    MC, status = ops.morton_keys_f64(V, _as_min3(minV), float(width), int(depth))
    if int(status.item()) & L.ST_OUT_OF_BOUNDS:
        raise ValueError("RAHT_param:OutOfBounds - voxel index out of bounds")      # RAHT_param.py:26-27
    return ops.tree_from_keys(MC, validate=True)


def _sequences(tree, one_based=False):
    cache = {}
    return (ops.LevelSequence(tree, 0, one_based, cache), ops.LevelSequence(tree, 1, False, cache),
            ops.LevelSequence(tree, 2, False, cache))


@torch.no_grad()
def RAHT_param(V, minV, width, depth, return_one_based: bool = False):
    """List, Flags, weights = RAHT_param(V, minV, width, depth)   (RAHT_param.py:6-83)."""
    tree = _validated_tree(V, minV, width, depth)
    return _sequences(tree, return_one_based)


@torch.no_grad()
def RAHT_param_reorder_fast(V, minV, width, depth):
    """List, Flags, weights, order_RAGFT = RAHT_param_reorder_fast(V, minV, width, depth)  (:191-279)."""
    tree = _validated_tree(V, minV, width, depth)
    Ls, Fs, Ws = _sequences(tree)
    order = tree.order() if tree.n > 1 else None          # N == 1: the reference returns None (:233-236)
    if order is not None:
        order._raht_tree = tree
    return Ls, Fs, Ws, order


RAHT_param_reorder = RAHT_param_reorder_fast               # :86-187, same outputs

"""Drop-in for cuda/merge_cluster_cuda/__init__.py of Daaohame/RAHT-3DGS-codec.

merge_gaussian_clusters_with_indices (reference cuda/merge_cluster_cuda/__init__.py:149-204),
merge_gaussian_clusters (:77-146), prepare_cluster_data (:30-74); the torch-extension `_C` and its
kernel (cuda/merge_cluster_wrapper.cu, cuda/merge_cluster.cu) are replaced by raht_merge_segments
of libraht_b200.so.  Errors follow the reference: RuntimeError for non-CUDA / shape / dtype problems.
"""
from typing import Tuple

import torch

from raht_b200 import ops

__all__ = ['merge_gaussian_clusters', 'merge_gaussian_clusters_with_indices', 'prepare_cluster_data']


def prepare_cluster_data(cluster_labels: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """labels [N] -> (cluster_indices int32 [N] grouped by ascending label, cluster_offsets int32 [K+1])."""
    if not cluster_labels.is_cuda:
        raise RuntimeError("cluster_labels must be a CUDA tensor")
    labels = cluster_labels.contiguous().long()
    lo = int(labels.min().item()) if labels.numel() else 0
    keys = labels - lo                                        # non-negative keys for the radix sort
    hi = int(keys.max().item()) if labels.numel() else 0
    keys_sorted, idx, _ = ops.radix_sort_pairs(keys, max(hi.bit_length(), 1))
    _, offsets, _, _ = ops.voxel_segments(keys_sorted, want_unique=False)
    return idx, offsets.clone()


def _check(means, quats, scales, opacities, colors, cluster_indices, cluster_offsets):
    names = ("cluster_indices", "cluster_offsets", "means", "quats", "scales", "opacities", "colors")
    for n, t in zip(names, (cluster_indices, cluster_offsets, means, quats, scales, opacities, colors)):
        if not t.is_cuda:
            raise RuntimeError(f"{n} must be a CUDA tensor")                      # merge_cluster_wrapper.cu:22-28
    if means.dim() != 2 or means.size(1) != 3:
        raise RuntimeError("means must be [N, 3]")                               # :38-42
    if quats.dim() != 2 or quats.size(1) != 4:
        raise RuntimeError("quats must be [N, 4]")
    if scales.dim() != 2 or scales.size(1) != 3:
        raise RuntimeError("scales must be [N, 3]")
    if opacities.dim() != 1:
        raise RuntimeError("opacities must be [N]")
    if colors.dim() != 2:
        raise RuntimeError("colors must be [N, color_dim]")
    N = means.size(0)
    for n, t in (("quats", quats), ("scales", scales), ("opacities", opacities), ("colors", colors)):
        if t.size(0) != N:
            raise RuntimeError(f"{n} must have same length as means")             # :45-48


def merge_gaussian_clusters_with_indices(means, quats, scales, opacities, colors, cluster_indices, cluster_offsets,
                                         weight_by_opacity: bool = True):
    means = means.contiguous().float()
    quats = quats.contiguous().float()
    scales = scales.contiguous().float()
    opacities = opacities.contiguous().float()
    colors = colors.contiguous().float()
    cluster_indices = cluster_indices.contiguous().int()
    cluster_offsets = cluster_offsets.contiguous().int()
    _check(means, quats, scales, opacities, colors, cluster_indices, cluster_offsets)
    outs, _ = ops.merge_segments(cluster_indices, cluster_offsets, means, quats, scales, opacities, colors,
                                 weight_by_opacity, separate=True)
    return tuple(outs)


def merge_gaussian_clusters(means, quats, scales, opacities, colors, cluster_labels, weight_by_opacity: bool = True):
    for t in (means, quats, scales, opacities, colors, cluster_labels):
        assert t.is_cuda, "All inputs must be CUDA tensors"
    N = means.shape[0]
    assert cluster_labels.shape[0] == N, "cluster_labels must have same length as inputs"
    cluster_indices, cluster_offsets = prepare_cluster_data(cluster_labels)
    return merge_gaussian_clusters_with_indices(means, quats, scales, opacities, colors, cluster_indices,
                                                cluster_offsets, weight_by_opacity)

"""Drop-in for python/voxelize_pc.py of Daaohame/RAHT-3DGS-codec.

get_morton_code       (reference python/voxelize_pc.py:25-59)
voxelize_pc_batched   (:62-172)
voxelize_pc           (:175-283, dict interface; file output needs open3d and is not provided)

Same arguments and returns.  torch.sort (:101) is replaced by a stable radix sort, so Gaussians that
share a voxel keep their input order (the reference's tie order is unspecified).
"""
from typing import Dict, Optional, Tuple

import torch

from raht_b200 import ops
from raht_b200 import _lib as L
from raht_b200._lib import lib, check, ptr, stream


def get_morton_code(V: torch.Tensor, J: int) -> torch.Tensor:
    if not V.is_cuda:
        raise RuntimeError("get_morton_code: V must be a CUDA tensor (this build has no CPU path)")
    return ops.morton_keys_i64(V, J)


@torch.no_grad()
def voxelize_pc_batched(PC: torch.Tensor, vmin: Optional[torch.Tensor] = None, width: Optional[float] = None,
                        J: int = 10, device: str = 'cuda') -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor, Dict]:
    if not str(device).startswith('cuda'):
        raise RuntimeError("voxelize_pc_batched: CUDA only (this build has no CPU path)")
    PC = PC.to(device)
    if PC.dtype != torch.float32:
        PC = PC.float()       # the reference runs in the dtype of PC; 3DGS means are fp32
    PC = PC.contiguous()
    dev = PC.device
    N, ld = PC.shape
    d = ld - 3
    st = stream()

    bounds = torch.empty(6, dtype=torch.float32, device=dev)
    params4 = torch.empty(4, dtype=torch.float32, device=dev)
    params_d2 = torch.empty(2, dtype=torch.float64, device=dev)
    check(lib.raht_bounds3_f32(ptr(PC), N, ld, ptr(bounds), st), "raht_bounds3_f32")
    import ctypes as C
    vm = None
    if vmin is not None:
        vm = (C.c_float * 3)(*[float(x) for x in vmin.detach().cpu().reshape(-1).tolist()])
    check(lib.raht_voxel_params(ptr(bounds), J, 1 if vmin is not None else 0, vm, 1 if width is not None else 0,
                                float(width) if width is not None else 0.0, ptr(params4), ptr(params_d2), st),
          "raht_voxel_params")
    keys = torch.empty(N, dtype=torch.int64, device=dev)
    check(lib.raht_morton_keys_f32(ptr(PC), N, ld, ptr(params4), J, ptr(keys), None, st), "raht_morton_keys_f32")
    keys_sorted, idx32, status = ops.radix_sort_pairs(keys, 3 * J)
    voxel_indices, _, unique_keys, Nvox = ops.voxel_segments(keys_sorted)
    if int(status.item()) & L.ST_SORT_STALL:
        raise RuntimeError("radix sort look-back stalled (internal error)")

    PCvox = torch.empty(Nvox, ld, dtype=torch.float32, device=dev)
    if d == 0:
        check(lib.raht_morton_decode_f32(ptr(unique_keys), Nvox, J, ptr(PCvox), st), "raht_morton_decode_f32")
    else:
        Vvox = torch.empty(Nvox, 3, dtype=torch.float32, device=dev)
        check(lib.raht_morton_decode_f32(ptr(unique_keys), Nvox, J, ptr(Vvox), st), "raht_morton_decode_f32")
        PCvox[:, :3] = Vvox
    PCsorted = torch.empty(N, ld, dtype=torch.float32, device=dev)
    DeltaPC = torch.empty(N, ld, dtype=torch.float32, device=dev)
    check(lib.raht_gather_rows_f32(ptr(PC), N, ld, ptr(idx32), ptr(params4), ptr(PCsorted), ptr(DeltaPC), ld, st),
          "raht_gather_rows_f32")
    if d > 0:
        check(lib.raht_voxel_attr_mean_f32(ptr(PC), N, ld, d, ptr(idx32), ptr(voxel_indices), Nvox, ptr(PCvox), ld, 3,
                                           ptr(DeltaPC), ld, st), "raht_voxel_attr_mean_f32")
    pd = params_d2.cpu().tolist()
    info = {
        'Nvox': Nvox,
        'voxel_size': pd[1],
        'vmin': vmin.to(dev) if vmin is not None else params4[:3].clone(),
        'width': pd[0],
        'N': N,
        'sort_idx': idx32.to(torch.int64),
    }
    return PCvox, PCsorted, voxel_indices, DeltaPC, info


def voxelize_pc(PC: torch.Tensor, param: dict):
    vmin = param.get('vmin')
    width = param.get('width')
    J = param.get('J')
    if param.get('writeFileOut', False):
        raise ImportError("open3d is required for file writing; not available in this build")   # voxelize_pc.py:229-230
    if vmin is not None and not isinstance(vmin, torch.Tensor):
        vmin = torch.tensor(vmin, dtype=PC.dtype, device=PC.device)
    PCvox, PCsorted, voxel_indices, DeltaPC, _ = voxelize_pc_batched(PC=PC, vmin=vmin, width=width, J=J,
                                                                     device=str(PC.device))
    return PCvox, PCsorted, voxel_indices, DeltaPC

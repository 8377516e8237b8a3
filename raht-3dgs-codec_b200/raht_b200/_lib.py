"""ctypes binding of libraht_b200.so (the C ABI declared in include/raht_b200.h).

This is the only place the package touches the native library.  There is no CPU fallback and no
alternative backend: if the library is missing, importing any operator module fails loudly.
PyTorch is used for device memory, streams and (multi-GPU) torch.distributed only.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(_HERE, "libraht_b200.so")

RAHT_F32, RAHT_F64 = 0, 1
ST_OUT_OF_BOUNDS, ST_NOT_SORTED_UNIQUE, ST_INT32_OVERFLOW, ST_STREAM_OVERFLOW, ST_SORT_STALL = 1, 2, 4, 8, 16
FLAG_NO_TMA = 1

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `make -C {_HERE} -j` (or __graft_entry__.build()). "
        "raht-3dgs-codec_b200 has no CPU fallback."
    )

lib = C.CDLL(LIB_PATH)

_p, _i, _i64, _sz, _d = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_double

# name -> (restype, argtypes); mirrors include/raht_b200.h exactly (tests check every symbol resolves)
SIGNATURES = {
    "raht_abi_version": (_i, []),
    "raht_last_error": (C.c_char_p, []),
    "raht_launch_count": (C.c_longlong, []),
    "raht_device_info": (_i, [C.POINTER(_i), C.POINTER(_i)]),
    "raht_publish_host": (_i, [_p, _p, _sz, _p]),
    "raht_bounds3_f32": (_i, [_p, _i64, _i64, _p, _p]),
    "raht_voxel_params": (_i, [_p, _i, _i, C.POINTER(C.c_float), _i, _d, _p, _p, _p]),
    "raht_morton_keys_f32": (_i, [_p, _i64, _i64, _p, _i, _p, _p, _p]),
    "raht_morton_keys_i64": (_i, [_p, _i64, _i, _p, _p]),
    "raht_morton_keys_f64": (_i, [_p, _i64, C.POINTER(_d), _d, _i, _p, _p, _p]),
    "raht_morton_decode_f32": (_i, [_p, _i64, _i, _p, _p]),
    "raht_radix_sort_workspace_bytes": (_sz, [_i64, _i, _i]),
    "raht_radix_sort_pairs": (_i, [_p, _p, _p, _p, _i64, _i, _i, _p, _p, _sz, _p]),
    "raht_voxel_segments_workspace_bytes": (_sz, [_i64]),
    "raht_voxel_segments": (_i, [_p, _i64, _p, _p, _p, _p, _p, _sz, _p]),
    "raht_gather_rows_f32": (_i, [_p, _i64, _i64, _p, _p, _p, _p, _i64, _p]),
    "raht_voxel_attr_mean_f32": (_i, [_p, _i64, _i64, _i, _p, _p, _i64, _p, _i64, _i, _p, _i64, _p]),
    "raht_merge_segments": (_i, [_p, _p, _i64, _p, _p, _p, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p, _i, _p]),
    "raht_tree_levels": (_i, [_p, _i64, _p, _p, _p]),
    "raht_tree_level_workspace_bytes": (_sz, [_i64]),
    "raht_tree_level": (_i, [_p, _i64, _i, _p, _p, _p, _p, _p, _sz, _p]),
    "raht_tree_max_level": (_i, [_p, _i64, _p, _p]),
    "raht_order_ragft_workspace_bytes": (_sz, [_i64]),
    "raht_order_ragft": (_i, [_p, _i64, _p, _p, _p, _sz, _p]),
    "raht_plan_workspace_bytes": (_sz, [_i64, _i]),
    "raht_plan_create": (_i, [_p, _i64, _i, _p, _p, _sz, C.POINTER(_p), _p]),
    "raht_plan_destroy": (None, [_p]),
    "raht_plan_info": (_i, [_p, C.POINTER(_i64), _i]),
    "raht_forward": (_i, [_p, _p, _i, _i, _p, _i, _i, _p]),
    "raht_inverse": (_i, [_p, _p, _i, _i, _p, _i, _p]),
    "raht_forward_quant": (_i, [_p, _p, _i, _i, _p, _i, _p, _p, _p, _i64, _p, _i, _p]),
    "raht_dequant_inverse": (_i, [_p, _p, _i64, _p, _p, _i, _p, _i, _p, _i, _p]),
    "raht_quant_reorder": (_i, [_p, _i, _i64, _i, _p, _p, _p, _i64, _p, _i, _p]),
    "raht_dequant_unreorder": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _i, _p]),
    "raht_quantize_f64": (_i, [_p, _i64, _d, _p, _p]),
    "raht_dequantize_f64": (_i, [_p, _i64, _d, _p, _p]),
    "raht_rlgr_encode_workspace_bytes": (_sz, [_i64, _i]),
    "raht_shard_top": (_i, [_p, _p, _p, _i, _p, _i, _i, _i, _p, _p, _p]),
    "raht_rlgr_sync_entries": (_i64, [_i64]),
    "raht_rlgr_encode": (_i, [_p, _i64, _i64, _i, _i, _p, _i64, _p, _p, _p, _p, _sz, _p]),
    "raht_rlgr_encode_stats": (_i, [_p, _i64, _i, C.POINTER(C.c_ulonglong), _p]),
    "raht_rlgr_decode": (_i, [_p, _i64, _p, _p, _i64, _i64, _i, _i, _p, _p]),
    "raht_rlgr_build_index": (_i, [_p, _i64, _p, _i64, _i, _p, _p]),
    "raht_sse_workspace_bytes": (_sz, [_i64, _i]),
    "raht_sse_f64": (_i, [_p, _p, _i64, _i, _i, _i, _p, _p, _sz, _p]),
    "raht_channel_minmax_workspace_bytes": (_sz, [_i64, _i]),
    "raht_channel_minmax": (_i, [_p, _i, _i64, _i, _p, _p, _sz, _p]),
    "raht_channel_minmax_static_max_channels": (_i, []),
    "raht_channel_minmax_static": (_i, [_p, _i, _i64, _i, _p, _p]),
    "raht_sse": (_i, [_p, _i, _p, _i, _i64, _i, _i, _i, _p, _p, _sz, _p]),
    "raht_geometry_symbols": (_i, [_p, _i64, _i64, _p, _p, _p]),
    "raht_geometry_keys_workspace_bytes": (_sz, [_i64]),
    "raht_geometry_keys": (_i, [_p, _i64, _i64, _p, _p, _sz, _p]),
    "raht_streams_pack": (_i, [_p, _i64, _p, _i, _p, _i64, _p, _p, _p]),
    "raht_streams_unpack": (_i, [_p, _p, _i, _p, _i64, _p]),
    "raht_shard_top_cells": (_i, [_p, _p, _p, _i, _p, _i, _i, _i, _p, _p, _p]),
    "raht_ply_pack_vertices": (_i, [_p, _p, _p, _p, _p, _i, _i64, _p, _p]),
    "raht_ply_unpack_vertices": (_i, [_p, _i, _i64, _p, _p, _p, _p]),
    "raht_splat_workspace_bytes": (_sz, [_i64]),
    "raht_splat_project": (_i, [_p, _p, _p, _p, _p, _i, _i, _i64, C.POINTER(C.c_float), C.c_float, C.c_float, C.c_float, C.c_float,
                                _i, _i, _p, _sz, _p, _p]),
    "raht_splat_tile_keys": (_i, [_p, _sz, _i64, _i, _i, _p, _p, _i64, _p]),
    "raht_splat_tile_ranges": (_i, [_p, _i64, _i, _p, _p]),
    "raht_splat_rasterize": (_i, [_p, _sz, _i64, _p, _p, _i, _i, C.POINTER(C.c_float), _p, _p]),
    "raht_rgb_to_yuv": (_i, [_p, _i, _i64, _p, _p]),
    "raht_histogram_i64": (_i, [_p, _i64, _i, _i, _p, _p]),
    "raht_pack_rows": (_i, [_p, _i64, _p, _p, _p, _p, _p, _i, _p, _p]),
    "raht_unpack_rows": (_i, [_p, _i64, _i, _p, _p, _p, _p, _p, _p]),
    "raht_scene_front_workspace_bytes": (_sz, [_i64, _i]),
    "raht_scene_front": (_i, [_p, _i64, _i, _i, C.POINTER(C.c_float), _d, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p, _p]),
    "raht_scene_tree_workspace_bytes": (_sz, [_i64, _i]),
    "raht_scene_encode_workspace_bytes": (_sz, [_i64, _i, _i, _i]),
    "raht_scene_encode": (_i, [_p, _p, _i64, _p, _p, _p, _p, _p, _i, _p, C.POINTER(_d), _i, _i, _p, _p, _i64, _p, _p, _p, _p,
                               C.POINTER(_p), C.POINTER(_i64), _p, _sz, _p, _sz, _p, _p]),
    "raht_scene_decode_workspace_bytes": (_sz, [_i64, _i, _i]),
    "raht_scene_decode": (_i, [_p, _i64, _i, _p, _i64, _p, _p, _i, C.POINTER(_d), _i, _i, _p, _p, _p, _p, C.POINTER(_p),
                               C.POINTER(_i64), _p, _sz, _p, _sz, _p, _p]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _f = getattr(lib, _name)          # AttributeError here = header and library out of sync
    _f.restype = _res
    _f.argtypes = _args


class RahtError(RuntimeError):
    pass


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib.raht_last_error().decode("utf-8", "replace")
        raise RahtError(f"{what or 'libraht_b200'} failed ({rc}): {msg}")


def ptr(t):
    """Device pointer of a tensor (None -> NULL).  The library launches on the CURRENT device's current stream (stream()
    below; it never calls cudaSetDevice), so a tensor of another GPU is refused here instead of faulting in a kernel:
    multi-GPU processes select the device first (`torch.cuda.set_device` / `with torch.cuda.device(d)`)."""
    if t is None:
        return None
    if t.is_cuda and t.device.index != torch.cuda.current_device():
        raise RuntimeError(f"tensor on {t.device} but the current CUDA device is cuda:{torch.cuda.current_device()}: "
                           "select the device (torch.cuda.set_device / torch.cuda.device) before calling raht_b200")
    return C.c_void_p(t.data_ptr())


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(t: torch.Tensor, name: str):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (raht-3dgs-codec_b200 has no CPU path)")


def workspace(nbytes: int, device) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


_WS_CACHE = {}


def workspace_cached(tag: str, nbytes: int, device) -> torch.Tensor:
    """Scratch memory that lives across calls, one buffer per (purpose, device, stream): the big workspaces (the RLGR
    encoder's is 1.3 GB for a 1 M x 59 scene) are not handed back to the caching allocator between calls, where another
    request may split them and the next encode pays a cudaMalloc in the middle of its stage.  Stream-ordered use only:
    the buffer belongs to the stream that is current when it is requested."""
    dev = torch.device(device)
    key = (tag, dev.index if dev.index is not None else torch.cuda.current_device(), torch.cuda.current_stream(dev).cuda_stream)
    buf = _WS_CACHE.get(key)
    need = max(int(nbytes), 256)
    if buf is None or buf.numel() < need:
        _WS_CACHE.pop(key, None)
        buf = torch.empty(need, dtype=torch.uint8, device=dev)
        _WS_CACHE[key] = buf
    return buf


def release_workspaces():
    """Drop the cached workspaces (they are otherwise kept for the life of the process)."""
    _WS_CACHE.clear()


def dtype_tag(dt: torch.dtype) -> int:
    if dt == torch.float64:
        return RAHT_F64
    if dt == torch.float32:
        return RAHT_F32
    raise TypeError(f"unsupported dtype {dt}")


_SCRATCH = {}


def read_host(*tensors):
    """Small device tensors -> numpy copies on the host with ONE synchronisation of the current stream and no copy
    engine (raht_publish_host): the sizes and status words the host has to wait for."""
    import numpy as np
    dev = tensors[0].device
    need = sum((t.numel() * t.element_size() + 15) // 16 * 16 for t in tensors)
    buf = _SCRATCH.get(dev)
    if buf is None or buf.numel() < need:
        buf = torch.empty(max(need, 1 << 16), dtype=torch.uint8).pin_memory()
        _SCRATCH[dev] = buf
    st, o, views = stream(), 0, []
    for t in tensors:
        t = t.contiguous()
        nb = t.numel() * t.element_size()
        check(lib.raht_publish_host(ptr(t), C.c_void_p(buf.data_ptr() + o), nb, st), "raht_publish_host")
        views.append((o, nb, t.dtype, tuple(t.shape)))
        o += (nb + 15) // 16 * 16
    torch.cuda.current_stream(dev).synchronize()
    out = []
    for o, nb, dt, shape in views:
        out.append(buf[o:o + nb].view(dt).reshape(shape).numpy().copy())
    return out

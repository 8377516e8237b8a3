"""End-to-end RAHT-3DGS encode / decode on one GPU: the call sequence of the reference drivers
(python/test_voxelize_3dgs.py:198-265 for voxelise+merge, python/encode_3dgs.py:144-276 for
RAHT_param -> RAHT -> quantise -> reorder -> RLGR and back) with every step on the device.

Channel layout of the coded matrix (D = 11 + color_dim, 59 for 3DGS with 48 SH coefficients):
[merged_means 3 | quats 4 | scales 3 | opacity 1 | colors].
"""
import os
from dataclasses import dataclass, field
from typing import Optional

import torch

from . import ops
from . import _lib as L
from ._lib import lib, check, ptr, stream


@dataclass
class EncodedScene:
    J: int
    n_in: int
    nvox: int
    D: int
    step: object                      # scalar or per-channel steps
    vmin: torch.Tensor                # (3,) fp32
    voxel_size: float
    width: float
    keys: torch.Tensor                # (nvox,) int64 Morton codes of the occupied voxels (geometry side info)
    streams: ops.RlgrStreams          # D closed RLGR streams
    transform_dtype: torch.dtype = torch.float64
    tree: Optional[ops.RahtTree] = field(default=None, repr=False)   # encoder-side cache (decoder rebuilds from keys)
    packed: Optional[torch.Tensor] = field(default=None, repr=False)  # merged attributes (nvox, D) fp32 when keep_packed

    def payload_bytes(self) -> int:
        """Bytes of the D reference-format RLGR streams."""
        return self.streams.total_bytes()

    def index_bytes(self) -> int:
        """Bytes of the optional sync index (side information for chunk-parallel decoding)."""
        return self.streams.index_bytes()


# The tree / plan kernels (RAHT_param) are chains of small, latency-bound launches that need nothing but the voxel keys:
# they run on a side stream next to the stage that is independent of them -- the segment merge in the encoder, the RLGR
# decoder in the decoder -- and the main stream waits for them before the first butterfly.  OFF by default (RAHT_OVERLAP=1
# turns it on): measured on the 1 M bench scene it LOSES -- the plan's small launches, interleaved with the co-running
# kernel's CTAs, stretch it by more than they take alone (merge 0.31 -> 0.43 ms, RLGR decode 0.57 -> 1.00 ms; step
# 4.09 -> 4.34 ms).  Kept, with its parity test, for scenes where the co-running stage is longer.
OVERLAP = os.environ.get("RAHT_OVERLAP") is not None
_SIDE = {}


def _side_stream(device) -> "torch.cuda.Stream":
    dev = torch.device(device)
    key = dev.index if dev.index is not None else torch.cuda.current_device()
    s = _SIDE.get(key)
    if s is None:
        s = _SIDE[key] = torch.cuda.Stream(device=dev)
    return s


def _build_tree(keys, status, tile_rows, want_position):
    tree = ops.tree_from_keys(keys, validate=True, status=status)
    tree.order(want_position=want_position)
    tree.plan(tile_rows)
    return tree


def start_tree(keys, status, tile_rows=None, want_position=True):
    """Queue tree_from_keys + order_RAGFT + the transform plan on the side stream (after everything queued on the current
    stream so far).  Returns (tree, side stream): `torch.cuda.current_stream().wait_stream(side)` before the tree is used.
    Memory: what the side stream allocates is consumed on the main stream and freed after it; the next side-stream job
    starts with a wait for the main stream, so the caching allocator's per-stream reuse stays ordered."""
    side = _side_stream(keys.device)
    side.wait_stream(torch.cuda.current_stream(keys.device))
    with torch.cuda.stream(side):
        tree = _build_tree(keys, status, tile_rows, want_position)
    return tree, side


class StageTimer:
    """CUDA-event timer per pipeline stage (events on the current stream)."""

    def __init__(self, enabled=False):
        self.enabled = enabled
        self.marks = []

    def mark(self, name):
        if self.enabled:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.marks.append((name, e))

    def handles(self, names):
        """Event handles for the stages of a scene-level C call (raht_scene_*: `events`), or None when timing is off.
        Returns (ctypes array of cudaEvent_t, commit): commit() appends the events to the marks once the call is queued."""
        if not self.enabled:
            return None, (lambda: None)
        import ctypes as C
        evs = []
        for _ in names:
            e = torch.cuda.Event(enable_timing=True)
            e.record()                                    # (a torch event gets its handle on its first record)
            evs.append(e)
        arr = (C.c_void_p * len(names))(*[C.c_void_p(e.cuda_event) for e in evs])
        return arr, (lambda: self.marks.extend(zip(names, evs)))

    def result(self):
        if not self.enabled or len(self.marks) < 2:
            return {}
        torch.cuda.synchronize()
        out = {}
        for (_, e0), (name, e1) in zip(self.marks[:-1], self.marks[1:]):
            out[name] = out.get(name, 0.0) + e0.elapsed_time(e1)
        return out


def _voxel_params(means, J, given=None):
    """params4 = (vmin xyz, voxel_size) fp32 and params_d2 = (width, voxel_size) fp64 on the device, from the data
    (voxelize_pc.py:87-97) or from a given (vmin (3,) tensor, width float) -- the `vmin`, `width` arguments of
    voxelize_pc_batched, used by the sharded encoder for the bounding box of the whole scene."""
    import ctypes as C
    dev = means.device
    st = stream()
    bounds = torch.empty(6, dtype=torch.float32, device=dev)
    params4 = torch.empty(4, dtype=torch.float32, device=dev)
    params_d2 = torch.empty(2, dtype=torch.float64, device=dev)
    if given is None:
        check(lib.raht_bounds3_f32(ptr(means), means.shape[0], 3, ptr(bounds), st), "raht_bounds3_f32")
        check(lib.raht_voxel_params(ptr(bounds), J, 0, None, 0, 0.0, ptr(params4), ptr(params_d2), st), "raht_voxel_params")
    else:
        vmin, width = given
        vm = (C.c_float * 3)(*[float(x) for x in vmin.detach().cpu().tolist()])
        bounds.zero_()
        check(lib.raht_voxel_params(ptr(bounds), J, 1, vm, 1, float(width), ptr(params4), ptr(params_d2), st), "raht_voxel_params")
    return params4, params_d2


@torch.no_grad()
def morton_keys_given(means, J, vmin, width):
    """Morton codes of the Gaussian centres for a given bounding box (no sort)."""
    means = means.contiguous().float()
    params4, _ = _voxel_params(means, J, (vmin, width))
    keys = torch.empty(means.shape[0], dtype=torch.int64, device=means.device)
    check(lib.raht_morton_keys_f32(ptr(means), means.shape[0], 3, ptr(params4), J, ptr(keys), None, stream()), "raht_morton_keys_f32")
    return keys


@torch.no_grad()
def voxelize_merge(means, quats, scales, opacities, colors, J, packed_dtype=torch.float32, timer=None, status=None, given=None,
                   on_keys=None):
    """Voxelise the Gaussian centres and merge the Gaussians of each voxel.
    Returns (packed (nvox, D), unique_keys int64 (nvox), info dict).  One synchronisation (the voxel
    count must reach the host); with a shared `status` word the sort status is left to the caller.
    given = (vmin, width): quantise against this bounding box instead of the data's."""
    timer = timer or StageTimer(False)
    dev = means.device
    N = means.shape[0]
    st = stream()
    means = means.contiguous().float()
    params4, params_d2 = _voxel_params(means, J, given)
    keys = torch.empty(N, dtype=torch.int64, device=dev)
    check(lib.raht_morton_keys_f32(ptr(means), N, 3, ptr(params4), J, ptr(keys), None, st), "raht_morton_keys_f32")
    timer.mark("morton")
    deferred = status is not None
    keys_sorted, idx32, status = ops.radix_sort_pairs(keys, 3 * J, status=status)
    timer.mark("sort")
    _, offsets, unique_keys, nvox = ops.voxel_segments(keys_sorted)
    timer.mark("segments")
    if not deferred:
        ops.raise_for_status(int(L.read_host(status)[0][0]))
    if on_keys is not None:                                        # the caller's work that needs the voxel keys only (side stream)
        on_keys(unique_keys)
    _, packed = ops.merge_segments(idx32, offsets, means, quats.contiguous().float(), scales.contiguous().float(),
                                   opacities.contiguous().float(), colors.contiguous().float(), True, separate=False,
                                   packed_dtype=packed_dtype)
    timer.mark("merge")
    info = {"Nvox": nvox, "N": N, "vmin": params4[:3], "params_d2": params_d2, "sort_idx": idx32}   # params_d2 = (width, voxel_size), device
    return packed, unique_keys, info


# Visual-importance weights of the attribute groups (encode_3dgs_debug.py:336-341: 1 / rendering PSNR of the ablation of
# each group) and the levels budget (:346).  The reference codes 56 channels (no positions); with the merged means in
# front (D = 59) they get the weight of the quaternions (a choice of this build: the reference does not code them).
IMPORTANCE = {"means": 1.0 / 21.93, "quats": 1.0 / 21.93, "scales": 1.0 / 26.36, "opacity": 1.0 / 42.22, "colors": 1.0 / 38.67}
LEVELS_BUDGET = 1024


def attribute_ranges(D):
    """Channel ranges of the attribute groups (encode_3dgs_debug.py:327-332) for the 56-channel reference layout, or the
    59-channel layout of this build (merged means in front)."""
    o = 3 if D >= 59 else 0
    r = {"means": (0, 3)} if o else {}
    r.update({"quats": (o, o + 4), "scales": (o + 4, o + 7), "opacity": (o + 7, o + 8), "colors": (o + 8, D)})
    return r


@torch.no_grad()
def per_attribute_steps(T, ranges=None, weights=None, budget=LEVELS_BUDGET):
    """Quantisation step per channel from the coefficient range of its attribute group and the group's share of a levels
    budget (encode_3dgs_debug.py:343-366): levels = max(int(budget * w / sum(w)), 2), step = max(range / max(levels - 1, 1),
    1e-6), range = max - min over the group's channels.  Returns (list of D steps, {group: {step, levels, range, channels}}).
    The range is reduced on the device (raht_channel_minmax); the rule is host arithmetic, as in the reference."""
    n, D = T.shape
    ranges = attribute_ranges(D) if ranges is None else ranges
    weights = {g: IMPORTANCE[g] for g in ranges} if weights is None else weights
    lo, hi = ops.channel_minmax(T)
    lo, hi = L.read_host(lo, hi)
    total = sum(weights[g] for g in weights)              # (the reference sums every weight it lists, :344)
    steps, info = [0.0] * D, {}
    for g, (a, b) in ranges.items():
        if a >= D:
            continue
        b = min(b, D)
        rng = float(hi[a:b].max() - lo[a:b].min())
        levels = max(int(budget * weights[g] / total), 2)
        step = max(rng / max(levels - 1, 1), 1e-6)
        info[g] = {"step": step, "levels": levels, "range": rng, "channels": (a, b)}
        for c in range(a, b):
            steps[c] = step
    if any(s == 0.0 for s in steps):
        raise ValueError("per_attribute_steps: the attribute ranges do not cover every channel")
    return steps, info


@torch.no_grad()
def encode_attributes(keys, Cmat, step, transform_dtype=torch.float64, tile_rows=None, timer=None, flags=0, status=None, fused=False,
                      tree_job=None):
    """RAHT_param -> RAHT -> quantise -> reorder -> RLGR for rows that are Morton-sorted unique voxels.
    fused: transform and quantisation in one pass (ops.forward_quant; same symbols, bit for bit), else the two
    reference steps one after the other.  Off by default -- measured, it does not pay on either kind of scene: on the
    sparse 3DGS scenes (1 M bench scene: consecutive voxels fall into different level buckets of order_RAGFT) the fused pass
    scatters 4-byte symbols from the transform tiles where the reorder kernel moves whole rows (forward -0.04 ms, inverse
    +0.05 ms per step), and on the dense point cloud of encode_ply.py (777 K voxels, 11 quantisation steps per frame,
    bench.py --workload c1) the two-pass path transforms ONCE and quantises eleven times, while the fused one repeats the
    transform per step: 10.14 ms against 10.40 ms per frame.
    Device-side validation (sorted/unique keys, int32 overflow, sort stall) accumulates in one status
    word that is read once, after the entropy coder."""
    timer = timer or StageTimer(False)
    status = ops._status_word(keys.device) if status is None else status
    if isinstance(step, str):
        if step != "per_attribute":
            raise ValueError(f"unknown step rule {step!r}")
        fused = False                                              # the rule needs the coefficients before they are quantised
    if tree_job is not None:                                       # (tree, side stream) from start_tree(): already queued
        tree, side = tree_job
        torch.cuda.current_stream(keys.device).wait_stream(side)
    else:
        tree = _build_tree(keys, status, tile_rows, True)
    order, _ = tree.order(want_position=True)
    timer.mark("raht_param")
    if fused:
        q = ops.forward_quant(tree, Cmat, step, transform_dtype, tile_rows, flags, status)
        timer.mark("raht_quant")
    else:
        T = ops.forward(tree, Cmat, transform_dtype, tile_rows, flags)
        timer.mark("raht")
        if isinstance(step, str):                                  # "per_attribute": steps from the coefficient ranges
            step, _ = per_attribute_steps(T)
            tree.steps = step
        q = ops.quant_reorder(T, order, step, status=status)
        timer.mark("quant_reorder")
    streams = ops.rlgr_encode(q, 1, status=status)
    timer.mark("rlgr_encode")
    return tree, streams


@torch.no_grad()
def decode_attributes(keys, streams, D, step, transform_dtype=torch.float64, tree=None, tile_rows=None, timer=None, flags=0, fused=False):
    """RLGR^-1 -> dequantise -> inverse reorder -> iRAHT (fused: the last three in one pass, ops.dequant_inverse)."""
    timer = timer or StageTimer(False)
    status = None
    side = None
    if tree is None:
        status = ops._status_word(keys.device)
        if OVERLAP:
            tree, side = start_tree(keys, status, tile_rows, want_position=fused)
        else:
            tree = _build_tree(keys, status, tile_rows, fused)
            timer.mark("raht_param_dec")
    q = ops.rlgr_decode(streams, 1)
    timer.mark("rlgr_decode")
    if side is not None:
        torch.cuda.current_stream(keys.device).wait_stream(side)
        timer.mark("raht_param_dec")                               # what is left of the tree / plan after the overlap
    order = tree.order()
    if fused:
        Crec = ops.dequant_inverse(tree, q, step, transform_dtype, tile_rows, flags)
        timer.mark("dequant_iraht")
    else:
        T = ops.dequant_unreorder(q, order, step, transform_dtype)
        timer.mark("dequant_unreorder")
        Crec = ops.inverse(tree, T, tile_rows, flags)
        timer.mark("iraht")
    if status is not None:
        ops.raise_for_status(int(L.read_host(status)[0][0]))      # the decoder's one synchronisation: were the keys a valid voxel list?
    return Crec


# ---- scene-level calls (csrc/scene.cu): one leg of the pipeline per host call -------------------------------------------
_PINNED = {}


_PIN_RING = 16


def _pinned_results(device, nbytes):
    """A small pinned host buffer for the results a scene-level call publishes (voxel count, stream lengths, status words).
    A ring of _PIN_RING buffers per device and stream: a deferred check (`checks=` of encode_scene / decode_scene) reads
    its buffer after later calls have been issued, so consecutive calls must not share one."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), torch.cuda.current_stream(device).cuda_stream)
    ring = _PINNED.get(key)
    if ring is None or ring[0][0].numel() < nbytes:
        ring = _PINNED[key] = ([torch.empty(max(nbytes, 4096), dtype=torch.uint8).pin_memory() for _ in range(_PIN_RING)], [0])
    bufs, nxt = ring
    buf = bufs[nxt[0] % _PIN_RING]
    nxt[0] += 1
    return buf


def _steps_host(step, D):
    import ctypes as C
    if isinstance(step, torch.Tensor):
        step = step.detach().cpu().reshape(-1).tolist()
    if isinstance(step, (int, float)):
        vals = [float(step)] * D
    else:
        vals = [float(x) for x in step]
        if len(vals) == 1:
            vals = vals * D
    if len(vals) != D or any(not (v > 0.0) for v in vals):
        raise ValueError(f"need one positive quantisation step, or {D}")
    return (C.c_double * D)(*vals)


def _tree_from_workspace(tree_ws, layout, nvox, tile_rows, plan_handle):
    """ops.RahtTree over the tree workspace a scene-level call filled (no copies)."""
    lv = tree_ws[layout[0]:layout[0] + nvox]
    tree = ops.RahtTree(lv, nvox)
    tree._order = tree_ws[layout[1]:layout[1] + 8 * nvox].view(torch.int64)
    tree._position = tree_ws[layout[2]:layout[2] + 4 * nvox].view(torch.int32)
    tree._w = tree_ws[layout[3]:layout[3] + 8 * nvox].view(torch.float64)
    tree._plans[tile_rows] = ops._PlanHandle(plan_handle, tree_ws)
    return tree


@torch.no_grad()
def _encode_scene_calls(means, quats, scales, opacities, colors, J, step, transform_dtype, tile_rows, timer, keep_packed, given=None,
                        checks=None, min_slot=0):
    """encode_scene through raht_scene_front + raht_scene_encode: two host calls and one synchronisation between them (the
    voxel count sizes the second leg), one more at the end (stream lengths and status word)."""
    import ctypes as C
    dev = means.device
    N = means.shape[0]
    st = stream()
    tile_rows = tile_rows or ops.DEFAULT_TILE_ROWS
    means, quats, scales = means.contiguous().float(), quats.contiguous().float(), scales.contiguous().float()
    opacities, colors = opacities.contiguous().float(), colors.contiguous().float()
    cd = colors.shape[1]
    D = 11 + cd
    steps = _steps_host(step, D)
    tag = L.dtype_tag(transform_dtype)
    # outputs of the front that outlive the call, carved out of one allocation
    o_keys, o_idx = 256, 256 + (N * 8 + 255) // 256 * 256
    o_off = o_idx + (N * 4 + 255) // 256 * 256
    front = torch.empty(o_off + (N + 1) * 4 + 256, dtype=torch.uint8, device=dev)
    params4 = front[0:16].view(torch.float32)
    params_d2 = front[16:32].view(torch.float64)
    nvox_dev = front[32:40].view(torch.int64)
    status = front[40:44].view(torch.int32)
    front[32:48].zero_()
    keys_all = front[o_keys:o_keys + N * 8].view(torch.int64)
    idx32 = front[o_idx:o_idx + N * 4].view(torch.int32)
    offsets = front[o_off:o_off + (N + 1) * 4].view(torch.int32)
    ws_bytes = lib.raht_scene_front_workspace_bytes(N, J)
    ws = L.workspace_cached("scene_front", ws_bytes, dev)
    pin = _pinned_results(dev, 64 + 8 * D + 64)
    ev, commit = timer.handles(("morton", "sort", "segments"))
    if given is None:
        use_given, vm, width = 0, None, 0.0
    else:
        use_given, vm, width = 1, (C.c_float * 3)(*[float(x) for x in given[0].detach().cpu().tolist()]), float(given[1])
    check(lib.raht_scene_front(ptr(means), N, J, use_given, vm, width, ptr(params4), ptr(params_d2), ptr(keys_all), ptr(idx32),
                               ptr(offsets), ptr(nvox_dev), C.c_void_p(pin.data_ptr()), ptr(status), ptr(ws), ws.numel(), ev, st),
          "raht_scene_front")
    commit()
    # Everything the second leg needs is allocated while the front runs, sized for nvox <= N: between the synchronisation
    # below and the first kernel of the second leg the GPU idles, so that stretch holds nothing but the one host call.
    slot = max((N * 4 + 64 + 7) // 8 * 8, int(min_slot))
    buf = torch.empty(D * slot, dtype=torch.uint8, device=dev)
    nbytes = torch.empty(D, dtype=torch.int64, device=dev)
    index_all = torch.empty(D * int(lib.raht_rlgr_sync_entries(N)), 2, dtype=torch.int32, device=dev)
    packed_all = torch.empty(N, D, dtype=torch.float32, device=dev) if keep_packed else None
    tree_ws = torch.empty(lib.raht_scene_tree_workspace_bytes(N, tile_rows), dtype=torch.uint8, device=dev)
    enc_ws = L.workspace_cached("scene_encode", lib.raht_scene_encode_workspace_bytes(N, D, tag, 0 if keep_packed else 1), dev)
    layout = (C.c_int64 * 11)()
    plan = C.c_void_p()
    ev, commit = timer.handles(("merge", "raht_param", "raht", "quant_reorder", "rlgr_encode"))
    pin_enc = C.c_void_p(pin.data_ptr() + 64)
    a = (ptr(idx32), ptr(offsets), ptr(means), ptr(quats), ptr(scales), ptr(opacities), ptr(colors), ptr(keys_all), ptr(packed_all),
         ptr(buf), ptr(nbytes), ptr(index_all), ptr(status), ptr(tree_ws), tree_ws.numel(), ptr(enc_ws), enc_ws.numel())
    torch.cuda.current_stream(dev).synchronize()                      # the voxel count sizes everything that follows
    nvox = int(pin[:8].view(torch.int64)[0])
    check(lib.raht_scene_encode(a[0], a[1], nvox, a[2], a[3], a[4], a[5], a[6], cd, a[7], steps, tag, tile_rows, a[8], a[9], slot, a[10],
                                a[11], a[12], pin_enc, C.byref(plan), layout, a[13], a[14], a[15], a[16], ev, st), "raht_scene_encode")
    commit()
    head = pin[:24].numpy()
    width_h, voxel_size_h = (float(x) for x in head[8:24].view("float64"))
    keys = keys_all[:nvox]
    entries = int(lib.raht_rlgr_sync_entries(nvox))
    index = index_all[:D * entries]
    packed = packed_all[:nvox] if keep_packed else None
    tree = _tree_from_workspace(tree_ws, list(layout), nvox, tile_rows, plan)
    streams = ops.RlgrStreams(buf, slot, nbytes, nvox, D, index)
    enc = EncodedScene(J=J, n_in=N, nvox=nvox, D=D, step=step, vmin=params4[:3], voxel_size=voxel_size_h, width=width_h, keys=keys,
                       streams=streams, transform_dtype=transform_dtype, tree=tree, packed=packed)
    done = torch.cuda.Event()
    done.record(torch.cuda.current_stream(dev))

    def finish():
        """stream lengths and the status word of this encode (published to `pin` by the call); a stream that outgrew its
        slot (more than 4 bytes per symbol) is coded again into wider slots"""
        done.synchronize()
        res = pin[64:64 + 8 * D + 4].numpy()
        sizes = [int(x) for x in res[:8 * D].view("int64")]
        bits = int(res[8 * D:8 * D + 4].view("int32")[0])
        ops.raise_for_status(bits)
        if bits & L.ST_STREAM_OVERFLOW:
            again = _encode_scene_calls(means, quats, scales, opacities, colors, J, step, transform_dtype, tile_rows, StageTimer(False),
                                        keep_packed, given, None, (max(sizes) + 64 + 7) // 8 * 8)
            enc.streams, enc.tree, enc.packed = again.streams, again.tree, again.packed
        else:
            streams._sizes = sizes

    if checks is None:
        finish()
    else:
        checks.append(finish)
    return enc


@torch.no_grad()
def _decode_scene_calls(enc, tile_rows, timer, checks=None):
    """decode_scene through raht_scene_decode: one host call; the tree is rebuilt from the voxel keys."""
    import ctypes as C
    dev = enc.keys.device
    nvox, D = enc.nvox, enc.D
    tile_rows = tile_rows or ops.DEFAULT_TILE_ROWS
    tag = L.dtype_tag(enc.transform_dtype)
    steps = _steps_host(enc.step, D)
    S = enc.streams
    Crec = torch.empty(nvox, D, dtype=enc.transform_dtype, device=dev)
    coords = torch.empty(nvox, 3, dtype=torch.float32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    tree_ws = L.workspace_cached("scene_decode_tree", lib.raht_scene_tree_workspace_bytes(nvox, tile_rows), dev)
    ws = L.workspace_cached("scene_decode", lib.raht_scene_decode_workspace_bytes(nvox, D, tag), dev)
    pin = _pinned_results(dev, 64 + 8 * D + 64)
    plan = C.c_void_p()
    ev, commit = timer.handles(("raht_param_dec", "rlgr_decode", "dequant_unreorder", "iraht", "geometry"))
    check(lib.raht_scene_decode(ptr(enc.keys), nvox, enc.J, ptr(S.buf), S.slot_bytes, ptr(S.nbytes), ptr(S.sync_index), D, steps, tag,
                                tile_rows, ptr(Crec), ptr(coords), ptr(status), C.c_void_p(pin.data_ptr() + 64 + 8 * D + 8), C.byref(plan),
                                None, ptr(tree_ws), tree_ws.numel(), ptr(ws), ws.numel(), ev, stream()), "raht_scene_decode")
    commit()
    lib.raht_plan_destroy(plan)                                       # (host handle only; the kernels are queued with copies of it)
    done = torch.cuda.Event()
    done.record(torch.cuda.current_stream(dev))

    def finish():                                                     # the decoder's one synchronisation: were the keys a valid voxel list?
        done.synchronize()
        ops.raise_for_status(int(pin[64 + 8 * D + 8:64 + 8 * D + 12].numpy().view("int32")[0]))

    if checks is None:
        finish()
    else:
        checks.append(finish)
    return coords, Crec


# RAHT_SCENE_CALLS=0 (or codec.SCENE_CALLS = False): encode_scene / decode_scene go through the per-operator entry points,
# one host call per kernel, as in round 1 -- kept as the reference point of tests/test_gpu_parity.py
# (test_scene_level_calls_match_the_operator_path) and for timing comparisons.
SCENE_CALLS = os.environ.get("RAHT_SCENE_CALLS", "1") != "0"


def _plain_step(step):
    return SCENE_CALLS and not isinstance(step, str)


@torch.no_grad()
def encode_scene(means, quats, scales, opacities, colors, J, step, transform_dtype=torch.float64, tile_rows=None,
                 timer=None, flags=0, keep_packed=False, fused=False, checks=None) -> EncodedScene:
    """checks: None -- the call returns when its status word and stream lengths have been read and checked (one more
    synchronisation).  A list -- that last step is appended to it as a callable instead and the call returns with its
    kernels queued: the caller runs the callables later (e.g. one scene behind), which raises what the call would have
    raised and completes `streams.sizes()`; until then the GPU never waits for the host between encode and decode."""
    timer = timer or StageTimer(False)
    timer.mark("start")
    if not fused and not OVERLAP and flags == 0 and _plain_step(step) and means.shape[0] > 0:
        return _encode_scene_calls(means, quats, scales, opacities, colors, J, step, transform_dtype, tile_rows, timer, keep_packed,
                                   checks=checks)
    status = ops._status_word(means.device)
    job = []
    on_keys = (lambda k: job.append(start_tree(k, status, tile_rows))) if OVERLAP else None
    packed, keys, info = voxelize_merge(means, quats, scales, opacities, colors, J, torch.float32, timer, status, on_keys=on_keys)
    tree, streams = encode_attributes(keys, packed, step, transform_dtype, tile_rows, timer, flags, status, fused,
                                      tree_job=job[0] if job else None)
    width, voxel_size = L.read_host(info["params_d2"])[0].tolist()          # after the coder's synchronisation: no extra wait
    step = getattr(tree, "steps", step)                            # "per_attribute": the steps the rule picked
    return EncodedScene(J=J, n_in=info["N"], nvox=info["Nvox"], D=packed.shape[1], step=step, vmin=info["vmin"],
                        voxel_size=voxel_size, width=width, keys=keys, streams=streams,
                        transform_dtype=transform_dtype, tree=tree, packed=packed if keep_packed else None)


@torch.no_grad()
def decode_scene(enc: EncodedScene, reuse_tree=False, tile_rows=None, timer=None, flags=0, fused=False, checks=None):
    """Returns (voxel coordinates (nvox,3) fp32 integer-valued, attributes (nvox, D)).  checks: as in encode_scene."""
    timer = timer or StageTimer(False)
    timer.mark("start_dec")
    if (SCENE_CALLS and not fused and not OVERLAP and flags == 0 and not reuse_tree and enc.nvox > 0 and enc.streams.nch == enc.D
            and enc.streams.n == enc.nvox):
        return _decode_scene_calls(enc, tile_rows, timer, checks)
    Crec = decode_attributes(enc.keys, enc.streams, enc.D, enc.step, enc.transform_dtype,
                             enc.tree if reuse_tree else None, tile_rows, timer, flags, fused)
    coords = torch.empty(enc.nvox, 3, dtype=torch.float32, device=enc.keys.device)
    check(lib.raht_morton_decode_f32(ptr(enc.keys), enc.nvox, enc.J, ptr(coords), stream()), "raht_morton_decode_f32")
    timer.mark("geometry")
    return coords, Crec


PLY_COLOR_STEPS = [1, 2, 4, 6, 8, 12, 16, 20, 24, 32, 64]        # encode_ply.py:29


@torch.no_grad()
def encode_ply_frame(V, Crgb, J, steps=None, fused=False, timer=None, keep_reconstruction=False):
    """One frame of the reference's point-cloud driver (python/encode_ply.py:104-214; BASELINE configs[0]):
    colours -> YUV (utils.rgb_to_yuv), RAHT_param + RAHT once, then for every quantisation step of the list
    (encode_ply.py:29) quantise, PSNR of the luma coefficients (:151-152), reorder, RLGR of the three channels, and the
    decoder: RLGR^-1, dequantise, inverse reorder, iRAHT.  V: (N, 3) Morton-sorted unique voxel coordinates (what
    read_ply_file returns for a voxelised cloud), Crgb: (N, 3) in [0, 255].
    fused: transform + quantisation in one pass per step (ops.forward_quant) instead of one transform and a quantisation
    per step.  Returns (per-step list of {step, bytes, bpp, psnr_y[, C_rec]}, Cyuv)."""
    import math
    import utils as shim_utils
    timer = timer or StageTimer(False)
    steps = PLY_COLOR_STEPS if steps is None else steps
    dev = V.device
    timer.mark("start")
    C = shim_utils.rgb_to_yuv(Crgb)                                             # encode_ply.py:108
    timer.mark("rgb_to_yuv")
    status = ops._status_word(dev)
    keys, status = ops.morton_keys_f64(V, (0.0, 0.0, 0.0), float(2 ** J), J)    # :113-115 origin 0, width 2^J
    tree = _build_tree(keys, status, None, True)
    order, _ = tree.order(want_position=True)
    timer.mark("raht_param")
    N = C.shape[0]
    T = None if fused else ops.forward(tree, C, torch.float64)                  # :121
    timer.mark("raht")
    out = []
    for step in steps:
        q = ops.forward_quant(tree, C, step, torch.float64, status=status) if fused else ops.quant_reorder(T, order, step, status=status)
        timer.mark("quant_reorder")
        streams = ops.rlgr_encode(q, 1, status=status)                          # :158-176 (reads the status word: the step's one sync)
        timer.mark("rlgr_encode")
        qd = ops.rlgr_decode(streams, 1)                                        # :179-189
        timer.mark("rlgr_decode")
        Td = ops.dequant_unreorder(qd, order, step, torch.float64)              # :199-206
        timer.mark("dequant_unreorder")
        rec = {"step": step, "bytes": streams.total_bytes(), "bpp": streams.total_bytes() * 8.0 / max(N, 1)}
        # :149-152  MSE of the quantised luma COEFFICIENTS over 255^2 (orthonormal transform: equals the luma MSE)
        y = Td[:, 0] if fused else None
        ref_y = T[:, 0] if T is not None else None
        if ref_y is not None:
            mse = float(L.read_host(ops.sse(ref_y.reshape(-1, 1).contiguous(), Td[:, :1].contiguous()))[0][0]) / (N * 255.0 ** 2)
            rec["psnr_y"] = -10.0 * math.log10(mse) if mse > 0 else float("inf")
        Crec = ops.inverse(tree, Td)                                            # :207
        timer.mark("iraht")
        if keep_reconstruction:
            rec["C_rec"] = Crec
            rec["streams"] = streams
        del y
        out.append(rec)
    return out, C


def psnr(Cref: torch.Tensor, Crec: torch.Tensor, c0=0, c1=None) -> float:
    """-10 log10(mean((C - C_rec)^2) + 1e-10)   (encode_3dgs.py:298-299)."""
    import math
    c1 = Cref.shape[1] if c1 is None else c1
    s = float(L.read_host(ops.sse(Cref, Crec, c0, c1))[0][0])
    return -10.0 * math.log10(s / (Cref.shape[0] * (c1 - c0)) + 1e-10)


class ScenePipeline:
    """Streams scenes from pinned HOST buffers through encode + decode with the copies off the compute stream
    ("overlap H2D of scene k+1 with compute of scene k on a second stream"): the H2D of the next
    scene runs under the current scene's kernels and the D2H of the current results under the next scene's, so a
    stream of scenes moves at max(H2D, compute, D2H) per scene instead of their sum (PCIe is full duplex).

        pipe = ScenePipeline(J, step)
        pipe.submit(host_scene)            # pinned (means, quats, scales, opacities, colors); starts its H2D
        for next_scene in scenes:
            pipe.submit(next_scene)        # prefetch
            out = pipe.run()               # encode + decode of the oldest submitted scene; D2H queued on the copy-out stream
        out = pipe.run(); pipe.finish()    # drain

    `run()` returns a dict of pinned host tensors (streams end to end, each padded to 8 bytes, their sizes, sync index, keys, and -- result="attributes" -- the
    decoded attributes, or -- result="psnr", what the reference's rate-distortion loop reads back,
    encode_3dgs.py:298-310 -- the sum of squared errors of the reconstruction) that are valid after `finish()` (or
    `out["done"].synchronize()`)."""

    def __init__(self, J, step, transform_dtype=torch.float64, tile_rows=None, device=None, result="attributes"):
        self.J, self.step, self.dtype, self.tile_rows = J, step, transform_dtype, tile_rows
        assert result in ("attributes", "psnr")
        self.result = result
        self.dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.s_in = torch.cuda.Stream(self.dev)
        self.s_out = torch.cuda.Stream(self.dev)
        self.queue = []                  # (device tensors, H2D-done event, input set)
        self.inputs = [None, None]       # two sets of device input tensors, reused (no allocation per scene)
        self.input_free = [None, None]   # event: the encode that read the set has finished
        self.n_submitted = 0
        self.pending = []                # outputs whose D2H may still be running (keeps the device tensors alive)
        self.pinned = {}                 # reusable pinned output buffers, two sets (ping-pong)
        self.flip = 0

    def submit(self, host_scene):
        """Start the H2D copy of a pinned host scene into one of the pipeline's two device input sets (the set is free
        once the encode that read it two submissions ago has finished: awaited on the copy-in stream, not on the host)."""
        slot = self.n_submitted & 1
        self.n_submitted += 1
        bufs = self.inputs[slot]
        if bufs is None or any(b.shape != h.shape or b.dtype != h.dtype for b, h in zip(bufs, host_scene)):
            bufs = [torch.empty(h.shape, dtype=h.dtype, device=self.dev) for h in host_scene]
            self.inputs[slot] = bufs
        with torch.cuda.stream(self.s_in):
            if self.input_free[slot] is not None:
                self.s_in.wait_event(self.input_free[slot])
            for b, h in zip(bufs, host_scene):
                b.copy_(h, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.s_in)
        self.queue.append((bufs, ev, slot))
        return sum(h.numel() * h.element_size() for h in host_scene)

    def _pin(self, name, numel, dtype):
        key = (name, self.flip)
        buf = self.pinned.get(key)
        if buf is None or buf.numel() < numel or buf.dtype != dtype:
            buf = torch.empty(max(numel, 1), dtype=dtype).pin_memory()
            self.pinned[key] = buf
        return buf[:numel]

    @torch.no_grad()
    def run(self):
        if len(self.queue) > 2:
            raise RuntimeError("ScenePipeline: at most two scenes may be in flight (submit, then run)")
        dv, ev, slot = self.queue.pop(0)
        main = torch.cuda.current_stream(self.dev)
        main.wait_event(ev)
        enc = encode_scene(*dv, J=self.J, step=self.step, transform_dtype=self.dtype, tile_rows=self.tile_rows,
                           keep_packed=self.result == "psnr")
        free = torch.cuda.Event()                        # the inputs have been read (merge is long done; the encoder has synchronised)
        free.record(main)
        self.input_free[slot] = free
        sizes = enc.streams.sizes()                      # (the encoder has synchronised: no extra wait)
        compact, _ = enc.streams.packed()                # end to end, each padded to 8 bytes
        coords, Crec = decode_scene(enc, reuse_tree=False, tile_rows=self.tile_rows)
        if self.result == "psnr":                        # -10 log10(sse / (nvox D) + 1e-10) on the host from one double
            final = ("sse", ops.sse(enc.packed, Crec))
        else:
            final = ("attributes", Crec.reshape(-1))
        done_compute = torch.cuda.Event()
        done_compute.record(main)
        idx = enc.streams.sync_index.view(torch.uint8).reshape(-1) if enc.streams.sync_index is not None else None
        outs = {"sizes": sizes, "nvox": enc.nvox}
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(done_compute)
            for name, t in (("streams", compact), ("index", idx), ("keys", enc.keys), final):
                if t is None:
                    continue
                t.record_stream(self.s_out)
                dst = self._pin(name, t.numel(), t.dtype)
                dst.copy_(t, non_blocking=True)
                outs[name] = dst
            done = torch.cuda.Event()
            done.record(self.s_out)
        outs["done"] = done
        outs["d2h_bytes"] = sum(outs[k].numel() * outs[k].element_size() for k in ("streams", "index", "keys", "attributes", "sse") if k in outs)
        self.pending.append(outs)
        if len(self.pending) > 2:                        # the ping-pong buffers of two steps ago are about to be reused
            self.pending.pop(0)["done"].synchronize()
        self.flip ^= 1
        return outs

    def finish(self):
        for o in self.pending:
            o["done"].synchronize()
        self.pending.clear()

"""Host-side operators over the C ABI: one Python function per entry point family.

Everything here takes and returns CUDA tensors; the arithmetic happens in libraht_b200.so.  The
reference-named shim modules (RAHT_param.py, RAHT.py, iRAHT.py, voxelize_pc.py, merge_cluster_cuda,
rlgr.py) and raht_b200.codec are built on these.
"""
import ctypes as C
from collections.abc import Sequence

import torch

from . import _lib as L
from ._lib import lib, check, ptr, stream, workspace

DEFAULT_TILE_ROWS = 128


def _status_word(device):
    return torch.zeros(1, dtype=torch.int32, device=device)


def raise_for_status(bits: int):
    """Map the device-side status bits (include/raht_b200.h RAHT_ST_*) to the reference's exceptions."""
    if bits & L.ST_SORT_STALL:
        raise RuntimeError("radix sort look-back stalled (internal error)")
    if bits & L.ST_OUT_OF_BOUNDS:
        raise ValueError("RAHT_param:OutOfBounds: voxel index out of [0, 2^depth - 1]")
    if bits & L.ST_NOT_SORTED_UNIQUE:
        raise ValueError("RAHT_param: rows must be Morton-sorted and unique (duplicate or unsorted voxels)")
    if bits & L.ST_INT32_OVERFLOW:
        raise OverflowError("quantised coefficient does not fit int32 (encode_3dgs.py:215 would truncate)")


# ------------------------------------------------------------------------------------------------
# voxelisation front end
# ------------------------------------------------------------------------------------------------
def morton_keys_i64(V: torch.Tensor, J: int) -> torch.Tensor:
    """get_morton_code (voxelize_pc.py:25-59)."""
    L.require_cuda(V, "V")
    V = V.to(torch.int64).contiguous()
    keys = torch.empty(V.shape[0], dtype=torch.int64, device=V.device)
    check(lib.raht_morton_keys_i64(ptr(V), V.shape[0], J, ptr(keys), stream()), "raht_morton_keys_i64")
    return keys


def morton_keys_f64(V: torch.Tensor, minV, width: float, depth: int):
    """Morton codes of RAHT_param.py:21-36; returns (keys int64, status int32[1])."""
    L.require_cuda(V, "V")
    V = V.to(torch.float64).contiguous()
    n = V.shape[0]
    keys = torch.empty(n, dtype=torch.int64, device=V.device)
    status = _status_word(V.device)
    mv = (C.c_double * 3)(*[float(x) for x in minV])
    check(lib.raht_morton_keys_f64(ptr(V), n, mv, float(width), depth, ptr(keys), ptr(status), stream()),
          "raht_morton_keys_f64")
    return keys, status


def radix_sort_pairs(keys: torch.Tensor, end_bit: int, vals: torch.Tensor = None, begin_bit: int = 0, status=None):
    """Stable LSD radix sort of 64-bit keys with 32-bit payloads (identity when vals is None)."""
    L.require_cuda(keys, "keys")
    n = keys.shape[0]
    dev = keys.device
    keys_out = torch.empty_like(keys)
    vals_out = torch.empty(n, dtype=torch.int32, device=dev)
    status = _status_word(dev) if status is None else status
    if n == 0:
        return keys_out, vals_out, status
    ws_bytes = lib.raht_radix_sort_workspace_bytes(n, begin_bit, end_bit)
    ws = L.workspace_cached("radix_sort", ws_bytes, dev)
    check(lib.raht_radix_sort_pairs(ptr(keys), ptr(keys_out), ptr(vals), ptr(vals_out), n, begin_bit, end_bit,
                                    ptr(status), ptr(ws), ws_bytes, stream()), "raht_radix_sort_pairs")
    return keys_out, vals_out, status


def voxel_segments(sorted_keys: torch.Tensor, want_unique=True):
    """Voxel boundaries of sorted keys -> (voxel_indices int64[nvox], cluster_offsets int32[nvox+1],
    unique_keys, nvox).  Synchronises (nvox must reach the host)."""
    n = sorted_keys.shape[0]
    dev = sorted_keys.device
    vi = torch.empty(n, dtype=torch.int64, device=dev)
    off = torch.empty(n + 1, dtype=torch.int32, device=dev)
    uk = torch.empty(n, dtype=sorted_keys.dtype, device=dev) if want_unique else None
    nvox = torch.zeros(1, dtype=torch.int64, device=dev)
    ws_bytes = lib.raht_voxel_segments_workspace_bytes(n)
    ws = workspace(ws_bytes, dev)
    check(lib.raht_voxel_segments(ptr(sorted_keys), n, ptr(vi), ptr(off), ptr(uk), ptr(nvox), ptr(ws), ws_bytes,
                                  stream()), "raht_voxel_segments")
    nv = int(L.read_host(nvox)[0][0])
    return vi[:nv], off[:nv + 1], (uk[:nv] if want_unique else None), nv


def merge_segments(cluster_indices, cluster_offsets, means, quats, scales, opacities, colors,
                   weight_by_opacity=True, separate=True, packed_dtype=None):
    """merge_clusters_cuda (cuda/merge_cluster_wrapper.cu:11-116).  Returns (5 tensors or None, packed or None)."""
    dev = means.device
    nclu = cluster_offsets.shape[0] - 1
    cd = colors.shape[1]
    outs = None
    if separate:
        outs = (torch.zeros(nclu, 3, dtype=torch.float32, device=dev), torch.zeros(nclu, 4, dtype=torch.float32, device=dev),
                torch.zeros(nclu, 3, dtype=torch.float32, device=dev), torch.zeros(nclu, dtype=torch.float32, device=dev),
                torch.zeros(nclu, cd, dtype=torch.float32, device=dev))
    packed = None
    tag = L.RAHT_F32
    if packed_dtype is not None:
        packed = torch.empty(nclu, 11 + cd, dtype=packed_dtype, device=dev)
        tag = L.dtype_tag(packed_dtype)
    o = outs or (None,) * 5
    check(lib.raht_merge_segments(ptr(cluster_indices), ptr(cluster_offsets), nclu, ptr(means), ptr(quats), ptr(scales),
                                  ptr(opacities), ptr(colors), cd, 1 if weight_by_opacity else 0, ptr(o[0]), ptr(o[1]),
                                  ptr(o[2]), ptr(o[3]), ptr(o[4]), ptr(packed), tag, stream()), "raht_merge_segments")
    return outs, packed


# ------------------------------------------------------------------------------------------------
# tree
# ------------------------------------------------------------------------------------------------
class _PlanHandle:
    def __init__(self, handle, ws):
        self.handle = handle
        self.ws = ws            # the plan lives in this workspace tensor

    def __del__(self):
        h, self.handle = self.handle, None
        if h:
            try:
                lib.raht_plan_destroy(h)
            except Exception:
                pass


class RahtTree:
    """Compact device tree: level[] (uint8 per row) + lazily built transform plans.

    Everything RAHT_param returns is derived from level[] on demand."""

    def __init__(self, level: torch.Tensor, n: int):
        self.level = level
        self.n = n
        self.device = level.device
        self._plans = {}
        self._w = None
        self._max_level = None
        self._order = None
        self._position = None

    # -- sizes ------------------------------------------------------------------------------------
    @property
    def max_level(self) -> int:
        if self._max_level is None:
            out = torch.empty(1, dtype=torch.int32, device=self.device)
            check(lib.raht_tree_max_level(ptr(self.level), self.n, ptr(out), stream()), "raht_tree_max_level")
            self._max_level = int(L.read_host(out)[0].reshape(-1)[0])
        return self._max_level

    @property
    def nlevels(self) -> int:
        """len(List) == len(Flags) == len(weights) of RAHT_param.py (one level when N == 1)."""
        return max(self.max_level, 0) + 1

    # -- plan -------------------------------------------------------------------------------------
    def plan(self, tile_rows: int = None):
        tile_rows = tile_rows or DEFAULT_TILE_ROWS
        if tile_rows not in self._plans:
            w = torch.empty(self.n, dtype=torch.float64, device=self.device)
            h = C.c_void_p()
            ws_bytes = lib.raht_plan_workspace_bytes(self.n, tile_rows)
            ws = workspace(ws_bytes, self.device)
            check(lib.raht_plan_create(ptr(self.level), self.n, tile_rows, ptr(w), ptr(ws), ws_bytes, C.byref(h), stream()),
                  "raht_plan_create")
            self._plans[tile_rows] = _PlanHandle(h, ws)
            self._w = w
        return self._plans[tile_rows].handle

    def plan_info(self, tile_rows: int = None):
        h = self.plan(tile_rows)
        buf = (C.c_int64 * 48)()
        check(lib.raht_plan_info(h, buf, 48), "raht_plan_info")
        st = int(buf[2])
        return {"n": int(buf[0]), "tile_rows": int(buf[1]), "stages": st, "bytes": int(buf[3]),
                "stage_tiles": [int(buf[4 + k]) for k in range(st)]}

    @property
    def w(self) -> torch.Tensor:
        if self._w is None:
            self.plan()
        return self._w

    # -- one level of List / Flags / weights ---------------------------------------------------------
    def level_arrays(self, l: int):
        n, dev = self.n, self.device
        lst = torch.empty(n, dtype=torch.int64, device=dev)
        flg = torch.empty(n, dtype=torch.uint8, device=dev)
        wts = torch.empty(n, dtype=torch.int64, device=dev)
        cnt = torch.zeros(1, dtype=torch.int64, device=dev)
        ws_bytes = lib.raht_tree_level_workspace_bytes(n)
        ws = workspace(ws_bytes, dev)
        check(lib.raht_tree_level(ptr(self.level), n, l, ptr(lst), ptr(flg), ptr(wts), ptr(cnt), ptr(ws), ws_bytes,
                                  stream()), "raht_tree_level")
        c = int(L.read_host(cnt)[0].reshape(-1)[0])
        return lst[:c].clone(), flg[:c].to(torch.bool), wts[:c].clone()

    # -- order_RAGFT ------------------------------------------------------------------------------
    def order(self, want_position=False):
        if self._order is None:
            n, dev = self.n, self.device
            order = torch.empty(n, dtype=torch.int64, device=dev)
            pos = torch.empty(n, dtype=torch.int32, device=dev)
            ws_bytes = lib.raht_order_ragft_workspace_bytes(n)
            ws = workspace(ws_bytes, dev)
            check(lib.raht_order_ragft(ptr(self.level), n, ptr(order), ptr(pos), ptr(ws), ws_bytes, stream()),
                  "raht_order_ragft")
            self._order, self._position = order, pos
        return (self._order, self._position) if want_position else self._order


def tree_from_keys(MC: torch.Tensor, validate=True, status=None) -> RahtTree:
    """status: a shared device status word; when given the sorted/unique check is left to the caller
    (raise_for_status at its next synchronisation point) and this call does not synchronise."""
    n = MC.shape[0]
    level = torch.empty(n, dtype=torch.uint8, device=MC.device)
    deferred = status is not None
    status = _status_word(MC.device) if status is None else status
    check(lib.raht_tree_levels(ptr(MC), n, ptr(level), ptr(status), stream()), "raht_tree_levels")
    if validate and not deferred:
        s = int(L.read_host(status)[0][0])
        if s & L.ST_NOT_SORTED_UNIQUE:
            raise ValueError("RAHT_param: rows must be Morton-sorted and unique (duplicate or unsorted voxels)")
    return RahtTree(level, n)


class LevelSequence(Sequence):
    """Lazy stand-in for the per-level Python lists of RAHT_param.py (List / Flags / weights): a
    level is materialised on the device only when it is indexed.  Materialised tensors carry a
    reference to the tree, so the transform finds its plan even after a caller rebuilds the list
    with `[t.to(device) for t in List]` (encode_3dgs.py:152-154)."""

    def __init__(self, tree: RahtTree, kind: int, one_based=False, cache=None):
        self._raht_tree = tree
        self._kind = kind
        self._one = one_based
        self._cache = cache if cache is not None else {}

    def __len__(self):
        return self._raht_tree.nlevels

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        n = len(self)
        if i < 0:
            i += n
        if not 0 <= i < n:
            raise IndexError(i)
        if i not in self._cache:
            self._cache[i] = self._raht_tree.level_arrays(i)
        t = self._cache[i][self._kind]
        if self._kind == 0 and self._one:
            t = t + 1
        t._raht_tree = self._raht_tree
        return t


def tree_of(List, Flags=None, weights=None, one_based=False) -> RahtTree:
    """Recover the RahtTree behind the (List, Flags, weights) a caller passes to RAHT()/iRAHT().
    Plain lists of tensors that did not come from this package are converted by rebuilding level[]
    from Flags: row List[l][k+1] is a right sibling at level l iff Flags[l][k]."""
    t = getattr(List, "_raht_tree", None)
    if t is None and len(List) > 0:
        t = getattr(List[0], "_raht_tree", None)
    if t is not None:
        return t
    if Flags is None:
        raise ValueError("cannot recover the RAHT tree: Flags required")
    first = List[0]
    L.require_cuda(first, "List[0]")
    n = int(first.shape[0])
    level = torch.full((n,), 255, dtype=torch.uint8, device=first.device)
    for l in range(len(Flags)):
        lst = List[l].to(first.device).long()
        if one_based:
            lst = lst - 1
        f = Flags[l].to(first.device).bool()
        right = lst[1:][f[:-1]]
        level[right] = l
    return RahtTree(level, n)


# ------------------------------------------------------------------------------------------------
# transform / quantisation / entropy coding
# ------------------------------------------------------------------------------------------------
def forward(tree: RahtTree, Cmat: torch.Tensor, out_dtype=torch.float64, tile_rows=None, flags=0) -> torch.Tensor:
    L.require_cuda(Cmat, "C")
    if Cmat.dtype not in (torch.float32, torch.float64):
        Cmat = Cmat.to(torch.float64)
    Cmat = Cmat.contiguous()
    n, D = Cmat.shape
    if n != tree.n:
        raise ValueError(f"C has {n} rows, tree has {tree.n}")
    T = torch.empty(n, D, dtype=out_dtype, device=Cmat.device)
    check(lib.raht_forward(tree.plan(tile_rows), ptr(Cmat), L.dtype_tag(Cmat.dtype), D, ptr(T), L.dtype_tag(out_dtype),
                           flags, stream()), "raht_forward")
    return T


def inverse(tree: RahtTree, T: torch.Tensor, tile_rows=None, flags=0) -> torch.Tensor:
    L.require_cuda(T, "T")
    if T.dtype not in (torch.float32, torch.float64):
        T = T.to(torch.float64)
    T = T.contiguous()
    n, D = T.shape
    if n != tree.n:
        raise ValueError(f"T has {n} rows, tree has {tree.n}")
    Cmat = torch.empty_like(T)
    check(lib.raht_inverse(tree.plan(tile_rows), ptr(T), L.dtype_tag(T.dtype), D, ptr(Cmat), flags, stream()),
          "raht_inverse")
    return Cmat


def shard_top(first_keys, prev_last_keys, offsets, rows, inverse=False):
    """The butterflies that join P Morton-prefix ranges (raht_shard_top): rows (P, D) = the ranges' low-pass rows,
    transformed in place.  Returns (rows, level (P,) uint8 = level of each range's first row in the whole scene)."""
    L.require_cuda(rows, "rows")
    P, D = rows.shape
    rows = rows.contiguous()
    level = torch.empty(P, dtype=torch.uint8, device=rows.device)
    status = _status_word(rows.device)
    check(lib.raht_shard_top(ptr(first_keys.contiguous()), ptr(prev_last_keys.contiguous()), ptr(offsets.contiguous()), P, ptr(rows),
                             L.dtype_tag(rows.dtype), D, 1 if inverse else 0, ptr(level), ptr(status), stream()), "raht_shard_top")
    return rows, level


def _steps_tensor(step, D, device):
    if isinstance(step, torch.Tensor):
        s = step.to(device=device, dtype=torch.float64).reshape(-1)
        if s.numel() == 1:
            s = s.expand(D)
        return s.contiguous()
    if isinstance(step, (int, float)):
        return torch.full((D,), float(step), dtype=torch.float64, device=device)
    return torch.tensor([float(x) for x in step], dtype=torch.float64, device=device)


def quant_reorder(T: torch.Tensor, order, step, check_overflow=True, status=None, division="exact") -> torch.Tensor:
    """q[c, p] = int32(floor(T[order[p], c] / step_c + 0.5)), channel-major (D, n).
    status: shared device status word (no synchronisation here; the caller checks it later).
    division: "exact" = true division (torch on the CPU: the parity convention); "reciprocal" = T * (1 / step), what
    torch evaluates on CUDA for `Coeff / step` with a Python scalar (encode_3dgs.py:204 on the reference's production
    device) -- the two differ in a handful of symbols when the step is not a power of two."""
    if division not in ("exact", "reciprocal"):
        raise ValueError(f"quant_reorder: unknown division mode {division!r}")
    n, D = T.shape
    dev = T.device
    T = T.contiguous()
    steps = _steps_tensor(step, D, dev)
    ld = (n + 3) // 4 * 4
    q = torch.empty(D, ld, dtype=torch.int32, device=dev)
    deferred = status is not None
    status = _status_word(dev) if status is None else status
    check(lib.raht_quant_reorder(ptr(T), L.dtype_tag(T.dtype), n, D, ptr(order), ptr(steps), ptr(q), ld, ptr(status),
                                 1 if division == "reciprocal" else 0, stream()), "raht_quant_reorder")
    if check_overflow and not deferred and int(L.read_host(status)[0][0]) & L.ST_INT32_OVERFLOW:
        raise OverflowError("quantised coefficient does not fit int32 (encode_3dgs.py:215 would truncate)")
    return q[:, :n]


def dequant_unreorder(q: torch.Tensor, order, step, out_dtype=torch.float64) -> torch.Tensor:
    """T[order[p], c] = q[c, p] * step_c."""
    D, n = q.shape
    dev = q.device
    if q.stride(1) != 1:
        q = q.contiguous()
    ld = q.stride(0) if D > 1 else max(n, 1)
    steps = _steps_tensor(step, D, dev)
    T = torch.empty(n, D, dtype=out_dtype, device=dev)
    check(lib.raht_dequant_unreorder(ptr(q), ld, n, D, ptr(order), ptr(steps), ptr(T), L.dtype_tag(out_dtype), stream()),
          "raht_dequant_unreorder")
    return T


def forward_quant(tree: RahtTree, Cmat: torch.Tensor, step, out_dtype=torch.float64, tile_rows=None, flags=0, status=None) -> torch.Tensor:
    """quant_reorder(forward(tree, C), order_RAGFT, step) in one pass (raht_forward_quant): the coefficients of the rows a
    stage-0 tile finishes are quantised out of shared memory and never written as fp64.  Same q, bit for bit."""
    L.require_cuda(Cmat, "C")
    if Cmat.dtype not in (torch.float32, torch.float64):
        Cmat = Cmat.to(torch.float64)
    if Cmat.dtype == torch.float64 and out_dtype == torch.float32:
        Cmat = Cmat.to(torch.float32)
    Cmat = Cmat.contiguous()
    n, D = Cmat.shape
    if n != tree.n:
        raise ValueError(f"C has {n} rows, tree has {tree.n}")
    dev = Cmat.device
    _, pos = tree.order(want_position=True)
    steps = _steps_tensor(step, D, dev)
    ld = max((n + 3) // 4 * 4, 1)
    q = torch.empty(D, ld, dtype=torch.int32, device=dev)
    T = torch.empty(n, D, dtype=out_dtype, device=dev)                  # scratch: only the deferred rows are written
    deferred = status is not None
    status = _status_word(dev) if status is None else status
    check(lib.raht_forward_quant(tree.plan(tile_rows), ptr(Cmat), L.dtype_tag(Cmat.dtype), D, ptr(T), L.dtype_tag(out_dtype), ptr(pos),
                                 ptr(steps), ptr(q), ld, ptr(status), flags, stream()), "raht_forward_quant")
    if not deferred and int(L.read_host(status)[0][0]) & L.ST_INT32_OVERFLOW:
        raise OverflowError("quantised coefficient does not fit int32 (encode_3dgs.py:215 would truncate)")
    return q[:, :n]


def dequant_inverse(tree: RahtTree, q: torch.Tensor, step, out_dtype=torch.float64, tile_rows=None, flags=0) -> torch.Tensor:
    """inverse(tree, dequant_unreorder(q, order_RAGFT, step)) in one pass (raht_dequant_inverse)."""
    L.require_cuda(q, "q")
    D, n = q.shape
    if n != tree.n:
        raise ValueError(f"q has {n} columns, tree has {tree.n}")
    dev = q.device
    if q.dtype != torch.int32 or q.stride(1) != 1:
        q = q.to(torch.int32).contiguous()
    ld = q.stride(0) if D > 1 else max(n, 1)
    _, pos = tree.order(want_position=True)
    steps = _steps_tensor(step, D, dev)
    T = torch.empty(n, D, dtype=out_dtype, device=dev)                  # scratch (deferred rows)
    Cmat = torch.empty(n, D, dtype=out_dtype, device=dev)
    check(lib.raht_dequant_inverse(tree.plan(tile_rows), ptr(q), ld, ptr(pos), ptr(steps), D, ptr(T), L.dtype_tag(out_dtype), ptr(Cmat),
                                   flags, stream()), "raht_dequant_inverse")
    return Cmat


class RlgrStreams:
    """nch closed RLGR streams in one device buffer (slot per channel).

    sync_index (optional): the encoder's side information for parallel decoding -- 8 bytes per range
    of 256 symbols (bit offset, coder state, position of the first non-zero); the streams themselves
    are the reference bitstreams and do not depend on it."""

    def __init__(self, buf, slot_bytes, nbytes, n, nch, sync_index=None):
        self.buf, self.slot_bytes, self.nbytes, self.n, self.nch = buf, slot_bytes, nbytes, n, nch
        self.sync_index = sync_index
        self.stats = None
        self._sizes = None

    def sizes(self):
        """Stream lengths in bytes on the host (one D2H, cached)."""
        if self._sizes is None:
            self._sizes = [int(x) for x in L.read_host(self.nbytes)[0]]
        return self._sizes

    def total_bytes(self):
        return int(sum(self.sizes()))

    def packed(self, out=None, status=None):
        """The nch streams end to end, each padded with zero bytes to a multiple of 8 (raht_streams_pack) ->
        (uint8 tensor, device offsets (nch + 1) int64).  `out`: an 8-byte-aligned uint8 slice to fill, else allocated."""
        total = sum((int(b) + 7) // 8 * 8 for b in self.sizes())
        if out is None:
            out = torch.empty(total, dtype=torch.uint8, device=self.buf.device)
        offsets = torch.empty(self.nch + 1, dtype=torch.int64, device=self.buf.device)
        check(lib.raht_streams_pack(ptr(self.buf), self.slot_bytes, ptr(self.nbytes), self.nch, ptr(out), out.numel(),
                                    ptr(offsets), ptr(status), stream()), "raht_streams_pack")
        return out, offsets

    def index_bytes(self):
        return 0 if self.sync_index is None else self.sync_index.numel() * self.sync_index.element_size()

    def to_bytes(self):
        """list[bytes], one per channel (D2H)."""
        sizes = self.sizes()
        host = self.buf.cpu().numpy()
        return [host[c * self.slot_bytes: c * self.slot_bytes + sizes[c]].tobytes() for c in range(self.nch)]


def rlgr_encode(q: torch.Tensor, flag_signed=1, slot_bytes=None, with_index=True, want_stats=False, status=None) -> RlgrStreams:
    """q: (nch, n) int32 channel-major (row stride >= n).
    status: shared device status word of the preceding pipeline steps; it is read here (the one
    synchronisation of the encoder after the voxel count) and any earlier failure is raised."""
    L.require_cuda(q, "q")
    nch, n = q.shape
    dev = q.device
    if q.dtype != torch.int32 or q.stride(1) != 1:
        q = q.to(torch.int32).contiguous()
    ld = q.stride(0) if nch > 1 else max(n, 1)
    slot = slot_bytes or ((n * 4 + 64 + 7) // 8 * 8)
    entries = int(lib.raht_rlgr_sync_entries(n))
    ws_bytes = lib.raht_rlgr_encode_workspace_bytes(n, nch)
    ws = L.workspace_cached("rlgr_encode", ws_bytes, dev)
    while True:
        buf = torch.empty(nch * slot, dtype=torch.uint8, device=dev)
        nbytes = torch.zeros(nch, dtype=torch.int64, device=dev)
        index = torch.empty(nch * entries, 2, dtype=torch.int32, device=dev) if with_index else None
        status = _status_word(dev) if status is None else status
        check(lib.raht_rlgr_encode(ptr(q), n, ld, nch, flag_signed, ptr(buf), slot, ptr(nbytes), ptr(index), ptr(status),
                                   ptr(ws), ws_bytes, stream()), "raht_rlgr_encode")
        st_host, nb_host = L.read_host(status, nbytes)                # the encoder's one synchronisation
        bits = int(st_host[0])
        raise_for_status(bits)
        if bits & L.ST_STREAM_OVERFLOW:
            slot = (int(nb_host.max()) + 64 + 7) // 8 * 8             # exact size is known even on overflow
            status.zero_()
            continue
        out = RlgrStreams(buf, slot, nbytes, n, nch, index)
        out._sizes = [int(x) for x in nb_host]
        if want_stats and n > 0 and nch > 0:
            st = (C.c_ulonglong * 6)()
            check(lib.raht_rlgr_encode_stats(ptr(ws), n, nch, st, stream()), "raht_rlgr_encode_stats")
            out.stats = {"solves": int(st[0]), "rounds": int(st[1]), "repairs": int(st[2]), "repair_rounds": int(st[4]), "table_hits": int(st[5]),
                         "chunks": nch * entries,
                         "slowest_chunk": {"cycles": int(st[3]) >> 24, "channel": (int(st[3]) >> 8) & 0xffff, "rounds": int(st[3]) & 0xff}}
        return out


def rlgr_build_index(streams: RlgrStreams) -> torch.Tensor:
    """Sync index (nch, entries, 2) uint32 of streams that carry none (reference-coder streams): raht_rlgr_build_index."""
    n, nch = streams.n, streams.nch
    entries = lib.raht_rlgr_sync_entries(n)
    index = torch.empty(nch, max(entries, 1), 2, dtype=torch.int32, device=streams.buf.device)
    check(lib.raht_rlgr_build_index(ptr(streams.buf), streams.slot_bytes, ptr(streams.nbytes), n, nch, ptr(index), stream()),
          "raht_rlgr_build_index")
    return index[:, :entries]


def rlgr_decode(streams: RlgrStreams, flag_signed=1, use_index=True, serial=False) -> torch.Tensor:
    """Parallel over 256-symbol ranges through the sync index: the encoder's, or -- for streams without one (written by the
    reference coder) or with use_index=False -- one rebuilt by a walk of the streams (rlgr_build_index).  serial=True:
    the plain one-thread-per-channel decoder (kept as the second opinion of the tests)."""
    dev = streams.buf.device
    n, nch = streams.n, streams.nch
    ld = (n + 3) // 4 * 4
    q = torch.empty(nch, max(ld, 1), dtype=torch.int32, device=dev)
    index = streams.sync_index if use_index else None
    if index is None and not serial and n > 0 and nch > 0:
        index = rlgr_build_index(streams)
    check(lib.raht_rlgr_decode(ptr(streams.buf), streams.slot_bytes, ptr(streams.nbytes), ptr(index), n, max(ld, 1), nch,
                               flag_signed, ptr(q), stream()), "raht_rlgr_decode")
    return q[:, :n]


def channel_minmax(T: torch.Tensor):
    """(min (D,), max (D,)) fp64 device tensors of the columns of an (n, D) fp32 / fp64 matrix (raht_channel_minmax)."""
    L.require_cuda(T, "T")
    T = T.contiguous()
    n, D = T.shape
    if n < 1:
        raise ValueError("channel_minmax of an empty matrix")
    out = torch.empty(2, D, dtype=torch.float64, device=T.device)
    ws_bytes = lib.raht_channel_minmax_workspace_bytes(n, D)
    ws = workspace(ws_bytes, T.device)
    check(lib.raht_channel_minmax(ptr(T), L.dtype_tag(T.dtype), n, D, ptr(out), ptr(ws), ws_bytes, stream()), "raht_channel_minmax")
    return out[0], out[1]


def sse(A: torch.Tensor, B: torch.Tensor, c0=0, c1=None) -> torch.Tensor:
    """Sum of squared differences of two (n, D) matrices (fp32 or fp64 each), accumulated in fp64 -> device double."""
    n, D = A.shape
    c1 = D if c1 is None else c1
    out = torch.empty(1, dtype=torch.float64, device=A.device)
    ws_bytes = lib.raht_sse_workspace_bytes(n, D)
    ws = workspace(ws_bytes, A.device)
    A, B = A.contiguous(), B.contiguous()
    check(lib.raht_sse(ptr(A), L.dtype_tag(A.dtype), ptr(B), L.dtype_tag(B.dtype), n, D, c0, c1, ptr(out), ptr(ws), ws_bytes, stream()),
          "raht_sse")
    return out

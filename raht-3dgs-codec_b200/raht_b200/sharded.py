"""Morton-prefix sharding of ONE scene over the GPUs of a box (BASELINE config 4).

The reference is single-GPU; this module keeps its arithmetic and its bitstream.  The bounding cube is cut into
2^B CELLS by the top B bits of the Morton code (aligned subtrees of the RAHT tree); rank r of P owns a CONTIGUOUS run of
cells, chosen from the all-reduced cell histogram so that every rank gets about N / P Gaussians whatever the scene looks
like -- and whatever P is (a rank's range is a union of aligned subtrees by construction; the first version cut the prefix
space evenly, which is not aligned for P = 5, 7, ...).  The concatenation of the ranks' Morton-sorted voxels in rank order
IS the sorted voxel list of the scene.  One process per GPU, `torch.distributed` (NCCL over NVLink) for the exchanges:

  encode   1 all_reduce   scene bounding box (6 floats)                                  [voxelize_pc.py:87-95]
           2 all_reduce   cell histogram (2^B counts) -> the same owner table on every rank
           3 all_to_all   the raw Gaussians to their owner, as ONE exchange of packed 236-byte rows (raht_pack_rows)
           --             voxelise + merge, then the RAHT of the rank's cells as a FOREST: every butterfly below the cell
                          roots; each cell's low-pass is left in its first row
           4 all_reduce   one row per cell (2^B x D, the owner contributes, the others add zero bits) + per cell
                          (rows, first code, last code); every rank joins the cells (raht_shard_top_cells)
           5 all_gather   rows and the 22 order-bucket counts per rank
           6 all_to_all   quantised coefficients, row-sharded -> channel-sharded in order_RAGFT order
           --             rank k entropy-codes channels k, k + P, ... of the WHOLE scene
  decode   the mirror image.

Result: coefficients, order and RLGR bytes are bit-identical to the single-GPU encode of the whole scene
(tests/test_sharded.py, world sizes 2, 3, 5 and 8).  The compute goes through a small backend object: `DeviceBackend`
(the CUDA library) is the only one shipped; the CPU test-suite plugs its own checker backend in to exercise the host
logic (partition, cell levels, exchanges) over gloo.
"""
from dataclasses import dataclass, field
from typing import List, Optional

import torch
import torch.distributed as dist

N_BUCKETS = 22           # order_RAGFT groups: root, then octree levels 20 .. 0 (tree.cu OR_BUCKETS)


# NCCL's all_to_all is send / recv pairs over its point-to-point channels; with its defaults the two large exchanges of
# this module (the scene's rows, the quantised coefficients: GBs per rank) ran at 210-250 GB/s per rank over NVLink 5.
# Measured on 4 x B200 (50 M Gaussians): these settings take the row exchange 13.8 -> 10.0 ms, the coefficient exchange
# 12.5 -> 8.7 and 8.9 -> 4.7 ms, the step 81.3 -> 70.4 ms; neither half alone (more channels / more channels per peer)
# changes anything, larger buffers or chunks do not help.  Must be in the environment before the process group is created.
NCCL_EXCHANGE_ENV = {"NCCL_MIN_P2P_NCHANNELS": "64", "NCCL_MAX_P2P_NCHANNELS": "64", "NCCL_NCHANNELS_PER_PEER": "32"}


def tune_nccl_env():
    """Set NCCL_EXCHANGE_ENV where the user has not set the variables; call before dist.init_process_group('nccl')."""
    import os
    for k, v in NCCL_EXCHANGE_ENV.items():
        os.environ.setdefault(k, v)


def order_bucket(level: torch.Tensor) -> torch.Tensor:
    """Bucket of order_RAGFT for every row from its level (tree.cu or_bucket)."""
    lv = level.to(torch.int64)
    return torch.where(lv == 255, torch.zeros_like(lv), 1 + (20 - torch.clamp(lv, max=62) // 3))


CELL_BITS_PER_RANK = 5   # ~32 cells per rank on average: the balance error of a rank is one cell
CELL_BITS_MAX = 10       # raht_shard_top_cells joins up to 1024 cells


def cell_bits(world: int, J: int) -> int:
    """B: the top B bits of the 3J-bit Morton code number the cells."""
    return max(0, min(3 * J, CELL_BITS_MAX, (max(world, 1) - 1).bit_length() + CELL_BITS_PER_RANK)) if world > 1 else 0


def owner_table(hist: torch.Tensor, world: int) -> torch.Tensor:
    """Owner rank of every cell from the global histogram (device tensor, no synchronisation): a cell goes to the rank in
    whose share [r N / P, (r + 1) N / P) its first Gaussian falls -- contiguous runs of cells, balanced to within one cell."""
    h = hist.to(torch.int64)
    start = torch.cumsum(h, 0) - h
    total = torch.clamp(h.sum(), min=1)
    return torch.clamp((start * world) // total, max=world - 1)


class DeviceBackend:
    """The CUDA library (raht_b200.ops); tensors live on the current CUDA device."""

    def __init__(self, tile_rows=None):
        from . import ops, codec
        self.ops, self.codec, self.tile_rows = ops, codec, tile_rows

    def bounds(self, means):
        from ._lib import lib, check, ptr, stream
        b = torch.empty(6, dtype=torch.float32, device=means.device)
        check(lib.raht_bounds3_f32(ptr(means), means.shape[0], 3, ptr(b), stream()), "raht_bounds3_f32")
        return b[:3], b[3:]

    def keys(self, means, J, vmin, width):
        return self.codec.morton_keys_given(means, J, vmin, width)

    def histogram(self, values, shift, nbins):
        """counts (nbins,) int64 of values >> shift (raht_histogram_i64: shared-memory counters, no contended global atomics)."""
        from ._lib import lib, check, ptr, stream
        out = torch.empty(nbins, dtype=torch.int64, device=values.device)
        check(lib.raht_histogram_i64(ptr(values.contiguous()), values.shape[0], shift, nbins, ptr(out), stream()), "raht_histogram_i64")
        return out

    def pack_by_owner(self, scene, owner, world):
        """Gaussians as packed (n, D) fp32 rows grouped by owner rank (stable): one hand-written radix pass over the owner
        ids gives the permutation, raht_pack_rows gathers the five attribute arrays through it."""
        from ._lib import lib, check, ptr, stream
        means, quats, scales, opac, colors = [t.contiguous().float() for t in scene]
        n, cd = means.shape[0], colors.shape[1]
        out = torch.empty(n, 11 + cd, dtype=torch.float32, device=means.device)
        if n == 0:
            return out
        _, perm, _ = self.ops.radix_sort_pairs(owner.to(torch.int64).contiguous(), max((world - 1).bit_length(), 1))
        check(lib.raht_pack_rows(ptr(perm), n, ptr(means), ptr(quats), ptr(scales), ptr(opac), ptr(colors), cd, ptr(out), stream()),
              "raht_pack_rows")
        return out

    def unpack_rows(self, rows, color_dim):
        from ._lib import lib, check, ptr, stream
        n, dev = rows.shape[0], rows.device
        outs = [torch.empty(n, 3, dtype=torch.float32, device=dev), torch.empty(n, 4, dtype=torch.float32, device=dev),
                torch.empty(n, 3, dtype=torch.float32, device=dev), torch.empty(n, dtype=torch.float32, device=dev),
                torch.empty(n, color_dim, dtype=torch.float32, device=dev)]
        check(lib.raht_unpack_rows(ptr(rows.contiguous()), n, color_dim, *[ptr(o) for o in outs], stream()), "raht_unpack_rows")
        return tuple(outs)

    def voxelize_merge(self, scene, J, vmin, width):
        packed, keys, _ = self.codec.voxelize_merge(*scene, J=J, given=(vmin, width))
        return packed, keys

    def levels(self, keys, cut):
        """level[] of the rank's rows with every cell a tree of its own: rows that start a cell (their level in the whole
        scene is >= cut, the lowest prefix bit) become roots (255)."""
        level = self.ops.tree_from_keys(keys).level
        if keys.shape[0]:
            level[level >= cut] = 255
        return level

    def forward_local(self, keys, level, packed, dtype):
        tree = self.ops.RahtTree(level, level.shape[0])
        return self.ops.forward(tree, packed, dtype, self.tile_rows)

    def inverse_local(self, keys, level, T):
        tree = self.ops.RahtTree(level, level.shape[0])
        return self.ops.inverse(tree, T, self.tile_rows)

    def shard_top_cells(self, first, last, counts, rows, inverse):
        """raht_shard_top_cells: rows (C, D) in place -> (rows, level of every cell's first row in the whole scene)."""
        import ctypes as C
        from ._lib import lib, check, ptr, stream, dtype_tag
        ncell, D = rows.shape
        level = torch.empty(ncell, dtype=torch.uint8, device=rows.device)
        status = torch.zeros(1, dtype=torch.int32, device=rows.device)
        check(lib.raht_shard_top_cells(ptr(first.contiguous()), ptr(last.contiguous()), ptr(counts.contiguous()), ncell, ptr(rows),
                                       dtype_tag(rows.dtype), D, 1 if inverse else 0, ptr(level), ptr(status), stream()),
              "raht_shard_top_cells")
        return rows, level

    def order(self, level_global):
        return self.ops.RahtTree(level_global, level_global.shape[0]).order()

    def quant_reorder(self, T, order, step):
        return self.ops.quant_reorder(T, order, step)

    def dequant_unreorder(self, q, order, step, dtype):
        return self.ops.dequant_unreorder(q, order, step, dtype)

    def rlgr_encode(self, q):
        return self.ops.rlgr_encode(q, 1)

    def rlgr_decode(self, streams):
        return self.ops.rlgr_decode(streams, 1)

    def stream_bytes(self, streams) -> List[bytes]:
        return streams.to_bytes()


@dataclass
class ShardedEncoded:
    J: int
    D: int
    step: object
    world: int
    rank: int
    n_rows: List[int]                    # voxels per rank
    keys: torch.Tensor                   # this rank's voxels (Morton codes, sorted)
    cell_bits: int                       # B: the scene is cut into 2^B cells
    cell_meta: torch.Tensor              # (2^B, 3) int64: voxels, first and last Morton code of every cell of the SCENE
    cell_level: torch.Tensor             # (2^B,) uint8: level of every cell's first row in the whole scene
    bucket_counts: torch.Tensor          # (P, 22)
    channels: List[int]                  # channels whose whole-scene stream this rank holds
    streams: object                      # the backend's stream object for those channels
    vmin: torch.Tensor = None
    width: float = 0.0
    transform_dtype: torch.dtype = torch.float64
    collective_bytes: dict = field(default_factory=dict)      # bytes this rank put into each exchange of the encode

    @property
    def n_total(self):
        return sum(self.n_rows)


class ShardedCodec:
    def __init__(self, backend=None, group=None):
        self.be = backend or DeviceBackend()
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    # ---- collectives (plain tensors; NCCL on the GPU, gloo in the CPU tests) --------------------------
    def _staged(self, t):
        """gloo moves host memory only: CUDA tensors are staged through the host (tests on one GPU); NCCL takes them as is."""
        return t.is_cuda and dist.get_backend(self.group) == "gloo"

    def _all_gather(self, t: torch.Tensor) -> torch.Tensor:
        if self.world == 1:
            return t.unsqueeze(0).clone()
        if self._staged(t):
            h = t.cpu()
            out = [torch.empty_like(h) for _ in range(self.world)]
            dist.all_gather(out, h.contiguous(), group=self.group)
            return torch.stack(out).to(t.device)
        out = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(out, t.contiguous(), group=self.group)
        return torch.stack(out)

    def _all_reduce(self, t, op):
        if self._staged(t):
            h = t.cpu()
            dist.all_reduce(h, op=op, group=self.group)
            t.copy_(h)
        else:
            dist.all_reduce(t, op=op, group=self.group)

    def _exchange(self, send: torch.Tensor, send_sizes: List[int], recv_sizes: List[int]) -> torch.Tensor:
        """all_to_all of flat buffers: elements [sum(send_sizes[:k]), +send_sizes[k]) of `send` go to rank k; returns what
        every rank sent here, in rank order, as ONE flat tensor (no concatenation on either side)."""
        send = send.reshape(-1)
        if self.world == 1:
            return send
        recv = torch.empty(sum(recv_sizes), dtype=send.dtype, device=send.device)
        if dist.get_backend(self.group) == "gloo":              # host path of the tests: pairwise send / recv
            so = [0]
            for x in send_sizes:
                so.append(so[-1] + x)
            ro = [0]
            for x in recv_sizes:
                ro.append(ro[-1] + x)
            host_send = send.cpu()
            host_recv = torch.empty(sum(recv_sizes), dtype=send.dtype)
            reqs = []
            for k in range(self.world):
                if k == self.rank:
                    host_recv[ro[k]:ro[k + 1]] = host_send[so[k]:so[k + 1]]
                    continue
                reqs.append(dist.isend(host_send[so[k]:so[k + 1]].contiguous(), k, group=self.group))
                reqs.append(dist.irecv(host_recv[ro[k]:ro[k + 1]], k, group=self.group))
            for q in reqs:
                q.wait()
            recv.copy_(host_recv)
            return recv
        dist.all_to_all_single(recv, send, output_split_sizes=recv_sizes, input_split_sizes=send_sizes, group=self.group)
        return recv

    # ---- partition ---------------------------------------------------------------------------------
    def scene_bounds(self, means):
        """Bounding box of the whole scene: vmin (3,) fp32 and width (python float), voxelize_pc.py:87-95."""
        lo, hi = self.be.bounds(means) if means.shape[0] else (torch.full((3,), float("inf"), device=means.device),
                                                                torch.full((3,), float("-inf"), device=means.device))
        lo, hi = lo.clone(), hi.clone()
        if self.world > 1:
            self._all_reduce(lo, dist.ReduceOp.MIN)
            self._all_reduce(hi, dist.ReduceOp.MAX)
        width = float((hi - lo).max().item())                    # fp32 subtraction, as V0 = V - vmin; width = V0.max()
        return lo, width

    def cell_owner(self, means, J, vmin, width, B):
        """(owner rank of every local Gaussian, owner table of the cells): cells by the top B bits of the Morton code, owners
        from the all-reduced histogram -- device arithmetic only, the same table on every rank."""
        ncell = 1 << B
        keys = self.be.keys(means, J, vmin, width)
        cell = keys >> (3 * J - B) if B else torch.zeros_like(keys)
        hist = self.be.histogram(keys, 3 * J - B if B else 63, ncell)
        if self.world > 1:
            self._all_reduce(hist, dist.ReduceOp.SUM)
        table = owner_table(hist, self.world)
        return table[cell], table

    def repartition(self, scene, owner):
        """Every Gaussian to the rank that owns its cell: ONE all_to_all of packed rows.  Returns (scene, bytes sent)."""
        if self.world == 1:
            return scene, 0
        cd = scene[4].shape[1]
        D = 11 + cd
        counts = self.be.histogram(owner, 0, self.world)
        all_counts = self._all_gather(counts).cpu()               # (P, P): [sender][receiver] -- the split sizes must reach the host
        send_sizes = all_counts[self.rank].tolist()
        recv_sizes = all_counts[:, self.rank].tolist()
        rows = self.be.pack_by_owner(scene, owner, self.world)    # grouped by owner, stable
        rows_in = self._exchange(rows, [c * D for c in send_sizes], [c * D for c in recv_sizes]).reshape(-1, D)
        return self.be.unpack_rows(rows_in, cd), int(rows.numel() * 4)

    def _cell_meta(self, keys, J, B):
        """Per cell of the scene, from this rank's sorted voxel keys: (count, first key, last key) and the row at which
        the cell starts here (meaningful where count > 0).  Device arithmetic only."""
        ncell = 1 << B
        dev = keys.device
        n = keys.shape[0]
        if n and B:                                                # the keys are sorted: a cell is a contiguous run of rows
            lower = torch.arange(ncell + 1, dtype=torch.int64, device=dev) << (3 * J - B)
            edges = torch.searchsorted(keys, lower)
            cnt = edges[1:] - edges[:-1]
        else:
            cnt = torch.full((ncell,), n, dtype=torch.int64, device=dev)
        start = torch.cumsum(cnt, 0) - cnt
        live = cnt > 0
        zero = torch.zeros_like(cnt)
        if n:
            first = torch.where(live, keys[torch.clamp(start, max=n - 1)], zero)
            last = torch.where(live, keys[torch.clamp(start + cnt - 1, min=0, max=n - 1)], zero)
        else:
            first, last = zero, zero.clone()
        return torch.stack([cnt, first, last], 1), start, live

    def _join_cells(self, T, start, live, meta, inverse):
        """The butterflies above the cells: every rank contributes the low-pass rows of its cells (first row of each),
        an all_reduce over the bit patterns (the owner's row plus zeros: exact, signed zeros included) hands every rank all
        of them, raht_shard_top_cells joins them, and the rank's own rows go back.  Returns the cells' levels."""
        ncell, D = live.shape[0], T.shape[1]
        dev = T.device
        n = T.shape[0]
        rows = torch.zeros(ncell, D, dtype=T.dtype, device=dev)
        if n:
            mine = torch.nonzero(live).reshape(-1)                # (a few hundred cells at most)
            rows[mine] = T[start[mine]]
        if self.world > 1:
            as_int = rows.view(torch.int64 if T.dtype == torch.float64 else torch.int32)
            self._all_reduce(as_int, dist.ReduceOp.SUM)
        rows, cell_level = self.be.shard_top_cells(meta[:, 1].contiguous(), meta[:, 2].contiguous(), meta[:, 0].contiguous(), rows, inverse)
        if n:
            T[start[mine]] = rows[mine]
        return cell_level

    # ---- encode ------------------------------------------------------------------------------------
    @torch.no_grad()
    def encode(self, means, quats, scales, opacities, colors, J, step, transform_dtype=torch.float64, repartition=True, timer=None):
        be, P, r = self.be, self.world, self.rank
        mark = timer.mark if timer is not None else (lambda name: None)
        mark("start")
        sent = {}
        scene = (means, quats, scales, opacities, colors)
        vmin, width = self.scene_bounds(means)
        mark("bounds_allreduce")
        B = cell_bits(P, J)
        if repartition and P > 1:
            owner, _ = self.cell_owner(means, J, vmin, width, B)
            mark("cell_histogram")
            scene, sent["repartition_alltoall"] = self.repartition(scene, owner)
        mark("repartition_alltoall")
        packed, keys = be.voxelize_merge(scene, J, vmin, width)
        mark("voxelize_merge")
        n, D = packed.shape
        dev = packed.device
        # the rank's cells as a forest: every butterfly below the cell roots; each cell's low-pass is left in its first row
        level = be.levels(keys, 3 * J - B) if n else torch.empty(0, dtype=torch.uint8, device=dev)
        T = be.forward_local(keys, level, packed, transform_dtype) if n else torch.empty(0, D, dtype=transform_dtype, device=dev)
        mark("raht_local")
        meta_local, start, live = self._cell_meta(keys, J, B)
        meta = meta_local.clone()
        if P > 1:
            self._all_reduce(meta, dist.ReduceOp.SUM)             # (every cell has one owner: the others add zeros)
        cell_level = self._join_cells(T, start, live, meta, False)
        sent["cells_allreduce"] = int(meta.numel() * 8 + (1 << B) * D * T.element_size())
        mark("top_allreduce_join")
        level_g = level.clone()
        if n:
            mine = torch.nonzero(live).reshape(-1)
            level_g[start[mine]] = cell_level[mine]               # 255 on the first cell of the scene: its root
        # order_RAGFT of the scene restricted to this rank, and where each of its groups sits in the whole sequence
        order = be.order(level_g) if n else torch.empty(0, dtype=torch.int64, device=dev)
        counts = torch.zeros(N_BUCKETS + 1, dtype=torch.int64, device=dev)
        if n:                                                      # (order lists the rows bucket by bucket: the counts are its run lengths)
            bk = order_bucket(level_g[order])
            edges = torch.searchsorted(bk, torch.arange(N_BUCKETS + 1, dtype=torch.int64, device=dev))
            counts[:N_BUCKETS] = edges[1:] - edges[:-1]
        counts[N_BUCKETS] = n
        gathered = self._all_gather(counts).cpu()                 # (P, 23): one small host read for the exchange sizes
        bucket_counts = gathered[:, :N_BUCKETS].contiguous()
        n_rows = gathered[:, N_BUCKETS].tolist()
        q = be.quant_reorder(T, order, step) if n else torch.empty(D, 0, dtype=torch.int32, device=dev)
        q = q[:, :n]
        mark("order_quant")
        # row-sharded -> channel-sharded: rank k gets channels k, k + P, ... of every rank
        chans = [list(range(k, D, P)) for k in range(P)]
        by_dest = torch.tensor([c for k in range(P) for c in chans[k]], dtype=torch.int64, device=dev)
        send = q[by_dest] if P > 1 else q.contiguous()           # (D, n): the channels grouped by destination, one gather
        recv = self._exchange(send, [len(chans[k]) * n for k in range(P)], [len(chans[r]) * n_rows[k] for k in range(P)])
        blocks, o = [], 0
        for k in range(P):
            m = len(chans[r]) * n_rows[k]
            blocks.append(recv[o:o + m].reshape(len(chans[r]), n_rows[k]))
            o += m
        sent["coeff_alltoall"] = int(q.numel() * 4)
        mark("coeff_alltoall")
        qg = interleave_buckets(blocks, bucket_counts)            # (|C_r|, N) in order_RAGFT order of the whole scene
        mark("interleave")
        streams = be.rlgr_encode(qg)
        mark("rlgr_encode")
        return ShardedEncoded(J=J, D=D, step=step, world=P, rank=r, n_rows=n_rows, keys=keys, cell_bits=B, cell_meta=meta,
                              cell_level=cell_level, bucket_counts=bucket_counts, channels=chans[r], streams=streams,
                              vmin=vmin, width=width, transform_dtype=transform_dtype, collective_bytes=sent)

    # ---- decode ------------------------------------------------------------------------------------
    @torch.no_grad()
    def decode(self, enc: ShardedEncoded, timer=None):
        """Returns the reconstructed attributes (n_local, D) of this rank's voxels (enc.keys)."""
        be, P, r = self.be, self.world, self.rank
        mark = timer.mark if timer is not None else (lambda name: None)
        mark("start_dec")
        D, n_rows, J, B = enc.D, enc.n_rows, enc.J, enc.cell_bits
        n = n_rows[r]
        dev = enc.keys.device
        chans = [list(range(k, D, P)) for k in range(P)]
        qg = be.rlgr_decode(enc.streams)                          # (|C_r|, N)
        mark("rlgr_decode")
        Cr = len(chans[r])
        flat = torch.empty(Cr * sum(n_rows), dtype=torch.int32, device=dev)
        blocks, o = [], 0
        for k in range(P):                                        # one send buffer, a view per destination
            blocks.append(flat[o:o + Cr * n_rows[k]].reshape(Cr, n_rows[k]))
            o += Cr * n_rows[k]
        deinterleave_buckets(qg, enc.bucket_counts, n_rows, out=blocks)
        mark("deinterleave")
        recv = self._exchange(flat, [Cr * n_rows[k] for k in range(P)], [len(chans[k]) * n for k in range(P)])
        mark("coeff_alltoall_dec")
        by_src = torch.tensor([c for k in range(P) for c in chans[k]], dtype=torch.int64, device=dev)
        q = torch.empty(D, n, dtype=torch.int32, device=dev)
        q[by_src] = recv.reshape(D, n)                            # (grouped by source rank k = channels k, k + P, ...)
        level = be.levels(enc.keys, 3 * J - B) if n else torch.empty(0, dtype=torch.uint8, device=dev)
        _, start, live = self._cell_meta(enc.keys, J, B)
        level_g = level.clone()
        if n:
            mine = torch.nonzero(live).reshape(-1)
            level_g[start[mine]] = enc.cell_level[mine]
        order = be.order(level_g) if n else torch.empty(0, dtype=torch.int64, device=dev)
        T = be.dequant_unreorder(q, order, enc.step, enc.transform_dtype) if n else torch.empty(0, D, dtype=enc.transform_dtype, device=dev)
        self._join_cells(T, start, live, enc.cell_meta, True)
        mark("dequant_top_dec")
        if n:
            T = be.inverse_local(enc.keys, level, T)
        mark("iraht_local")
        return T


def _segments(bucket_counts: torch.Tensor):
    """For every (bucket, rank) in sequence order: (rank, local start, global start, length)."""
    bc = bucket_counts.cpu().tolist()
    P = len(bc)
    local = [0] * P
    g = 0
    out = []
    for b in range(N_BUCKETS):
        for k in range(P):
            c = bc[k][b]
            if c:
                out.append((k, local[k], g, c))
            local[k] += c
            g += c
    return out, g


def interleave_buckets(blocks: List[torch.Tensor], bucket_counts: torch.Tensor) -> torch.Tensor:
    """blocks[k]: (C, n_k) coefficients of rank k in ITS order_RAGFT order -> (C, N) in the scene's order: inside
    every group (root, then octree levels coarse to fine) the rows of rank 0, then rank 1, ... (ascending row index)."""
    segs, N = _segments(bucket_counts)
    C = blocks[0].shape[0]
    out = torch.empty(C, N, dtype=blocks[0].dtype, device=blocks[0].device)
    for k, ls, gs, c in segs:
        out[:, gs:gs + c] = blocks[k][:, ls:ls + c]
    return out


def deinterleave_buckets(qg: torch.Tensor, bucket_counts: torch.Tensor, n_rows: List[int], out=None) -> List[torch.Tensor]:
    segs, _ = _segments(bucket_counts)
    C = qg.shape[0]
    blocks = out if out is not None else [torch.empty(C, n_rows[k], dtype=qg.dtype, device=qg.device) for k in range(len(n_rows))]
    for k, ls, gs, c in segs:
        blocks[k][:, ls:ls + c] = qg[:, gs:gs + c]
    return blocks

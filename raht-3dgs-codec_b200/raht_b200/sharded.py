"""Morton-prefix sharding of ONE scene over the GPUs of a box (BASELINE config 4).

The reference is single-GPU; this module keeps its arithmetic and its bitstream.  Rank r of P owns the
Gaussians whose voxel falls into the r-th range of Morton prefixes (a spatial box: octant halves for
P = 2, 4, 8), so that the concatenation of the ranks' Morton-sorted voxels in rank order IS the sorted voxel
list of the scene.  One process per GPU, `torch.distributed` (NCCL over NVLink) for the exchanges:

  encode   1 all_reduce   scene bounding box (6 floats)                                  [voxelize_pc.py:87-95]
           (optional)     all_to_all of the raw Gaussians to the rank that owns their prefix
           2 all_gather   per rank: rows, first / last Morton code, 22 order-bucket counts
           --             voxelise + merge + RAHT of the local range (every butterfly inside the range)
           3 all_gather   the P low-pass rows (D values each); every rank joins them (raht_shard_top)
           4 all_to_all   quantised coefficients, row-sharded -> channel-sharded in order_RAGFT order
           --             rank k entropy-codes channels k, k + P, ... of the WHOLE scene
  decode   the mirror image.

Result: coefficients, order and RLGR bytes are bit-identical to the single-GPU encode of the whole scene
(tests/test_sharded.py).  The compute goes through a small backend object: `DeviceBackend` (the CUDA
library) is the only one shipped; the CPU test-suite plugs its own checker backend in to exercise the host
logic (partition, boundary levels, exchanges) over gloo.
"""
from dataclasses import dataclass, field
from typing import List, Optional

import torch
import torch.distributed as dist

N_BUCKETS = 22           # order_RAGFT groups: root, then octree levels 20 .. 0 (tree.cu OR_BUCKETS)


def order_bucket(level: torch.Tensor) -> torch.Tensor:
    """Bucket of order_RAGFT for every row from its level (tree.cu or_bucket)."""
    lv = level.to(torch.int64)
    return torch.where(lv == 255, torch.zeros_like(lv), 1 + (20 - torch.clamp(lv, max=62) // 3))


def owner_of_key(keys: torch.Tensor, J: int, world: int) -> torch.Tensor:
    """Rank that owns a Morton code: equal ranges of the top ceil(log2 P) prefix bits."""
    bits = max((world - 1).bit_length(), 0)
    if bits == 0:
        return torch.zeros_like(keys)
    prefix = keys >> (3 * J - bits)
    return (prefix * world) >> bits


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

    def voxelize_merge(self, scene, J, vmin, width):
        packed, keys, _ = self.codec.voxelize_merge(*scene, J=J, given=(vmin, width))
        return packed, keys

    def levels(self, keys):
        return self.ops.tree_from_keys(keys).level

    def forward_local(self, keys, level, packed, dtype):
        tree = self.ops.RahtTree(level, level.shape[0])
        return self.ops.forward(tree, packed, dtype, self.tile_rows)

    def inverse_local(self, keys, level, T):
        tree = self.ops.RahtTree(level, level.shape[0])
        return self.ops.inverse(tree, T, self.tile_rows)

    def shard_top(self, first, prev_last, offsets, rows, inverse):
        return self.ops.shard_top(first, prev_last, offsets, rows, inverse)

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
    first_keys: torch.Tensor             # (P,) first code of every non-empty rank, prev_last_keys likewise
    prev_last_keys: torch.Tensor
    bucket_counts: torch.Tensor          # (P, 22)
    channels: List[int]                  # channels whose whole-scene stream this rank holds
    streams: object                      # the backend's stream object for those channels
    vmin: torch.Tensor = None
    width: float = 0.0
    transform_dtype: torch.dtype = torch.float64
    timings: dict = field(default_factory=dict)

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

    def _all_to_all(self, chunks: List[torch.Tensor], recv_sizes: List[int]) -> List[torch.Tensor]:
        """chunks[k] goes to rank k; returns what every rank sent here (1-D tensors)."""
        if self.world == 1:
            return [chunks[0].reshape(-1).clone()]
        send = torch.cat([c.reshape(-1) for c in chunks])
        recv = torch.empty(sum(recv_sizes), dtype=send.dtype, device=send.device)
        if dist.get_backend(self.group) == "gloo":              # host path of the tests: pairwise send / recv
            dev = chunks[0].device
            outs = [None] * self.world
            reqs = []
            for k in range(self.world):
                if k == self.rank:
                    outs[k] = chunks[k].reshape(-1).clone()
                    continue
                outs[k] = torch.empty(recv_sizes[k], dtype=send.dtype)
                reqs.append(dist.isend(chunks[k].reshape(-1).cpu().contiguous(), k, group=self.group))
                reqs.append(dist.irecv(outs[k], k, group=self.group))
            for q in reqs:
                q.wait()
            return [o.to(dev) for o in outs]
        dist.all_to_all_single(recv, send, output_split_sizes=recv_sizes, input_split_sizes=[c.numel() for c in chunks],
                               group=self.group)
        return list(torch.split(recv, recv_sizes))

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

    def repartition(self, scene, J, vmin, width):
        """Send every Gaussian to the rank that owns its Morton prefix (all_to_all of the raw rows)."""
        if self.world == 1:
            return scene
        means = scene[0]
        owner = owner_of_key(self.be.keys(means, J, vmin, width), J, self.world)
        perm = torch.argsort(owner, stable=True)
        counts = torch.bincount(owner, minlength=self.world)
        all_counts = self._all_gather(counts)                     # (P, P): [sender][receiver]
        send_sizes = counts.tolist()
        recv_sizes = all_counts[:, self.rank].tolist()
        out = []
        for a in scene:
            cols = 1 if a.dim() == 1 else a.shape[1]
            rows = a[perm].reshape(-1, cols)
            parts = list(torch.split(rows, send_sizes))
            got = self._all_to_all(parts, [r * cols for r in recv_sizes])
            cat = torch.cat(got).reshape(-1, cols)
            out.append(cat.reshape(-1) if a.dim() == 1 else cat)
        return tuple(out)

    # ---- encode ------------------------------------------------------------------------------------
    @torch.no_grad()
    def encode(self, means, quats, scales, opacities, colors, J, step, transform_dtype=torch.float64, repartition=True, timer=None):
        be, P, r = self.be, self.world, self.rank
        mark = timer.mark if timer is not None else (lambda name: None)
        mark("start")
        scene = (means, quats, scales, opacities, colors)
        vmin, width = self.scene_bounds(means)
        mark("bounds_allreduce")
        if repartition:
            scene = self.repartition(scene, J, vmin, width)
        mark("repartition_alltoall")
        packed, keys = be.voxelize_merge(scene, J, vmin, width)
        mark("voxelize_merge")
        n, D = packed.shape
        dev = packed.device
        if n and P > 1:
            own = owner_of_key(keys[[0, -1]], J, P)
            if int(own[0]) != r or int(own[1]) != r:
                raise ValueError("sharded encode: this rank holds Gaussians outside its Morton-prefix range (use repartition=True)")
        # per rank: rows, first and last code
        meta = torch.zeros(3, dtype=torch.int64, device=dev)
        meta[0] = n
        if n:
            meta[1], meta[2] = keys[0], keys[-1]
        allmeta = self._all_gather(meta)                          # (P, 3)
        n_rows = allmeta[:, 0].tolist()
        nonempty = [k for k in range(P) if n_rows[k] > 0]
        first = allmeta[nonempty, 1].contiguous()
        last = allmeta[nonempty, 2]
        prev_last = torch.cat([first[:1], last[:-1]]).contiguous() if nonempty else first
        offs = torch.zeros(len(nonempty) + 1, dtype=torch.int64, device=dev)
        offs[1:] = torch.cumsum(allmeta[nonempty, 0], 0)
        # local transform: every butterfly inside this range; the range's low-pass is left in its first row
        level = be.levels(keys) if n else torch.empty(0, dtype=torch.uint8, device=dev)
        T = be.forward_local(keys, level, packed, transform_dtype) if n else torch.empty(0, D, dtype=transform_dtype, device=dev)
        mark("raht_local")
        root = T[0].clone() if n else torch.zeros(D, dtype=transform_dtype, device=dev)
        roots = self._all_gather(root)[nonempty].contiguous()     # (P', D)
        roots, top_level = be.shard_top(first, prev_last, offs, roots, False)       # the butterflies that join the ranges
        mark("top_allgather_join")
        level_g = level.clone()
        if n:
            me = nonempty.index(r)
            T[0] = roots[me]
            level_g[0] = top_level[me]                            # 255 on the first range: the root of the scene
        # order_RAGFT of the scene restricted to this range, and where each of its groups sits in the whole sequence
        order = be.order(level_g) if n else torch.empty(0, dtype=torch.int64, device=dev)
        counts = torch.bincount(order_bucket(level_g), minlength=N_BUCKETS) if n else torch.zeros(N_BUCKETS, dtype=torch.int64, device=dev)
        bucket_counts = self._all_gather(counts)                  # (P, 22)
        q = be.quant_reorder(T, order, step) if n else torch.empty(D, 0, dtype=torch.int32, device=dev)
        q = q[:, :n].contiguous()                                 # (D, n) channel-major, local bucket order
        mark("order_quant")
        # row-sharded -> channel-sharded: rank k gets channels k, k + P, ... of every rank
        chans = [list(range(k, D, P)) for k in range(P)]
        parts = [q[chans[k]] for k in range(P)]
        got = self._all_to_all(parts, [len(chans[r]) * n_rows[k] for k in range(P)])
        blocks = [got[k].reshape(len(chans[r]), n_rows[k]) for k in range(P)]
        mark("coeff_alltoall")
        qg = interleave_buckets(blocks, bucket_counts)            # (|C_r|, N) in order_RAGFT order of the whole scene
        mark("interleave")
        streams = be.rlgr_encode(qg)
        mark("rlgr_encode")
        return ShardedEncoded(J=J, D=D, step=step, world=P, rank=r, n_rows=n_rows, keys=keys, first_keys=first,
                              prev_last_keys=prev_last, bucket_counts=bucket_counts, channels=chans[r], streams=streams,
                              vmin=vmin, width=width, transform_dtype=transform_dtype)

    # ---- decode ------------------------------------------------------------------------------------
    @torch.no_grad()
    def decode(self, enc: ShardedEncoded, timer=None):
        """Returns the reconstructed attributes (n_local, D) of this rank's voxels (enc.keys)."""
        be, P, r = self.be, self.world, self.rank
        mark = timer.mark if timer is not None else (lambda name: None)
        mark("start_dec")
        D, n_rows = enc.D, enc.n_rows
        n = n_rows[r]
        dev = enc.keys.device
        chans = [list(range(k, D, P)) for k in range(P)]
        qg = be.rlgr_decode(enc.streams)                          # (|C_r|, N)
        mark("rlgr_decode")
        blocks = deinterleave_buckets(qg, enc.bucket_counts, n_rows)
        mark("deinterleave")
        got = self._all_to_all([b.contiguous() for b in blocks], [len(chans[k]) * n for k in range(P)])
        mark("coeff_alltoall_dec")
        q = torch.empty(D, n, dtype=torch.int32, device=dev)
        for k in range(P):
            q[chans[k]] = got[k].reshape(len(chans[k]), n)
        nonempty = [k for k in range(P) if n_rows[k] > 0]
        offs = torch.zeros(len(nonempty) + 1, dtype=torch.int64, device=dev)
        offs[1:] = torch.cumsum(torch.tensor([n_rows[k] for k in nonempty], dtype=torch.int64, device=dev), 0)
        level = be.levels(enc.keys) if n else torch.empty(0, dtype=torch.uint8, device=dev)
        # the levels of the ranges' first rows in the whole scene (no data needed: P codes)
        _, top_level = be.shard_top(enc.first_keys, enc.prev_last_keys, offs,
                                    torch.zeros(len(nonempty), D, dtype=enc.transform_dtype, device=dev), True)
        level_g = level.clone()
        if n:
            me = nonempty.index(r)
            level_g[0] = top_level[me]
        order = be.order(level_g) if n else torch.empty(0, dtype=torch.int64, device=dev)
        T = be.dequant_unreorder(q, order, enc.step, enc.transform_dtype) if n else torch.empty(0, D, dtype=enc.transform_dtype, device=dev)
        root = T[0].clone() if n else torch.zeros(D, dtype=enc.transform_dtype, device=dev)
        roots = self._all_gather(root)[nonempty].contiguous()
        roots, _ = be.shard_top(enc.first_keys, enc.prev_last_keys, offs, roots, True)
        mark("dequant_top_dec")
        if n:
            T[0] = roots[me]
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


def deinterleave_buckets(qg: torch.Tensor, bucket_counts: torch.Tensor, n_rows: List[int]) -> List[torch.Tensor]:
    segs, _ = _segments(bucket_counts)
    C = qg.shape[0]
    blocks = [torch.empty(C, n_rows[k], dtype=qg.dtype, device=qg.device) for k in range(len(n_rows))]
    for k, ls, gs, c in segs:
        blocks[k][:, ls:ls + c] = qg[:, gs:gs + c]
    return blocks

"""Morton-prefix sharding of one scene (raht_b200/sharded.py): P ranks must produce the
coefficients and the RLGR bytes of the single-process encode of the whole scene, bit for bit.

CPU (not gpu): world_size 2, 3 and 5 over gloo with the ORACLE plugged in as the compute backend -- checks the
host logic (partition by prefix, boundary levels, the butterflies that join the ranges, order_RAGFT across
ranks, the row-sharded -> channel-sharded exchange and its mirror).
GPU: the same with the CUDA library as backend, 2 / 3 / 8 processes sharing cuda:0 (gloo, staged through the host;
on a multi-GPU box the production path is NCCL, see bench.py --workload c4)."""
import os
import sys
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers
from oracle import oracle_ctypes as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "raht-3dgs-codec_b200")
J, STEP, N_SCENE = 8, 0.05, 16000          # defaults of the CPU runs; the GPU run also takes a BASELINE-sized scene


def _msb(x):
    return int(x).bit_length() - 1


class OracleBackend:
    """CPU restatement (oracle/) behind the backend interface of sharded.py -- TEST ONLY."""

    def bounds(self, means):
        return means.min(0).values, means.max(0).values

    def keys(self, means, J, vmin, width):
        vs = np.float32(width / 2 ** J)
        v0 = means.numpy() - vmin.numpy().astype(np.float32)
        q = np.clip(np.floor(v0 / vs).astype(np.int64), 0, 2 ** J - 1)
        return torch.from_numpy(orc.morton(q, J).astype(np.int64))

    def voxelize_merge(self, scene, J, vmin, width):
        self.J = J
        means, quats, scales, opac, colors = [a.numpy() for a in scene]
        n = means.shape[0]
        if n == 0:
            return torch.empty(0, 11 + colors.shape[1]), torch.empty(0, dtype=torch.int64)
        r = orc.voxelize(means, J, vmin=vmin.numpy().astype(np.float32), width=float(width), want_delta=False)
        coff = np.concatenate([r["voxel_indices"], [n]]).astype(np.int32)
        mm, mq, ms, mo, mc = orc.merge_clusters(r["sort_idx"].astype(np.int32), coff, means, quats, scales, opac, colors, True)
        packed = np.concatenate([mm, mq, ms, mo[:, None], mc], 1).astype(np.float32)
        keys = r["MC_sorted"][r["voxel_indices"]].astype(np.int64)
        return torch.from_numpy(packed), torch.from_numpy(keys)

    def histogram(self, values, shift, nbins):
        return torch.from_numpy(np.bincount(np.minimum(values.numpy() >> shift, nbins - 1), minlength=nbins).astype(np.int64))

    def pack_by_owner(self, scene, owner, world):
        means, quats, scales, opac, colors = [a.numpy() for a in scene]
        rows = np.concatenate([means, quats, scales, opac[:, None], colors], 1).astype(np.float32)
        return torch.from_numpy(rows[np.argsort(owner.numpy(), kind="stable")])

    def unpack_rows(self, rows, color_dim):
        r = rows.numpy()
        return tuple(torch.from_numpy(np.ascontiguousarray(x)) for x in (r[:, 0:3], r[:, 3:7], r[:, 7:10], r[:, 10], r[:, 11:11 + color_dim]))

    def levels(self, keys, cut):
        k = keys.numpy()
        lev = np.full(k.shape[0], 255, dtype=np.uint8)
        lev[1:] = [_msb(a ^ b) for a, b in zip(k[:-1], k[1:])]
        lev[lev >= cut] = 255                                   # every cell a tree of its own
        return torch.from_numpy(lev)

    def _forest(self, keys, level):
        roots = np.flatnonzero(level.numpy() == 255).tolist() + [keys.shape[0]]
        return [(a, b) for a, b in zip(roots[:-1], roots[1:])]

    def forward_local(self, keys, level, packed, dtype):
        T = np.empty(packed.shape, dtype=np.float64)
        for a, b in self._forest(keys, level):
            T[a:b], _ = orc.Tree(keys.numpy()[a:b], self.J).forward(packed.numpy()[a:b].astype(np.float64))
        return torch.from_numpy(T)

    def inverse_local(self, keys, level, T):
        C = np.empty(T.shape, dtype=np.float64)
        for a, b in self._forest(keys, level):
            C[a:b] = orc.Tree(keys.numpy()[a:b], self.J).inverse(T.numpy()[a:b])
        return torch.from_numpy(C)

    def shard_top_cells(self, first, last, counts, rows, inverse):
        f, la, cnt, R = first.numpy(), last.numpy(), counts.numpy(), rows.numpy().copy()
        live = [i for i in range(len(cnt)) if cnt[i] > 0]
        off = np.concatenate([[0], np.cumsum(cnt)])
        lev = np.zeros(len(cnt), dtype=np.uint8)
        for n, i in enumerate(live):
            lev[i] = 255 if n == 0 else _msb(int(la[live[n - 1]]) ^ int(f[i]))
        order = sorted(live[1:], key=lambda i: (lev[i], i))
        for i in (reversed(order) if inverse else order):
            s = max(j for j in live if j < i and lev[j] > lev[i])
            t = min([j for j in live if j > i and lev[j] > lev[i]] + [len(cnt)])
            w0, w1 = np.float64(off[i] - off[s]), np.float64(off[t] - off[i])
            a, b = np.sqrt(w0 / (w0 + w1)), np.sqrt(w1 / (w0 + w1))
            x0, x1 = R[s].copy(), R[i].copy()
            if inverse:
                R[s], R[i] = a * x0 - b * x1, b * x0 + a * x1
            else:
                R[s], R[i] = a * x0 + b * x1, (-b) * x0 + a * x1
        return torch.from_numpy(R), torch.from_numpy(lev)

    def order(self, level_global):
        from raht_b200.sharded import order_bucket
        return torch.from_numpy(np.argsort(order_bucket(level_global).numpy(), kind="stable"))

    def quant_reorder(self, T, order, step):
        return torch.from_numpy(np.ascontiguousarray(orc.quantize(T.numpy(), step)[order.numpy()].T.astype(np.int32)))

    def dequant_unreorder(self, q, order, step, dtype):
        T = np.empty((q.shape[1], q.shape[0]), dtype=np.float64)
        T[order.numpy()] = orc.dequantize(q.numpy().T.astype(np.float64), step)
        return torch.from_numpy(T)

    def rlgr_encode(self, q):
        q = np.ascontiguousarray(q.numpy())
        return orc.rlgr_encode_channels(q, 1) + (q.shape[1], q.shape[0])

    def rlgr_decode(self, streams):
        _, out, cap, nb, n, nch = streams
        return torch.from_numpy(orc.rlgr_decode_channels(out, cap, nb, n, nch, 1).astype(np.int32))

    def stream_bytes(self, streams):
        return list(streams[0])


def reference_encode(scene, J=J):
    """Single-process oracle encode + decode of the whole scene -> (keys, streams per channel, decoded attributes)."""
    means, quats, scales, opac, colors = scene
    n = means.shape[0]
    r = orc.voxelize(means, J, want_delta=False)
    coff = np.concatenate([r["voxel_indices"], [n]]).astype(np.int32)
    mm, mq, ms, mo, mc = orc.merge_clusters(r["sort_idx"].astype(np.int32), coff, means, quats, scales, opac, colors, True)
    packed = np.concatenate([mm, mq, ms, mo[:, None], mc], 1).astype(np.float32).astype(np.float64)
    keys = r["MC_sorted"][r["voxel_indices"]]
    t = orc.Tree(keys, J)
    T, _ = t.forward(packed)
    order = t.order()
    q = orc.quantize(T, STEP)[order].T.astype(np.int32)
    streams, _, _, _ = orc.rlgr_encode_channels(np.ascontiguousarray(q), 1)
    Tdec = orc.dequantize(q.T.astype(np.float64), STEP)[np.argsort(order, kind="stable")]
    return keys.astype(np.int64), list(streams), t.inverse(Tdec)


def _worker(rank, world, initfile, use_cuda, result_dir, n_scene=N_SCENE, J=J):
    for p in (ROOT, PKG, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", init_method=f"file://{initfile}", rank=rank, world_size=world)
    try:
        from raht_b200 import sharded
        scene = helpers.synthetic_gaussians(n_scene, 77)
        dev = "cuda:0" if use_cuda else "cpu"
        mine = [torch.from_numpy(np.ascontiguousarray(a[rank::world])).to(dev) for a in scene]      # arbitrary split: repartition moves the rows
        backend = sharded.DeviceBackend() if use_cuda else OracleBackend()
        codec = sharded.ShardedCodec(backend)
        enc = codec.encode(*mine, J=J, step=STEP)
        rec = codec.decode(enc)
        if isinstance(backend, OracleBackend):
            backend.J = J
        out = {"channels": enc.channels, "streams": backend.stream_bytes(enc.streams), "keys": enc.keys.cpu().numpy(),
               "rec": rec.cpu().numpy(), "n_rows": enc.n_rows, "sent": enc.collective_bytes}
        torch.save(out, os.path.join(result_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _run(world, use_cuda, n_scene=N_SCENE, J=J):
    with tempfile.TemporaryDirectory() as d:
        initfile = os.path.join(d, "init")
        mp.spawn(_worker, args=(world, initfile, use_cuda, d, n_scene, J), nprocs=world, join=True)
        outs = [torch.load(os.path.join(d, f"rank{r}.pt"), weights_only=False) for r in range(world)]
    scene = helpers.synthetic_gaussians(n_scene, 77)
    keys, streams, rec = reference_encode(scene, J)
    rows = [len(o["keys"]) for o in outs]
    assert max(rows) <= 1.35 * len(keys) / world + 64, f"unbalanced ranks: {rows}"           # balanced by the cell histogram
    assert sum(outs[0]["n_rows"]) == len(keys)
    assert np.array_equal(np.concatenate([o["keys"] for o in outs]), keys)              # rank order = Morton order of the scene
    seen = set()
    for o in outs:
        for c, s in zip(o["channels"], o["streams"]):
            assert s == streams[c], f"channel {c}: sharded RLGR stream differs from the single-process stream"
            seen.add(c)
    assert seen == set(range(len(streams)))
    assert np.array_equal(np.concatenate([o["rec"] for o in outs]), rec)                # decoded attributes, bit for bit


@pytest.mark.parametrize("world", [2, 3, 5])
def test_sharded_encode_matches_single_process_gloo_cpu(world):
    """World sizes that are not powers of two included (5: the first version gave rank 3 the prefixes 101 and 110, which are
    not siblings -- every rank's range is now a union of whole cells, each transformed as its own tree)."""
    _run(world, use_cuda=False)


def test_interleave_roundtrip():
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    from raht_b200 import sharded
    rng = np.random.default_rng(0)
    counts = torch.from_numpy(rng.integers(0, 5, size=(3, sharded.N_BUCKETS)))
    n_rows = counts.sum(1).tolist()
    blocks = [torch.from_numpy(rng.integers(-9, 9, size=(4, n))) for n in n_rows]
    g = sharded.interleave_buckets(blocks, counts)
    back = sharded.deinterleave_buckets(g, counts, n_rows)
    assert all(torch.equal(a, b) for a, b in zip(blocks, back))
    assert g.shape == (4, sum(n_rows))


@pytest.mark.gpu
@pytest.mark.parametrize("world,n_scene,J", [(2, N_SCENE, J), (3, 60000, 10), (8, 50000, 9), (2, 2_000_000, 16)])
def test_sharded_encode_matches_single_process_cuda(world, n_scene, J):
    """The CUDA library behind the sharded codec, `world` processes sharing cuda:0 (gloo, staged through the host), up to a
    2 M-Gaussian scene at depth 16: keys, every RLGR stream and the decoded attributes equal the single-process oracle's."""
    _run(world, use_cuda=True, n_scene=n_scene, J=J)


@pytest.mark.gpu
@pytest.mark.parametrize("ncell", [3, 4, 8, 64, 700])
def test_shard_top_cells_vs_restatement(ncell):
    """raht_shard_top_cells (the butterflies that join the cells, empty cells skipped) against the numpy restatement the
    CPU tests use, forward and inverse, bit for bit."""
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    from raht_b200 import sharded
    rng = np.random.default_rng(ncell)
    Jc, D = 12, 11
    counts = rng.integers(0, 50, size=ncell).astype(np.int64)
    counts[rng.random(ncell) < 0.3] = 0
    if counts.sum() == 0:
        counts[ncell // 2] = 7
    bits = max((ncell - 1).bit_length(), 1)
    first = np.zeros(ncell, dtype=np.int64)
    last = np.zeros(ncell, dtype=np.int64)
    for c in range(ncell):                                    # codes inside cell c of a 3 Jc-bit space
        lo = c << (3 * Jc - bits)
        ks = np.sort(rng.choice(1 << (3 * Jc - bits), size=max(int(counts[c]), 1), replace=False)) + lo
        first[c], last[c] = ks[0], ks[-1]
    rows = rng.normal(size=(ncell, D))
    rows[counts == 0] = 0.0
    be, ob = sharded.DeviceBackend(), OracleBackend()
    for inverse in (False, True):
        exp, lev = ob.shard_top_cells(torch.from_numpy(first), torch.from_numpy(last), torch.from_numpy(counts), torch.from_numpy(rows.copy()), inverse)
        got, glev = be.shard_top_cells(torch.from_numpy(first).cuda(), torch.from_numpy(last).cuda(), torch.from_numpy(counts).cuda(),
                                       torch.from_numpy(rows.copy()).cuda(), inverse)
        assert np.array_equal(got.cpu().numpy(), exp.numpy()) and np.array_equal(glev.cpu().numpy(), lev.numpy())

"""Morton-prefix sharding of one scene (raht_b200/sharded.py): P ranks must produce the
coefficients and the RLGR bytes of the single-process encode of the whole scene, bit for bit.

CPU (not gpu): world_size 2 and 3 over gloo with the ORACLE plugged in as the compute backend -- checks the
host logic (partition by prefix, boundary levels, the butterflies that join the ranges, order_RAGFT across
ranks, the row-sharded -> channel-sharded exchange and its mirror).
GPU: the same with the CUDA library as backend, two processes sharing cuda:0 (gloo, staged through the host;
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
J, STEP, N_SCENE = 8, 0.05, 24000


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

    def levels(self, keys):
        k = keys.numpy()
        lev = np.full(k.shape[0], 255, dtype=np.uint8)
        lev[1:] = [_msb(a ^ b) for a, b in zip(k[:-1], k[1:])]
        return torch.from_numpy(lev)

    def forward_local(self, keys, level, packed, dtype):
        T, _ = orc.Tree(keys.numpy(), self.J).forward(packed.numpy().astype(np.float64))
        return torch.from_numpy(T)

    def inverse_local(self, keys, level, T):
        return torch.from_numpy(orc.Tree(keys.numpy(), self.J).inverse(T.numpy()))

    def shard_top(self, first, prev_last, offsets, rows, inverse):
        f, pl, off, R = first.numpy(), prev_last.numpy(), offsets.numpy(), rows.numpy().copy()
        P = len(f)
        lev = [255] + [_msb(pl[i] ^ f[i]) for i in range(1, P)]
        order = sorted(range(1, P), key=lambda i: (lev[i], i))
        for i in (reversed(order) if inverse else order):
            s = max(j for j in range(i) if lev[j] > lev[i])
            t = min([j for j in range(i + 1, P) if lev[j] > lev[i]] + [P])
            w0, w1 = np.float64(off[i] - off[s]), np.float64(off[t] - off[i])
            a, b = np.sqrt(w0 / (w0 + w1)), np.sqrt(w1 / (w0 + w1))
            x0, x1 = R[s].copy(), R[i].copy()
            if inverse:
                R[s], R[i] = a * x0 - b * x1, b * x0 + a * x1
            else:
                R[s], R[i] = a * x0 + b * x1, (-b) * x0 + a * x1
        return torch.from_numpy(R), torch.tensor(lev, dtype=torch.uint8)

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


def reference_encode(scene):
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


def _worker(rank, world, initfile, use_cuda, result_dir):
    for p in (ROOT, PKG, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", init_method=f"file://{initfile}", rank=rank, world_size=world)
    try:
        from raht_b200 import sharded
        scene = helpers.synthetic_gaussians(N_SCENE, 77)
        dev = "cuda:0" if use_cuda else "cpu"
        mine = [torch.from_numpy(np.ascontiguousarray(a[rank::world])).to(dev) for a in scene]      # arbitrary split: repartition moves the rows
        backend = sharded.DeviceBackend() if use_cuda else OracleBackend()
        codec = sharded.ShardedCodec(backend)
        enc = codec.encode(*mine, J=J, step=STEP)
        rec = codec.decode(enc)
        out = {"channels": enc.channels, "streams": backend.stream_bytes(enc.streams), "keys": enc.keys.cpu().numpy(),
               "rec": rec.cpu().numpy(), "n_rows": enc.n_rows}
        torch.save(out, os.path.join(result_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _run(world, use_cuda):
    with tempfile.TemporaryDirectory() as d:
        initfile = os.path.join(d, "init")
        mp.spawn(_worker, args=(world, initfile, use_cuda, d), nprocs=world, join=True)
        outs = [torch.load(os.path.join(d, f"rank{r}.pt"), weights_only=False) for r in range(world)]
    scene = helpers.synthetic_gaussians(N_SCENE, 77)
    keys, streams, rec = reference_encode(scene)
    assert sum(outs[0]["n_rows"]) == len(keys)
    assert np.array_equal(np.concatenate([o["keys"] for o in outs]), keys)              # rank order = Morton order of the scene
    seen = set()
    for o in outs:
        for c, s in zip(o["channels"], o["streams"]):
            assert s == streams[c], f"channel {c}: sharded RLGR stream differs from the single-process stream"
            seen.add(c)
    assert seen == set(range(len(streams)))
    assert np.array_equal(np.concatenate([o["rec"] for o in outs]), rec)                # decoded attributes, bit for bit


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_encode_matches_single_process_gloo_cpu(world):
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
def test_sharded_encode_matches_single_process_cuda():
    _run(2, use_cuda=True)

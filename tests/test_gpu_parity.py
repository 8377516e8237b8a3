"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI / the reference-named Python
shims, against (i) the golden fixtures generated from the reference and (ii) the CPU oracle on seeded
inputs.  Integer / index / byte results and fp64 coefficients must be bit-exact."""
import numpy as np
import pytest
import torch

import helpers
from oracle import oracle_ctypes as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def cu(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    return t if dtype is None else t.to(dtype)


def test_pinning_assumption_cuda_sqrt_is_ieee():
    """The fixtures assume torch on CUDA (the reference's production device) rounds sqrt correctly."""
    x = np.random.default_rng(0).random(200000)
    assert np.array_equal(torch.sqrt(cu(x)).cpu().numpy(), np.sqrt(x))
    a = np.random.default_rng(1).integers(1, 10**6, size=200000).astype(np.float64)
    b = np.random.default_rng(2).integers(1, 10**6, size=200000).astype(np.float64)
    assert np.array_equal((cu(a) / (cu(a) + cu(b))).cpu().numpy(), a / (a + b))


# ---------------------------------------------------------------------------------------------------
# tree + transform + quantisation + RLGR against the golden fixtures (reference outputs)
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", helpers.RAHT_CASES)
@pytest.mark.parametrize("tile_rows", [128, 256])
def test_golden_raht_pipeline(name, tile_rows):
    import RAHT_param as P
    import RAHT as R
    import iRAHT as IR
    from raht_b200 import ops
    g = helpers.load(name)
    J = int(g["J"])
    V = cu(g["V"], torch.float64)
    origin = torch.zeros(3, dtype=torch.float64, device=DEV)
    List, Flags, weights, order = P.RAHT_param_reorder_fast(V, origin, 2 ** J, J)
    levels = helpers.split_levels(g)
    assert len(List) == len(Flags) == len(weights) == int(g["nlevels"])
    for l, (lst, fl, w) in enumerate(levels):
        assert List[l].dtype == torch.int64 and Flags[l].dtype == torch.bool and weights[l].dtype == torch.int64
        assert np.array_equal(List[l].cpu().numpy(), lst), f"List[{l}]"
        assert np.array_equal(Flags[l].cpu().numpy(), fl), f"Flags[{l}]"
        assert np.array_equal(weights[l].cpu().numpy(), w), f"weights[{l}]"
    assert np.array_equal(order.cpu().numpy(), g["order"])

    tree = ops.tree_of(List)
    C = cu(g["C"])
    T = ops.forward(tree, C, torch.float64, tile_rows)
    assert np.array_equal(T.cpu().numpy(), g["T"]), "T must be bit-exact with RAHT2_optimized"
    T2, w = R.RAHT2_optimized(C, List, Flags, weights)
    assert np.array_equal(T2.cpu().numpy(), g["T"]) and w.shape == (C.shape[0], 1)
    assert np.array_equal(w.cpu().numpy().reshape(-1), g["w"])
    # TMA and plain-load paths agree
    assert torch.equal(ops.forward(tree, C, torch.float64, tile_rows, flags=1), T)
    Cl = ops.inverse(tree, T, tile_rows)
    assert np.array_equal(Cl.cpu().numpy(), g["C_lossless"])
    assert torch.equal(ops.inverse(tree, T, tile_rows, flags=1), Cl)
    assert np.array_equal(IR.inverse_RAHT_optimized(T, List, Flags, weights).cpu().numpy(), g["C_lossless"])
    # fp32 input is widened exactly
    T32in = ops.forward(tree, C.float(), torch.float64, tile_rows)
    assert np.array_equal(T32in.cpu().numpy(), orc.Tree(helpers.morton_np(g["V"]), J).forward(g["C"].astype(np.float32).astype(np.float64))[0])

    for s in g["steps"]:
        s = int(s)
        q = ops.quant_reorder(T, order, s)
        assert np.array_equal(q.cpu().numpy(), g[f"q_step{s}"])
        streams = ops.rlgr_encode(q, 1)
        got = streams.to_bytes()
        exp = helpers.split_streams(g, s)
        for ch in range(len(exp)):
            assert got[ch] == exp[ch], f"step {s}: RLGR bytes of channel {ch} differ"
        qd = ops.rlgr_decode(streams, 1)
        assert torch.equal(qd, q)
        assert torch.equal(ops.rlgr_decode(streams, 1, use_index=False), q)      # channel-serial path (no sync index)
        Td = ops.dequant_unreorder(qd, order, s)
        Crec = ops.inverse(tree, Td, tile_rows)
        assert np.array_equal(Crec.cpu().numpy(), g[f"C_rec_step{s}"])


def test_plain_lists_are_accepted():
    """encode_3dgs.py:152-154 rebuilds the lists with [t.to(device) for t in List]; also lists that did
    not come from this package (here: from the golden file) must work."""
    import RAHT_param as P
    import RAHT as R
    g = helpers.load("raht_clustered_J7_D5")
    J = int(g["J"])
    V = cu(g["V"], torch.float64)
    List, Flags, weights, _ = P.RAHT_param_reorder_fast(V, torch.zeros(3, device=DEV, dtype=torch.float64), 2 ** J, J)
    L2 = [t.to(device=DEV, non_blocking=True) for t in List]
    F2 = [t.to(device=DEV, non_blocking=True) for t in Flags]
    W2 = [t.to(device=DEV, non_blocking=True) for t in weights]
    T, _ = R.RAHT2_optimized(cu(g["C"]), L2, F2, W2)
    assert np.array_equal(T.cpu().numpy(), g["T"])
    lv = helpers.split_levels(g)
    L3, F3, W3 = [cu(a) for a, _, _ in lv], [cu(b) for _, b, _ in lv], [cu(c) for _, _, c in lv]
    T3, w3 = R.RAHT2_optimized(cu(g["C"]), L3, F3, W3)
    assert np.array_equal(T3.cpu().numpy(), g["T"]) and np.array_equal(w3.cpu().numpy().reshape(-1), g["w"])
    L1 = [a + 1 for a in L3]
    T4, _ = R.RAHT2_optimized(cu(g["C"]), L1, F3, W3, one_based=True)
    assert np.array_equal(T4.cpu().numpy(), g["T"])


def test_raht_param_errors():
    import RAHT_param as P
    V = torch.tensor([[0, 0, 0], [0, 0, 1], [4, 0, 0]], dtype=torch.float64, device=DEV)
    with pytest.raises(ValueError, match="OutOfBounds"):
        P.RAHT_param(V, torch.zeros(3, device=DEV), 4, 2)
    V = torch.tensor([[0, 0, 1], [0, 0, 0]], dtype=torch.float64, device=DEV)
    with pytest.raises(ValueError, match="sorted"):
        P.RAHT_param_reorder_fast(V, torch.zeros(3, device=DEV), 4, 2)
    V = torch.tensor([[0, 0, 1], [0, 0, 1]], dtype=torch.float64, device=DEV)
    with pytest.raises(ValueError):
        P.RAHT_param_reorder_fast(V, torch.zeros(3, device=DEV), 4, 2)
    # N == 1 (RAHT_param.py:233-236): one level, order None
    V = torch.tensor([[1, 2, 3]], dtype=torch.float64, device=DEV)
    Ls, Fs, Ws, order = P.RAHT_param_reorder_fast(V, torch.zeros(3, device=DEV), 4, 2)
    assert len(Ls) == 1 and order is None and Ws[0].tolist() == [1] and Fs[0].tolist() == [False]
    import RAHT as R
    T, w = R.RAHT2_optimized(torch.tensor([[5.0, 6.0]], device=DEV), Ls, Fs, Ws)
    assert T.tolist() == [[5.0, 6.0]] and w.tolist() == [[1.0]]


# ---------------------------------------------------------------------------------------------------
# the same pipeline against the CPU oracle at sizes the oracle finishes in seconds
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,J,D,clustered,seed", [(60000, 10, 59, True, 1), (150000, 12, 59, False, 2), (300000, 16, 3, True, 3),
                                                  (200001, 21, 8, False, 4), (129, 4, 59, False, 5), (2, 3, 4, False, 6)])
def test_oracle_raht_pipeline(n, J, D, clustered, seed):
    from raht_b200 import ops
    import RAHT_param as P
    Pts, MC = helpers.unique_sorted_voxels(n, J, seed, clustered)
    N = len(MC)
    rng = np.random.default_rng(seed)
    C = (rng.normal(size=(N, D)) * np.linspace(0.5, 40, D)[None, :]).astype(np.float64)
    t = orc.Tree(MC, J)
    List, Flags, weights, order = P.RAHT_param_reorder_fast(cu(Pts, torch.float64), torch.zeros(3, device=DEV), 2 ** J, J)
    assert len(List) == t.nlevels
    for l in sorted(set([0, 1, t.nlevels // 2, t.nlevels - 1])):
        a, b, c = t.level(l)
        assert np.array_equal(List[l].cpu().numpy(), a) and np.array_equal(Flags[l].cpu().numpy(), b)
        assert np.array_equal(weights[l].cpu().numpy(), c)
    o_ref = t.order()
    if N > 4:
        assert np.array_equal(order.cpu().numpy(), o_ref)
    tree = ops.tree_of(List)
    To, wo = t.forward(C)
    for R_ in (128, 256 if D > 30 else 512):      # 512 rows x 59 fp64 channels exceed 227 KB of shared memory
        T = ops.forward(tree, cu(C), torch.float64, R_)
        assert np.array_equal(T.cpu().numpy(), To), f"tile_rows={R_}"
        assert np.array_equal(tree.w.cpu().numpy(), wo)
        assert np.array_equal(ops.inverse(tree, T, R_).cpu().numpy(), t.inverse(To))
    info = tree.plan_info(128)
    assert info["stage_tiles"][0] == (N + 127) // 128 and info["stage_tiles"][-1] == 1 and info["stages"] >= 1
    for s in (1, 12):
        q = ops.quant_reorder(T, order, s)
        q_ref = orc.quantize(To, s)[order.cpu().numpy()].T.astype(np.int32)
        assert np.array_equal(q.cpu().numpy(), q_ref)
        streams = ops.rlgr_encode(q, 1)
        got = streams.to_bytes()
        exp, _, _, _ = orc.rlgr_encode_channels(q_ref, 1)
        assert got == exp
        assert torch.equal(ops.rlgr_decode(streams, 1), q)
        assert torch.equal(ops.rlgr_decode(streams, 1, use_index=False), q)


@pytest.mark.parametrize("n,J,seed", [(700, 6, 31), (3000, 6, 32), (9000, 8, 33)])
def test_transform_calls_back_to_back_on_small_scenes(n, J, seed):
    """The stages of a transform are programmatic dependent launches: the kernels of a call start while the
    kernel before them drains, and some read their inputs ahead of their wait.  On small scenes every grid of the chain
    fits the GPU at once, so a missing dependency shows: forward -> inverse -> forward queued without a host
    synchronisation, many times over the same buffers, must give the oracle's bits every time."""
    from raht_b200 import ops
    Pts, MC = helpers.unique_sorted_voxels(n, J, seed, False)
    N = len(MC)
    C = np.random.default_rng(seed).normal(size=(N, 59)).astype(np.float64) * 7.0
    t = orc.Tree(MC, J)
    To, _ = t.forward(C)
    Co = t.inverse(To)
    tree = ops.tree_from_keys(cu(MC))
    Cd = cu(C)
    out = []
    for _ in range(25):
        T = ops.forward(tree, Cd, torch.float64)
        R = ops.inverse(tree, T)
        T2 = ops.forward(tree, R, torch.float64)
        out.append((T, R, T2))
    torch.cuda.synchronize()
    To2, _ = t.forward(Co)
    for T, R, T2 in out:
        assert np.array_equal(T.cpu().numpy(), To) and np.array_equal(R.cpu().numpy(), Co) and np.array_equal(T2.cpu().numpy(), To2)


def test_fp32_mode_tolerance():
    """Build-specific fp32 transform: <= 1e-5 relative error on coefficients, PSNR within 0.01 dB."""
    from raht_b200 import ops, codec
    Pts, MC = helpers.unique_sorted_voxels(100000, 12, 21, True)
    N = len(MC)
    C = np.random.default_rng(21).normal(size=(N, 59)).astype(np.float32)
    tree = ops.tree_from_keys(cu(MC))
    T64 = ops.forward(tree, cu(C), torch.float64)
    T32 = ops.forward(tree, cu(C), torch.float32)
    scale = T64.abs().max(dim=0).values
    err = ((T32.double() - T64).abs().max(dim=0).values / scale).max().item()
    assert err <= 1e-5, err
    order = tree.order()
    ps = []
    for T in (T64, T32):
        q = ops.quant_reorder(T, order, 0.05)
        Td = ops.dequant_unreorder(q, order, 0.05, T.dtype)
        Crec = ops.inverse(tree, Td)
        ps.append(codec.psnr(cu(C).double(), Crec.double()))
    assert abs(ps[0] - ps[1]) <= 0.01, ps


# ---------------------------------------------------------------------------------------------------
# voxelisation front end
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", helpers.VOX_CASES)
def test_golden_voxelize(name):
    import voxelize_pc as VX
    g = helpers.load(name)
    PC, J = g["PC"], int(g["J"])
    vmin = torch.from_numpy(g["vmin_in"]) if g["vmin_in"].size else None
    width = float(g["width_in"][0]) if g["width_in"].size else None
    PCvox, PCsorted, vi, DeltaPC, info = VX.voxelize_pc_batched(cu(PC), vmin=vmin, width=width, J=J, device=DEV)
    assert info["Nvox"] == int(g["Nvox"]) and info["N"] == PC.shape[0]
    assert info["voxel_size"] == float(g["voxel_size"]) and info["width"] == float(g["width"])
    assert np.array_equal(info["vmin"].cpu().numpy(), g["vmin"])
    assert info["sort_idx"].dtype == torch.int64 and np.array_equal(info["sort_idx"].cpu().numpy(), g["sort_idx"])
    assert vi.dtype == torch.int64 and np.array_equal(vi.cpu().numpy(), g["voxel_indices"])
    assert np.array_equal(PCsorted.cpu().numpy(), g["PCsorted"])
    assert np.array_equal(PCvox.cpu().numpy(), g["PCvox"])
    assert np.array_equal(DeltaPC.cpu().numpy(), g["DeltaPC"])
    out4 = VX.voxelize_pc(cu(PC), {"J": J, "vmin": None if vmin is None else vmin.tolist(), "width": width})
    assert torch.equal(out4[0], PCvox) and torch.equal(out4[2], vi)
    V0i = cu(g["PCvox"][:, :3].astype(np.int64))
    assert np.array_equal(VX.get_morton_code(V0i, J).cpu().numpy(), helpers.morton_np(g["PCvox"][:, :3].astype(np.int64)))


@pytest.mark.parametrize("n,bits,seed", [(1, 8, 0), (1000, 5, 1), (3072, 17, 2), (3073, 36, 3), (200000, 48, 4), (1500000, 63, 5),
                                         (777777, 9, 6)])
def test_radix_sort_stable(n, bits, seed):
    from raht_b200 import ops
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 2 ** bits, size=n, dtype=np.int64)
    ks, idx, status = ops.radix_sort_pairs(cu(keys), bits)
    order = np.argsort(keys, kind="stable")
    assert int(status.item()) == 0
    assert np.array_equal(ks.cpu().numpy(), keys[order])
    assert np.array_equal(idx.cpu().numpy().astype(np.int64), order)
    vals = rng.integers(0, 2 ** 31, size=n).astype(np.int32)
    ks2, v2, _ = ops.radix_sort_pairs(cu(keys), bits, cu(vals))
    assert np.array_equal(v2.cpu().numpy(), vals[order])


@pytest.mark.parametrize("n,J,seed", [(200000, 6, 1), (300000, 12, 2), (50000, 16, 3)])
def test_oracle_voxelize_merge(n, J, seed):
    import voxelize_pc as VX
    import merge_cluster_cuda as MG
    from raht_b200 import codec
    means, quats, scales, opac, colors = helpers.synthetic_gaussians(n, seed)
    r = orc.voxelize(means, J, want_delta=False)
    PCvox, PCsorted, vi, DeltaPC, info = VX.voxelize_pc_batched(cu(means), J=J, device=DEV)
    assert info["Nvox"] == r["Nvox"]
    assert np.array_equal(info["sort_idx"].cpu().numpy(), r["sort_idx"])
    assert np.array_equal(vi.cpu().numpy(), r["voxel_indices"])
    Vvox = r["Vint"][r["sort_idx"]][r["voxel_indices"]].astype(np.float32)
    assert np.array_equal(PCvox.cpu().numpy(), Vvox)
    # test_voxelize_3dgs.py:225-247
    cidx = info["sort_idx"].int()
    coff = torch.cat([vi, torch.tensor([n], device=DEV)]).int()
    exp = orc.merge_clusters(r["sort_idx"].astype(np.int32), np.concatenate([r["voxel_indices"], [n]]).astype(np.int32),
                             means, quats, scales, opac, colors, True)
    got = MG.merge_gaussian_clusters_with_indices(cu(means), cu(quats), cu(scales), cu(opac), cu(colors), cidx, coff, True)
    for a, b, nm in zip(got, exp, ("means", "quats", "scales", "opacities", "colors")):
        assert np.array_equal(a.cpu().numpy(), b), nm
    got_u = MG.merge_gaussian_clusters_with_indices(cu(means), cu(quats), cu(scales), cu(opac), cu(colors), cidx, coff, False)
    exp_u = orc.merge_clusters(r["sort_idx"].astype(np.int32), np.concatenate([r["voxel_indices"], [n]]).astype(np.int32),
                               means, quats, scales, opac, colors, False)
    for a, b in zip(got_u, exp_u):
        assert np.array_equal(a.cpu().numpy(), b)
    packed, keys, info2 = codec.voxelize_merge(cu(means), cu(quats), cu(scales), cu(opac), cu(colors), J)
    assert np.array_equal(keys.cpu().numpy(), r["MC_sorted"][r["voxel_indices"]])
    assert np.array_equal(packed.cpu().numpy(), np.concatenate([exp[0], exp[1], exp[2], exp[3][:, None], exp[4]], 1))
    # label interface
    labels = cu(r["MC"])
    got_l = MG.merge_gaussian_clusters(cu(means), cu(quats), cu(scales), cu(opac), cu(colors), labels, True)
    for a, b in zip(got_l, exp):
        assert np.array_equal(a.cpu().numpy(), b)


def test_merge_errors_and_empty_clusters():
    import merge_cluster_cuda as MG
    m = torch.zeros(4, 3, device=DEV); q = torch.zeros(4, 4, device=DEV); s = torch.zeros(4, 3, device=DEV)
    o = torch.ones(4, device=DEV); c = torch.ones(4, 6, device=DEV)
    idx = torch.arange(4, device=DEV).int()
    off = torch.tensor([0, 2, 2, 4], device=DEV).int()          # middle cluster empty -> zeros (merge_cluster.cu:27-29)
    out = MG.merge_gaussian_clusters_with_indices(m, q, s, o, c, idx, off)
    assert out[3].tolist() == [1.0, 0.0, 1.0] and out[1][0].tolist() == [0.0, 0.0, 0.0, 1.0] and out[1][1].tolist() == [0.0] * 4
    with pytest.raises(RuntimeError):
        MG.merge_gaussian_clusters_with_indices(m.cpu(), q, s, o, c, idx, off)
    with pytest.raises(RuntimeError):
        MG.merge_gaussian_clusters_with_indices(m, q[:, :3], s, o, c, idx, off)


# ---------------------------------------------------------------------------------------------------
# RLGR
# ---------------------------------------------------------------------------------------------------
def test_rlgr_kats_membuf_interface():
    import rlgr
    for seq, out, fs in helpers.rlgr_kats():
        m = rlgr.membuf()
        m.rlgrWrite(seq.tolist(), fs)
        m.close()
        assert bytes(m.get_buffer()) == out, seq[:8]
        r = rlgr.membuf(list(out))
        _, dec = r.rlgrRead(len(seq), fs)
        assert dec == seq.tolist()


@pytest.mark.parametrize("zfrac,scale,n", [(0.0, 3.0, 70000), (0.3, 50.0, 70000), (0.85, 2.0, 90000), (0.97, 5.0, 120000),
                                           (0.9995, 1.0, 300000), (0.5, 3e8, 20000)])
def test_rlgr_random_vs_oracle(zfrac, scale, n):
    from raht_b200 import ops
    rng = np.random.default_rng(int(scale) + n)
    nch = 5
    x = np.clip(np.rint(rng.laplace(0, scale, size=(nch, n))), -2**31, 2**31 - 1).astype(np.int32)
    x[rng.random((nch, n)) < zfrac] = 0
    x[1, -300:] = 0            # trailing zero run (tail rule)
    x[2, -1] = -1              # tail quirk: u-1 == 0 in run mode
    x[3, :] = 0
    streams = ops.rlgr_encode(cu(x), 1, want_stats=True)
    exp, _, _, _ = orc.rlgr_encode_channels(x, 1)
    assert streams.to_bytes() == exp
    assert np.array_equal(ops.rlgr_decode(streams, 1).cpu().numpy(), x)                       # chunk-parallel, sync index
    assert np.array_equal(ops.rlgr_decode(streams, 1, use_index=False).cpu().numpy(), x)      # index rebuilt by a walk of the streams
    assert np.array_equal(ops.rlgr_decode(streams, 1, use_index=False, serial=True).cpu().numpy(), x)   # channel-serial
    assert streams.stats["chunks"] == nch * ((n + 255) // 256)                                 # one sync entry per 256 symbols
    no_index = ops.rlgr_encode(cu(x), 1, with_index=False)
    assert no_index.sync_index is None and no_index.to_bytes() == exp


def test_rlgr_regime_changes_and_unsigned():
    """Dense / sparse / empty stretches inside one channel (speculation must be repaired exactly), lengths
    around the block (128) and chunk (3968) sizes, and the unsigned mapping (flagSigned = 0)."""
    from raht_b200 import ops
    rng = np.random.default_rng(99)
    parts = [np.rint(rng.laplace(0, s, 30000)) for s in (0.05, 5, 0.3, 0.01, 50, 0.0, 1, 0.002, 300)]
    x = np.concatenate(parts).astype(np.int32)[None, :]
    streams = ops.rlgr_encode(cu(x), 1, want_stats=True)
    exp, _, _, _ = orc.rlgr_encode_channels(x, 1)
    assert streams.to_bytes() == exp
    assert np.array_equal(ops.rlgr_decode(streams, 1).cpu().numpy(), x)
    for n in (1, 2, 127, 128, 129, 3967, 3968, 3969, 7936, 7937, 12001):
        for last in (None, 0, -1, 7):
            y = np.rint(rng.laplace(0, 0.7, size=(3, n))).astype(np.int32)
            y[1] = np.abs(y[1])
            y[2, n // 3:] = 0
            if last is not None:
                y[0, -1] = last
            for fs in (1, 0):
                z = np.abs(y) if fs == 0 else y
                st = ops.rlgr_encode(cu(z), fs)
                e, _, _, _ = orc.rlgr_encode_channels(z, fs)
                assert st.to_bytes() == e, (n, last, fs)
                assert np.array_equal(ops.rlgr_decode(st, fs).cpu().numpy(), z), (n, last, fs)
                assert np.array_equal(ops.rlgr_decode(st, fs, use_index=False).cpu().numpy(), z), (n, last, fs)
                if n in (1, 129, 7937):
                    assert np.array_equal(ops.rlgr_decode(st, fs, use_index=False, serial=True).cpu().numpy(), z), (n, last, fs)


def test_rlgr_one_signed_sparse_channels_use_the_tables():
    """Sparse channels whose non-zeros are all positive never meet the p = 0 event that re-synchronises k_RP, and in
    run mode k_P keeps the parity of its differences: the candidate records fail and the transition tables of
    rlgr.cu (128-state window per block) must settle the chunks -- with the reference's bytes."""
    from raht_b200 import ops
    rng = np.random.default_rng(2024)
    n, nch = 600000, 6
    x = np.zeros((nch, n), dtype=np.int32)
    for c in range(nch):
        dens = (0.04, 0.08, 0.12, 0.2, 0.3, 0.15)[c]
        mag = (1.2, 2.0, 3.0, 1.5, 6.0, 40.0)[c]
        nz = rng.random(n) < dens * (0.4 + 1.2 * np.linspace(0, 1, n) ** (c % 3))          # density drifts along the channel
        v = 1 + np.floor(rng.exponential(mag, n)).astype(np.int32)
        x[c, nz] = v[nz]
    x[5, n // 2:] *= -1                                                                    # one channel changes sign half-way
    streams = ops.rlgr_encode(cu(x), 1, want_stats=True)
    exp, _, _, _ = orc.rlgr_encode_channels(x, 1)
    assert streams.to_bytes() == exp
    assert streams.stats["table_hits"] > 0, streams.stats
    assert np.array_equal(ops.rlgr_decode(streams, 1).cpu().numpy(), x)
    assert np.array_equal(ops.rlgr_decode(streams, 1, use_index=False).cpu().numpy(), x)


# ---------------------------------------------------------------------------------------------------
# whole scene, size-independent properties at a BASELINE-sized input
# ---------------------------------------------------------------------------------------------------
def test_scene_roundtrip_properties_1m():
    from raht_b200 import codec, ops
    n, J = 1_000_000, 12
    means, quats, scales, opac, colors = helpers.synthetic_gaussians(n, 2001)
    args = [cu(a) for a in (means, quats, scales, opac, colors)]
    enc = codec.encode_scene(*args, J=J, step=1.0 / 256)
    assert enc.D == 59 and 0.9 * n < enc.nvox <= n
    keys = enc.keys.cpu().numpy()
    assert np.all(np.diff(keys) > 0)                                     # sorted, unique
    coords, Crec = codec.decode_scene(enc)
    packed, keys2, _ = codec.voxelize_merge(*args, J=J)
    assert torch.equal(keys2, enc.keys)
    # quantisation error bound: orthonormal transform => per-coefficient error <= step/2, so RMSE <= step/2
    err = (Crec - packed.double())
    assert err.abs().max().item() < 1.0 and (err ** 2).mean().sqrt().item() <= 0.5 / 256
    # lossless path: iRAHT(RAHT(C)) == C to rounding; energy preserved; DC = sqrt(N) mean
    tree = enc.tree
    T = ops.forward(tree, packed, torch.float64)
    assert torch.allclose(ops.inverse(tree, T), packed.double(), rtol=0, atol=1e-9)
    assert torch.allclose(T.norm(dim=0), packed.double().norm(dim=0), rtol=1e-12)
    assert torch.allclose(T[0], packed.double().mean(0) * np.sqrt(enc.nvox), rtol=1e-9, atol=1e-9)
    order = tree.order()
    assert torch.equal(torch.sort(order).values, torch.arange(enc.nvox, device=DEV))
    # decode without the encoder-side tree gives the same result
    _, Crec2 = codec.decode_scene(enc, reuse_tree=False)
    assert torch.equal(Crec, Crec2)
    # the chunk-parallel decoder (sync index) and the sequential one (the reference's way) agree on every symbol
    q = ops.rlgr_decode(enc.streams, 1)
    assert torch.equal(q, ops.rlgr_decode(enc.streams, 1, use_index=False))
    # fused transform + quantisation: the same bitstreams and the same reconstruction
    enc_f = codec.encode_scene(*args, J=J, step=1.0 / 256, fused=True)
    assert torch.equal(enc_f.streams.nbytes, enc.streams.nbytes) and enc_f.streams.to_bytes() == enc.streams.to_bytes()
    assert torch.equal(codec.decode_scene(enc_f, fused=True)[1], Crec)
    # through the container
    from raht_b200 import container
    box = container.pack(enc)
    dec = container.unpack(box)
    assert torch.equal(dec.keys, enc.keys) and torch.equal(codec.decode_scene(dec)[1], Crec)
    assert box.numel() < enc.payload_bytes() + enc.index_bytes() + enc.nvox * 3 + 4096     # geometry under 3 bytes per voxel


def test_scene_pipeline_matches_direct_calls():
    """codec.ScenePipeline (copies on side streams, prefetch) must hand back exactly what encode_scene / decode_scene
    produce, for every scene of a stream of different scenes."""
    import torch
    from raht_b200 import codec
    J, step = 10, 0.02
    scenes = [helpers.synthetic_gaussians(60000 + 7000 * k, 300 + k) for k in range(4)]
    hosts = [[torch.from_numpy(a).pin_memory() for a in sc] for sc in scenes]
    pipe = codec.ScenePipeline(J, step)
    outs = []
    pipe.submit(hosts[0])
    for k in range(len(hosts)):
        if k + 1 < len(hosts):
            pipe.submit(hosts[k + 1])
        o = pipe.run()
        o["done"].synchronize()
        outs.append({n: (o[n].clone() if hasattr(o[n], "clone") else o[n]) for n in ("streams", "keys", "attributes", "sizes", "nvox")})
    pipe.finish()
    for sc, o in zip(scenes, outs):
        enc = codec.encode_scene(*[cu(a) for a in sc], J=J, step=step)
        _, Crec = codec.decode_scene(enc)
        assert o["nvox"] == enc.nvox and o["sizes"] == enc.streams.sizes()
        raw, off, parts = o["streams"].numpy(), 0, []
        for sz in o["sizes"]:                                            # streams end to end, each padded to 8 bytes
            parts.append(bytes(raw[off:off + sz]))
            off += (sz + 7) // 8 * 8
        assert parts == enc.streams.to_bytes() and off == raw.size
        assert np.array_equal(o["keys"].numpy(), enc.keys.cpu().numpy())
        assert np.array_equal(o["attributes"].numpy().reshape(enc.nvox, -1), Crec.cpu().numpy())
    # result="psnr": the squared error instead of the attributes
    pipe = codec.ScenePipeline(J, step, result="psnr")
    pipe.submit(hosts[0])
    o = pipe.run()
    pipe.finish()
    packed, _, _ = codec.voxelize_merge(*[cu(a) for a in scenes[0]], J=J)
    enc = codec.encode_scene(*[cu(a) for a in scenes[0]], J=J, step=step)
    _, Crec = codec.decode_scene(enc)
    sse = float(((packed.double() - Crec) ** 2).sum().item())
    assert abs(float(o["sse"][0]) - sse) <= 1e-9 * sse and "attributes" not in o


@pytest.mark.gpu
def test_scene_side_stream_overlap_matches_single_stream(monkeypatch):
    """The tree / plan kernels run on a side stream next to the merge (encoder) and the RLGR decoder (decoder):
    streams, keys and decoded attributes must equal those of the single-stream order, scene after scene (the side
    stream's allocations are consumed on the main stream and recycled by the next scene)."""
    import torch
    from raht_b200 import codec
    assert codec.OVERLAP or True
    scenes = [helpers.synthetic_gaussians(40000 + 9000 * k, 900 + k) for k in range(5)]
    results = {}
    for overlap in (True, False):
        monkeypatch.setattr(codec, "OVERLAP", overlap)
        outs = []
        for rep in range(2):
            for sc in scenes:
                enc = codec.encode_scene(*[cu(a) for a in sc], J=10, step=0.03)
                _, Crec = codec.decode_scene(enc)
                outs.append((enc.nvox, enc.keys.clone(), enc.streams.to_bytes(), Crec.clone()))
        torch.cuda.synchronize()
        results[overlap] = outs
    for a, b in zip(results[True], results[False]):
        assert a[0] == b[0] and torch.equal(a[1], b[1]) and a[2] == b[2] and torch.equal(a[3], b[3])


@pytest.mark.gpu
@pytest.mark.parametrize("n,J,D,tile_rows,dtype", [(1, 3, 3, None, "f64"), (7, 2, 2, None, "f64"), (300, 5, 5, 32, "f64"),
                                                   (20000, 8, 59, None, "f64"), (20000, 8, 59, 64, "f32"), (70000, 10, 11, 128, "f64")])
def test_fused_quantisation_matches_the_two_reference_steps(n, J, D, tile_rows, dtype):
    """raht_forward_quant == raht_forward + raht_quant_reorder and raht_dequant_inverse == raht_dequant_unreorder +
    raht_inverse, bit for bit (encode_3dgs.py:198-210 and :255-276), whatever the number of plan stages."""
    import torch
    from raht_b200 import ops
    rng = np.random.default_rng(n + D)
    keys = np.sort(rng.choice(1 << (3 * J), size=min(n, 1 << (3 * J)), replace=False)).astype(np.int64)
    C = rng.normal(0, 3, size=(keys.size, D)).astype(np.float32)
    C[:, 0] *= 1000.0
    tdt = torch.float64 if dtype == "f64" else torch.float32
    steps = [0.03125 * (1 + (c % 3)) for c in range(D)]
    for step in (0.25, steps):
        tree = ops.tree_from_keys(torch.from_numpy(keys).cuda())
        order = tree.order()
        Cd = torch.from_numpy(C).cuda()
        T = ops.forward(tree, Cd, tdt, tile_rows)
        q_ref = ops.quant_reorder(T, order, step)
        q = ops.forward_quant(tree, Cd, step, tdt, tile_rows)
        assert torch.equal(q, q_ref)
        rec_ref = ops.inverse(tree, ops.dequant_unreorder(q_ref, order, step, tdt), tile_rows)
        rec = ops.dequant_inverse(tree, q_ref, step, tdt, tile_rows)
        assert torch.equal(rec, rec_ref)
    # int32 overflow is still reported
    big = torch.full((keys.size, D), 3.0e9, dtype=torch.float32).cuda()
    with pytest.raises(OverflowError):
        ops.forward_quant(ops.tree_from_keys(torch.from_numpy(keys).cuda()), big, 0.5, tdt, tile_rows)


@pytest.mark.gpu
@pytest.mark.parametrize("D", [56, 59])
def test_per_attribute_steps_match_the_reference_rule(D):
    """encode_3dgs_debug.py:326-381: range of the coefficients per attribute group -> levels -> step -> quantise."""
    import torch
    from raht_b200 import codec, ops
    rng = np.random.default_rng(D)
    J, n = 8, 15000
    keys = np.sort(rng.choice(1 << (3 * J), size=n, replace=False)).astype(np.int64)
    C = rng.normal(0, 1, size=(n, D)) * np.linspace(0.2, 5.0, D)
    tree_o = orc.Tree(keys, J)
    T_o, _ = tree_o.forward(C)
    tree = ops.tree_from_keys(torch.from_numpy(keys).cuda())
    T = ops.forward(tree, torch.from_numpy(C).cuda(), torch.float64)
    assert np.array_equal(T.cpu().numpy(), T_o)
    lo, hi = ops.channel_minmax(T)
    assert np.array_equal(lo.cpu().numpy(), T_o.min(0)) and np.array_equal(hi.cpu().numpy(), T_o.max(0))
    lo32, hi32 = ops.channel_minmax(T.float())
    assert np.array_equal(lo32.cpu().numpy(), T_o.astype(np.float32).min(0).astype(np.float64))
    ranges = codec.attribute_ranges(D)
    steps, info = codec.per_attribute_steps(T)
    want, enc_o = orc.per_attribute_steps(T_o, ranges, {g: codec.IMPORTANCE[g] for g in ranges})
    steps_o, info_o = codec.per_attribute_steps(T, ranges, {g: codec.IMPORTANCE[g] for g in ranges})
    for g, (a, b) in ranges.items():
        assert (info_o[g]["step"], info_o[g]["levels"], info_o[g]["range"]) == want[g]
    q = ops.quant_reorder(T, None, steps_o)
    assert np.array_equal(q.cpu().numpy().T.astype(np.float64), enc_o)
    assert steps == steps_o
    if D == 56:
        assert [info[g]["levels"] for g in ("quats", "scales", "opacity", "colors")] == [350, 291, 182, 198]


@pytest.mark.gpu
def test_scene_with_per_attribute_steps_roundtrips_through_the_container():
    import torch
    from raht_b200 import codec, container
    scene = [torch.from_numpy(a).cuda() for a in helpers.synthetic_gaussians(20000, 8)]
    enc = codec.encode_scene(*scene, J=8, step="per_attribute")
    assert isinstance(enc.step, list) and len(enc.step) == 59 and len(set(enc.step)) == 5
    _, ref = codec.decode_scene(enc)
    dec = container.unpack(container.pack(enc))
    assert dec.step == enc.step
    _, rec = codec.decode_scene(dec)
    assert torch.equal(rec, ref)
    packed, _, _ = codec.voxelize_merge(*scene, J=8)
    assert codec.psnr(packed, rec) > 20.0


@pytest.mark.gpu
@pytest.mark.parametrize("n,J,step", [(1, 4, 0.5), (3000, 6, 0.02), (120000, 11, 1.0 / 64), (250000, 16, [0.01 * (1 + c % 4) for c in range(59)])])
def test_scene_level_calls_match_the_operator_path(n, J, step, monkeypatch):
    """raht_scene_front / raht_scene_encode / raht_scene_decode (one host call per leg, csrc/scene.cu) against the same
    pipeline driven operator by operator: keys, merged attributes, order, streams, sync index and decoded attributes are
    the same bytes; the tree a scene-level encode returns is usable by the operators."""
    from raht_b200 import codec, ops
    scene = [cu(a) for a in helpers.synthetic_gaussians(n, 4242 + n)]
    monkeypatch.setattr(codec, "SCENE_CALLS", True)
    a = codec.encode_scene(*scene, J=J, step=step, keep_packed=True)
    ca, ra = codec.decode_scene(a)
    monkeypatch.setattr(codec, "SCENE_CALLS", False)
    b = codec.encode_scene(*scene, J=J, step=step, keep_packed=True)
    cb, rb = codec.decode_scene(b)
    assert a.nvox == b.nvox and a.D == b.D and a.voxel_size == b.voxel_size and a.width == b.width
    assert torch.equal(a.vmin, b.vmin) and torch.equal(a.keys, b.keys) and torch.equal(a.packed, b.packed)
    assert a.streams.sizes() == b.streams.sizes() and a.streams.to_bytes() == b.streams.to_bytes()
    assert torch.equal(a.streams.sync_index, b.streams.sync_index)
    assert torch.equal(ca, cb) and torch.equal(ra, rb)
    if a.nvox > 1:
        assert torch.equal(a.tree.order(), b.tree.order()) and torch.equal(a.tree.level, b.tree.level) and torch.equal(a.tree.w, b.tree.w)
        T = ops.forward(a.tree, a.packed, torch.float64)                       # the plan inside the tree workspace
        assert torch.equal(T, ops.forward(b.tree, b.packed, torch.float64))
    monkeypatch.setattr(codec, "SCENE_CALLS", True)
    _, rc = codec.decode_scene(b)                                              # streams of one path, decoder of the other
    assert torch.equal(rc, ra)
    assert torch.equal(codec.decode_scene(a, reuse_tree=True)[1], ra)


def test_deferred_checks_give_the_same_scene_and_still_raise():
    """encode_scene / decode_scene with checks=[...]: the calls return with their kernels queued, the callables do the
    status / length read-back later (bench.py runs them one step behind).  Same streams and attributes as the waiting
    calls; several scenes in flight do not disturb each other's results (ring of pinned result buffers); an invalid voxel
    list handed to the decoder raises when its check runs, not silently never."""
    from raht_b200 import codec
    scenes = [[cu(a) for a in helpers.synthetic_gaussians(30000 + 7000 * k, 900 + k)] for k in range(4)]
    eager = [codec.encode_scene(*s, J=10, step=1.0 / 16) for s in scenes]
    eager_dec = [codec.decode_scene(e)[1] for e in eager]
    checks, lazy, lazy_dec = [], [], []
    for s in scenes:                                                            # four scenes queued before any check runs
        e = codec.encode_scene(*s, J=10, step=1.0 / 16, checks=checks)
        lazy.append(e)
        lazy_dec.append(codec.decode_scene(e, checks=checks)[1])
    assert len(checks) == 8 and all(e.streams._sizes is None for e in lazy)
    for c in checks:
        c()
    for a, b, ra, rb in zip(eager, lazy, eager_dec, lazy_dec):
        assert a.nvox == b.nvox and b.streams._sizes == a.streams.sizes() and a.streams.to_bytes() == b.streams.to_bytes()
        assert torch.equal(a.keys, b.keys) and torch.equal(ra, rb)
    bad = lazy[0]
    bad.keys = torch.flip(bad.keys, [0]).contiguous()                           # not Morton-sorted any more
    checks = []
    codec.decode_scene(bad, checks=checks)
    with pytest.raises(ValueError):
        for c in checks:
            c()

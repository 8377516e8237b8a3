"""GPU parity at BASELINE sizes (-m gpu): the exact scenes bench.py times (C2: seed 2001, 1 M Gaussians, depth 12;
C3: seed 3001, 6 M Gaussians, depth 16) through the CUDA library, compared BIT FOR BIT with the CPU oracle --
voxel keys, order_RAGFT, the quantised symbols of all 59 channels, every RLGR stream (coded on the host by the
reference's own coder object code, oracle/_ref/librlgr_ref.so, when built; else the oracle port) and the decoded
attributes -- at the dense step bench.py uses (1/64) and at sparse steps, where the RLGR encoder leaves the
speculation records for its tables / helper warps (the regime-dependent paths DESIGN 4.2 describes)."""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest
import torch

import helpers
from oracle import oracle_ctypes as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_streams(q):
    """RLGR bytes of every channel of q (nch, n) int32 on the host: reference coder when built, else the oracle port."""
    if orc.ref_rlgr_available():
        with ThreadPoolExecutor(max_workers=_cores()) as pool:
            return list(pool.map(lambda c: orc.ref_rlgr_encode(q[c].astype(np.int64), 1)[0], range(q.shape[0])))
    return list(orc.rlgr_encode_channels(q, 1)[0])


def oracle_front(scene, J):
    means, quats, scales, opac, colors = scene
    n = means.shape[0]
    r = orc.voxelize(means, J, want_delta=False)
    coff = np.concatenate([r["voxel_indices"], [n]]).astype(np.int32)
    mm, mq, ms, mo, mc = orc.merge_clusters(r["sort_idx"].astype(np.int32), coff, means, quats, scales, opac, colors, True)
    packed = np.concatenate([mm, mq, ms, mo[:, None], mc], 1)
    keys = r["MC_sorted"][r["voxel_indices"]]
    return packed, keys


@pytest.mark.parametrize("name,n,J,seed,steps", [("c2", 1_000_000, 12, 2001, (1.0 / 64, 0.25, 2.0)),
                                                 ("c3", 6_000_000, 16, 3001, (1.0 / 64, 0.25)),
                                                 ("c2_rank3", 1_000_000, 12, 2004, (1.0 / 64,))])   # a replica rank's scene: 3 exact repairs in channel 0
def test_bench_scene_is_bit_exact_with_the_oracle(name, n, J, seed, steps):
    from raht_b200 import codec, ops
    orc.set_num_threads(_cores())
    scene = helpers.synthetic_gaussians(n, seed)
    args = [torch.from_numpy(a).to(DEV) for a in scene]
    packed_o, keys_o = oracle_front(scene, J)
    t = orc.Tree(keys_o, J)
    T_o, _ = t.forward(packed_o.astype(np.float64))
    order_o = t.order()
    inv_o = np.argsort(order_o, kind="stable")
    for step in steps:
        enc = codec.encode_scene(*args, J=J, step=step, keep_packed=True)
        assert enc.nvox == len(keys_o)
        assert np.array_equal(enc.keys.cpu().numpy(), keys_o), "voxel keys"
        assert np.array_equal(enc.packed.cpu().numpy(), packed_o), "merged attributes"
        assert np.array_equal(enc.tree.order().cpu().numpy(), order_o), "order_RAGFT"
        q_o = np.ascontiguousarray(orc.quantize(T_o, step)[order_o].T.astype(np.int32))
        q = ops.rlgr_decode(enc.streams, 1)
        assert np.array_equal(q.cpu().numpy(), q_o), f"step {step}: quantised symbols (through the indexed decoder)"
        del q
        got = enc.streams.to_bytes()
        exp = cpu_streams(q_o)
        for c in range(len(exp)):
            assert got[c] == exp[c], f"{name} step {step}: RLGR bytes of channel {c} differ from the host coder"
        del got, exp
        _, Crec = codec.decode_scene(enc)
        Tdec = orc.dequantize(q_o.T.astype(np.float64), step)[inv_o]
        assert np.array_equal(Crec.cpu().numpy(), t.inverse(Tdec)), f"step {step}: decoded attributes"
        del enc, Crec, Tdec, q_o
        torch.cuda.empty_cache()


def test_reference_format_streams_without_index_decode_at_scale():
    """A stream written by the reference's own coder carries no sync index: the decoder must still return every symbol
    (1 M-symbol channels, dense and sparse), and fast enough to be usable."""
    from raht_b200 import ops
    rng = np.random.default_rng(5)
    n, nch = 1_000_000, 6
    x = np.rint(rng.laplace(0, 4.0, size=(nch, n))).astype(np.int32)
    for c, z in enumerate((0.0, 0.5, 0.9, 0.99, 0.999, 1.0)):
        x[c, rng.random(n) < z] = 0
    streams = ops.rlgr_encode(torch.from_numpy(x).to(DEV), 1, with_index=False)
    assert streams.sync_index is None
    assert streams.to_bytes() == cpu_streams(x)
    t0, t1, t2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    t0.record()
    index = ops.rlgr_build_index(streams)                     # the serial walk (a warp per channel) ...
    t1.record()
    streams.sync_index = index
    got = ops.rlgr_decode(streams, 1)                         # ... then every 256-symbol range in parallel
    t2.record()
    torch.cuda.synchronize()
    assert np.array_equal(got.cpu().numpy(), x)
    streams.sync_index = None
    assert np.array_equal(ops.rlgr_decode(streams, 1, use_index=False, serial=True).cpu().numpy(), x)
    print(f"index-less decode of {nch} x {n}: walk {t0.elapsed_time(t1):.2f} ms + ranges {t1.elapsed_time(t2):.2f} ms")


def test_c1_point_cloud_frame_is_bit_exact_with_the_oracle():
    """BASELINE configs[0] -- the encode_ply.py workload at its size: ~750 K voxels on a sphere surface at depth 10, RGB ->
    YUV, RAHT, the reference's 11 quantisation steps (encode_ply.py:29), RLGR of Y / U / V and back.  The colour transform
    is a floating-point row (1e-10); from the SAME YUV input on, symbols, bytes and reconstructions are bit-exact, for
    the two-pass path and for the fused transform + quantisation."""
    from raht_b200 import codec, synth, ops
    import utils as shim_utils
    V, rgb = synth.sphere_point_cloud()
    assert 600_000 < V.shape[0] < 900_000
    J = 10
    Vd, rgbd = torch.from_numpy(V.astype(np.float64)).to(DEV), torch.from_numpy(rgb).to(DEV)
    yuv = shim_utils.rgb_to_yuv(rgbd)
    np.testing.assert_allclose(yuv.cpu().numpy(), orc.rgb_to_yuv(rgb), rtol=0, atol=1e-10)
    res, C = codec.encode_ply_frame(Vd, rgbd, J, keep_reconstruction=True)
    assert torch.equal(C, yuv)
    Cn = C.cpu().numpy()
    keys = orc.morton(V, J)
    t = orc.Tree(keys, J)
    T_o, _ = t.forward(Cn)
    order_o = t.order()
    inv_o = np.argsort(order_o, kind="stable")
    assert [r["step"] for r in res] == codec.PLY_COLOR_STEPS
    for r in res:
        step = r["step"]
        q_o = np.ascontiguousarray(orc.quantize(T_o, step)[order_o].T.astype(np.int32))
        exp = cpu_streams(q_o)
        assert r["streams"].to_bytes() == exp, f"step {step}: RLGR bytes"
        assert r["bytes"] == sum(len(b) for b in exp)
        Tdec = orc.dequantize(q_o.T.astype(np.float64), step)[inv_o]
        assert np.array_equal(r["C_rec"].cpu().numpy(), t.inverse(Tdec)), f"step {step}: reconstruction"
        y_hat = np.floor(T_o[:, 0] / step + 0.5) * step                                     # encode_ply.py:149-152
        psnr = -10 * np.log10(np.linalg.norm(T_o[:, 0] - y_hat) ** 2 / (len(keys) * 255.0 ** 2))
        assert abs(r["psnr_y"] - psnr) <= 1e-9 * abs(psnr)
    res_f, _ = codec.encode_ply_frame(Vd, rgbd, J, steps=[1, 12, 64], fused=True, keep_reconstruction=True)
    for rf in res_f:
        r = next(x for x in res if x["step"] == rf["step"])
        assert rf["streams"].to_bytes() == r["streams"].to_bytes() and torch.equal(rf["C_rec"], r["C_rec"])

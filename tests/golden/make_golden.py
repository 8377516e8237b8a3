"""Generates the golden fixtures in this directory from the REFERENCE ITSELF.

Run in the build container only (needs /root/reference and oracle/_ref/librlgr_ref.so):
    python tests/golden/make_golden.py
It imports the reference's Python operators unmodified (python/RAHT_param.py, RAHT.py, iRAHT.py,
voxelize_pc.py, utils.py -- torch on CPU) and calls the reference's own RLGR coder through
oracle/_ref/librlgr_ref.so (membuf.cpp compiled where it lies).  Two documented deviations from
running the reference drivers as they are:
  1. torch.sort is called with stable=True inside voxelize_pc_batched (the
     reference's tie order is unspecified);
  2. torch.sqrt is replaced by the IEEE-754 correctly rounded square root while the reference's RAHT /
     iRAHT run.  This container's torch-CPU build (MKL VML, AVX-512) returns a result 1 ulp off for
     ~0.7 % of float64 inputs (e.g. sqrt(0.5) -> 0.7071067811865475 instead of ...476), whereas torch
     on CUDA -- the reference's production device (encode_3dgs.py:22) -- and MATLAB round correctly.
     The IEEE result is the specification; tests/test_oracle.py::test_oracle_vs_reference_live shows
     both (bit-exact with the patch, <= 1e-12 without it).
The fixtures pin the CPU oracle (tests/test_oracle.py) and the CUDA path (tests/test_gpu_*.py).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, "/root/reference/python")
sys.path.insert(0, ROOT)

import RAHT_param as ref_param          # noqa: E402
import RAHT as ref_raht                 # noqa: E402
import iRAHT as ref_iraht               # noqa: E402
import voxelize_pc as ref_vox           # noqa: E402
import utils as ref_utils               # noqa: E402
from oracle import oracle_ctypes as orc  # noqa: E402  (only for the reference RLGR shim loader)

torch.set_num_threads(4)

_torch_sqrt = torch.sqrt


def ieee_sqrt(x):
    return torch.from_numpy(np.sqrt(x.detach().cpu().numpy()))


def morton_np(V, J):
    return ref_vox.get_morton_code(torch.from_numpy(V.astype(np.int64)), J).numpy()


def unique_sorted_cloud(rng, n, J, clustered=True):
    hi = 2 ** J
    if clustered:
        c = rng.integers(0, hi, size=(max(n // 40, 1), 3))
        P = c[rng.integers(0, len(c), size=n)] + rng.normal(0, hi / 40, size=(n, 3))
        P = np.clip(np.floor(P), 0, hi - 1).astype(np.int64)
    else:
        P = rng.integers(0, hi, size=(n, 3))
    mc = morton_np(P, J)
    _, first = np.unique(mc, return_index=True)
    P = P[first]
    mc = mc[first]
    o = np.argsort(mc, kind="stable")
    return P[o], mc[o]


def raht_case(name, V, J, C, steps):
    Vt = torch.from_numpy(V.astype(np.float64))
    origin = torch.zeros(3, dtype=torch.float64)
    List, Flags, weights, order = ref_param.RAHT_param_reorder_fast(Vt, origin, 2 ** J, J)
    L2, F2, W2 = ref_param.RAHT_param(Vt, origin, 2 ** J, J)
    assert len(L2) == len(List) and all(torch.equal(a, b) for a, b in zip(L2, List))
    assert all(torch.equal(a, b) for a, b in zip(F2, Flags)) and all(torch.equal(a, b) for a, b in zip(W2, weights))
    Ct = torch.from_numpy(C)
    torch.sqrt = ieee_sqrt                    # deviation 2 (see the module docstring)
    T, w = ref_raht.RAHT2_optimized(Ct, List, Flags, weights)
    Crec_lossless = ref_iraht.inverse_RAHT_optimized(T, List, Flags, weights)
    out = {
        "V": V.astype(np.int64), "J": np.int64(J), "C": C, "T": T.numpy(), "w": w.numpy().reshape(-1),
        "C_lossless": Crec_lossless.numpy(),
        "order": order.numpy() if order is not None else np.zeros(0, np.int64),
        "nlevels": np.int64(len(Flags)),
        "level_len": np.array([len(x) for x in List], dtype=np.int64),
        "List": np.concatenate([x.numpy() for x in List]),
        "Flags": np.concatenate([x.numpy().astype(np.uint8) for x in Flags]),
        "weights": np.concatenate([x.numpy() for x in weights]),
        "steps": np.array(steps, dtype=np.int64),
    }
    for s in steps:
        # encode_3dgs.py:204-217, 229-234, 255-274
        Coeff_enc = torch.floor(T / s + 0.5)
        reordered = Coeff_enc.index_select(0, order)
        q = reordered.to(torch.int32).numpy()                       # (N, D)
        streams = []
        for ch in range(q.shape[1]):
            buf, _ = orc.ref_rlgr_encode(q[:, ch].astype(np.int64), 1)
            dec, _ = orc.ref_rlgr_decode(buf, q.shape[0], 1)
            assert np.array_equal(dec, q[:, ch])
            streams.append(np.frombuffer(buf, dtype=np.uint8))
        Coeff_dec = torch.from_numpy(q).to(torch.float64) * s
        inv = torch.argsort(order)
        Coeff_dec = Coeff_dec[inv, :]
        C_rec = ref_iraht.inverse_RAHT_optimized(Coeff_dec, List, Flags, weights)
        out[f"q_step{s}"] = np.ascontiguousarray(q.T)             # channel-major (D, N)
        out[f"stream_len_step{s}"] = np.array([len(x) for x in streams], dtype=np.int64)
        out[f"streams_step{s}"] = np.concatenate(streams) if streams else np.zeros(0, np.uint8)
        out[f"C_rec_step{s}"] = C_rec.numpy()
    torch.sqrt = _torch_sqrt
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "N", V.shape[0], "levels", len(Flags), "bytes", {s: int(out[f'stream_len_step{s}'].sum()) for s in steps})


def voxelize_case(name, PC, J, vmin=None, width=None):
    orig_sort = torch.sort
    torch.sort = lambda x, *a, **k: orig_sort(x, *a, stable=True, **k)     # the one documented deviation
    try:
        PCvox, PCsorted, vi, DeltaPC, info = ref_vox.voxelize_pc_batched(
            torch.from_numpy(PC), vmin=None if vmin is None else torch.tensor(vmin, dtype=torch.float32),
            width=width, J=J, device="cpu")
    finally:
        torch.sort = orig_sort
    np.savez_compressed(os.path.join(HERE, name + ".npz"), PC=PC, J=np.int64(J),
                        vmin_in=np.zeros(0, np.float32) if vmin is None else np.array(vmin, np.float32),
                        width_in=np.array([] if width is None else [width], np.float64),
                        PCvox=PCvox.numpy(), PCsorted=PCsorted.numpy(), voxel_indices=vi.numpy(), DeltaPC=DeltaPC.numpy(),
                        Nvox=np.int64(info["Nvox"]), voxel_size=np.float64(info["voxel_size"]),
                        vmin=info["vmin"].numpy(), width=np.float64(info["width"]), sort_idx=info["sort_idx"].numpy())
    print(name, "N", PC.shape[0], "Nvox", info["Nvox"])


def main():
    assert orc.ref_rlgr_available(), "build oracle/_ref first: make -C oracle"
    rng = np.random.default_rng(20261104)

    # --- RLGR known-answer vectors from the reference coder (SURVEY App. B inputs + random) -----------
    kat_inputs = [
        [0], [1], [-1], [0] * 16, [0] * 100, [-4, -3, -2, -1, 0, 1, 2, 3, 4], [0, 0, 0, 0, 1], [0, 0, 0, 0, -1],
        [0, 0, 0, 0, 2], [0, 0, 0, 5, -3, 0, 0, 0, 0, 0, 0, 0, 1, 0, 0],
        [100000, -100000, 2147483647, -2147483648],
    ]
    for zfrac, scale, n in [(0.0, 3.0, 4000), (0.3, 40.0, 4000), (0.8, 2.0, 6000), (0.97, 5.0, 8000), (0.999, 1.0, 20000),
                            (0.5, 3e5, 3000)]:
        x = np.rint(rng.laplace(0, scale, size=n)).astype(np.int64)
        x[rng.random(n) < zfrac] = 0
        kat_inputs.append(x.tolist())
    kat_inputs.append([0] * 70000)
    kat_inputs.append(([0] * 5000 + [7]) * 3 + [0] * 4095)
    seqs, outs, flags = [], [], []
    for s in kat_inputs:
        buf, _ = orc.ref_rlgr_encode(np.array(s, dtype=np.int64), 1)
        seqs.append(np.array(s, dtype=np.int64)); outs.append(np.frombuffer(buf, np.uint8)); flags.append(1)
    u = [0, 3, 0, 0, 20, 1]
    buf, _ = orc.ref_rlgr_encode(np.array(u, dtype=np.int64), 0)
    seqs.append(np.array(u, dtype=np.int64)); outs.append(np.frombuffer(buf, np.uint8)); flags.append(0)
    np.savez_compressed(os.path.join(HERE, "rlgr_kat.npz"),
                        seq=np.concatenate(seqs), seq_len=np.array([len(x) for x in seqs], np.int64),
                        out=np.concatenate(outs), out_len=np.array([len(x) for x in outs], np.int64),
                        flag_signed=np.array(flags, np.int64))
    print("rlgr_kat", len(seqs), "vectors; hex of first ten:", [bytes(o).hex() for o in outs[:10]])

    # --- RAHT: SURVEY App. B toy + random clouds -------------------------------------------------------
    V = np.array([[0, 0, 0], [0, 0, 1], [0, 1, 0], [1, 2, 0], [2, 1, 2], [2, 2, 2], [3, 3, 3]], dtype=np.int64)
    C = np.stack([np.arange(1, 8, dtype=np.float64), 10.0 * np.arange(1, 8)], axis=1)
    raht_case("raht_toy_J2", V, 2, C, [1])
    V, _ = unique_sorted_cloud(rng, 3000, 7, clustered=True)
    C = rng.normal(0, 20, size=(V.shape[0], 5))
    raht_case("raht_clustered_J7_D5", V, 7, C, [1, 12])
    V, _ = unique_sorted_cloud(rng, 2500, 12, clustered=False)
    C = np.concatenate([rng.normal(0, 1, size=(V.shape[0], 8)), rng.integers(0, 256, size=(V.shape[0], 3)).astype(np.float64)], 1)
    raht_case("raht_sparse_J12_D11", V, 12, C, [1, 20])
    # D = 59 3DGS-like rows (small)
    V, _ = unique_sorted_cloud(rng, 700, 9, clustered=True)
    C = rng.normal(0, 1, size=(V.shape[0], 59)) * np.linspace(0.1, 30, 59)[None, :]
    raht_case("raht_3dgs_J9_D59", V, 9, C, [1, 4])
    # YUV point cloud (config 1 flavour): surface of a sphere, D = 3
    d = rng.normal(size=(6000, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    P = np.clip(np.floor(64 + 50 * d), 0, 127).astype(np.int64)
    mc = morton_np(P, 7); _, first = np.unique(mc, return_index=True); P = P[first]; P = P[np.argsort(mc[first], kind="stable")]
    rgb = rng.integers(0, 256, size=(P.shape[0], 3)).astype(np.float64)
    yuv = ref_utils.rgb_to_yuv(torch.from_numpy(rgb)).numpy()
    raht_case("raht_yuv_sphere_J7_D3", P, 7, yuv, [1, 8])
    np.savez_compressed(os.path.join(HERE, "rgb_to_yuv.npz"), rgb=rgb, yuv=yuv)

    # --- voxelisation ---------------------------------------------------------------------------------
    PC = rng.normal(0, 1, size=(4000, 3)).astype(np.float32)
    voxelize_case("voxelize_pos_J5", PC, 5)
    PC = np.concatenate([rng.uniform(-3, 7, size=(3000, 3)), rng.uniform(0, 255, size=(3000, 2))], 1).astype(np.float32)
    voxelize_case("voxelize_attr_J4", PC, 4)
    voxelize_case("voxelize_attr_J6_given", PC, 6, vmin=[-3.5, -3.25, -4.0], width=12.0)


if __name__ == "__main__":
    main()

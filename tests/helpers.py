"""Shared helpers for the test-suite: golden-fixture loading and seeded synthetic scenes."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

RAHT_CASES = ["raht_toy_J2", "raht_clustered_J7_D5", "raht_sparse_J12_D11", "raht_3dgs_J9_D59", "raht_yuv_sphere_J7_D3"]
VOX_CASES = ["voxelize_pos_J5", "voxelize_attr_J4", "voxelize_attr_J6_given"]


def load(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz")))


def split_levels(g):
    """-> list of (List, Flags, weights) per level from the concatenated golden arrays."""
    out, o = [], 0
    for L in g["level_len"]:
        L = int(L)
        out.append((g["List"][o:o + L], g["Flags"][o:o + L].astype(bool), g["weights"][o:o + L]))
        o += L
    return out


def split_streams(g, step):
    lens = g[f"stream_len_step{step}"]
    blob = g[f"streams_step{step}"]
    out, o = [], 0
    for n in lens:
        out.append(bytes(blob[o:o + int(n)]))
        o += int(n)
    return out


def rlgr_kats():
    g = load("rlgr_kat")
    seqs, outs = [], []
    o = 0
    for n in g["seq_len"]:
        seqs.append(g["seq"][o:o + int(n)]); o += int(n)
    o = 0
    for n in g["out_len"]:
        outs.append(bytes(g["out"][o:o + int(n)])); o += int(n)
    return list(zip(seqs, outs, [int(f) for f in g["flag_signed"]]))


def spread3(x):
    x = x.astype(np.uint64) & np.uint64(0x1fffff)
    x = (x | (x << np.uint64(32))) & np.uint64(0x1f00000000ffff)
    x = (x | (x << np.uint64(16))) & np.uint64(0x1f0000ff0000ff)
    x = (x | (x << np.uint64(8))) & np.uint64(0x100f00f00f00f00f)
    x = (x | (x << np.uint64(4))) & np.uint64(0x10c30c30c30c30c3)
    x = (x | (x << np.uint64(2))) & np.uint64(0x1249249249249249)
    return x


def morton_np(P):
    return (spread3(P[:, 2]) | (spread3(P[:, 1]) << np.uint64(1)) | (spread3(P[:, 0]) << np.uint64(2))).astype(np.int64)


def synthetic_gaussians(n, seed, color_dim=48):
    """3DGS-like scene of SURVEY 8(d): 70 % on ellipsoid shells, 30 % uniform; fp32."""
    rng = np.random.default_rng(seed)
    n_shell = int(0.7 * n)
    k = 64
    centres = rng.uniform(-0.7, 0.7, size=(k, 3))
    radii = rng.uniform(0.05, 0.3, size=(k, 3))
    which = rng.integers(0, k, size=n_shell)
    d = rng.normal(size=(n_shell, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    shell = centres[which] + radii[which] * d + rng.normal(0, 0.002, size=(n_shell, 3))
    uni = rng.uniform(-1, 1, size=(n - n_shell, 3))
    means = np.concatenate([shell, uni]).astype(np.float32)
    rng.shuffle(means)
    quats = rng.normal(size=(n, 4)).astype(np.float32)
    quats /= np.linalg.norm(quats, axis=1, keepdims=True)
    scales = np.exp(rng.normal(-4, 1, size=(n, 3))).astype(np.float32)
    opac = (1.0 / (1.0 + np.exp(-rng.normal(0, 2, size=n)))).astype(np.float32)
    colors = np.concatenate([0.5 * rng.normal(size=(n, 3)), 0.1 * rng.normal(size=(n, color_dim - 3))], 1).astype(np.float32)
    return means, quats, scales, opac, colors


def unique_sorted_voxels(n, J, seed, clustered=True):
    rng = np.random.default_rng(seed)
    hi = 2 ** J
    if clustered:
        c = rng.integers(0, hi, size=(max(n // 50, 1), 3))
        P = c[rng.integers(0, len(c), size=n)] + rng.normal(0, hi / 30, size=(n, 3))
        P = np.clip(np.floor(P), 0, hi - 1).astype(np.int64)
    else:
        P = rng.integers(0, hi, size=(n, 3)).astype(np.int64)
    mc = morton_np(P)
    mc_u, first = np.unique(mc, return_index=True)
    return P[first], mc_u

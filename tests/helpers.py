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


import sys as _sys
_PKG = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "raht-3dgs-codec_b200")
if _PKG not in _sys.path:
    _sys.path.insert(0, _PKG)
# the scene generators live in the package (bench.py times exactly the scenes the tests check); they are numpy-only,
# but importing the package loads libraht_b200.so -- which every test module needs anyway
from raht_b200.synth import synthetic_gaussians, morton_np, sphere_point_cloud, _spread3 as spread3  # noqa: E402,F401


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

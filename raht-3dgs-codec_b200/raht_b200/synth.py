"""Seeded synthetic workloads of SURVEY 8(d) (numpy only; no device work): the 3DGS-like scenes of BASELINE configs
C2-C5 and the voxelised RGB point cloud of config C1.  bench.py and the tests share these generators, so the scenes
that are timed are the scenes that are parity-tested.  (SURVEY 8(d) proposes torch.Generator; numpy's default_rng is
used instead -- equally deterministic on every box, and the CPU checker consumes numpy arrays directly.)"""
import numpy as np


def synthetic_gaussians(n, seed, color_dim=48):
    """3DGS-like scene: 70 % of the centres on 64 random ellipsoid shells, 30 % uniform in [-1, 1]^3; fp32.
    Returns (means (n,3), quats (n,4), scales (n,3), opacities (n,), colors (n,color_dim))."""
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


def _spread3(x):
    x = x.astype(np.uint64) & np.uint64(0x1fffff)
    x = (x | (x << np.uint64(32))) & np.uint64(0x1f00000000ffff)
    x = (x | (x << np.uint64(16))) & np.uint64(0x1f0000ff0000ff)
    x = (x | (x << np.uint64(8))) & np.uint64(0x100f00f00f00f00f)
    x = (x | (x << np.uint64(4))) & np.uint64(0x10c30c30c30c30c3)
    x = (x | (x << np.uint64(2))) & np.uint64(0x1249249249249249)
    return x


def morton_np(P):
    """Morton codes of integer coordinates (x is the most significant bit of each triple, voxelize_pc.py:25-59)."""
    return (_spread3(P[:, 2]) | (_spread3(P[:, 1]) << np.uint64(1)) | (_spread3(P[:, 0]) << np.uint64(2))).astype(np.int64)


def sphere_point_cloud(n_points=3_000_000, depth=10, radius=230.0, seed=1001):
    """BASELINE config C1 (the encode_ply.py workload): `n_points` unit vectors, P = floor(2^(depth-1) + R d) clamped to the
    grid, deduplicated by Morton code (sorted unique voxels) -- a dense SURFACE, ~750 K voxels at R = 230, depth 10 --
    with uniform random 8-bit RGB per voxel.  Returns (V (nvox,3) int64 Morton-sorted, rgb (nvox,3) float64 in 0..255)."""
    rng = np.random.default_rng(seed)
    d = rng.normal(size=(n_points, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    half = 2 ** (depth - 1)
    P = np.clip(np.floor(half + radius * (half / 512.0) * d), 0, 2 ** depth - 1).astype(np.int64)
    mc = morton_np(P)
    _, first = np.unique(mc, return_index=True)
    V = P[first]
    rgb = rng.integers(0, 256, size=(V.shape[0], 3)).astype(np.float64)
    return V, rgb


def synthetic_gaussians_device(n, seed, device, color_dim=48):
    """The same distribution generated ON THE DEVICE with torch (seeded torch.Generator; ~100x faster than numpy for the
    50 M-Gaussian scene of config C4 and the 64 x 2 M scenes of config C5, which no parity test needs bit-identical on
    the host).  Returns CUDA tensors (means, quats, scales, opacities, colors), fp32."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    f = dict(device=device, dtype=torch.float32, generator=g)
    n_shell = int(0.7 * n)
    k = 64
    centres = torch.rand(k, 3, **f) * 1.4 - 0.7
    radii = torch.rand(k, 3, **f) * 0.25 + 0.05
    which = torch.randint(0, k, (n_shell,), device=device, generator=g)
    d = torch.randn(n_shell, 3, **f)
    d = d / d.norm(dim=1, keepdim=True)
    shell = centres[which] + radii[which] * d + 0.002 * torch.randn(n_shell, 3, **f)
    uni = torch.rand(n - n_shell, 3, **f) * 2 - 1
    means = torch.cat([shell, uni])[torch.randperm(n, device=device, generator=g)].contiguous()
    quats = torch.randn(n, 4, **f)
    quats = quats / quats.norm(dim=1, keepdim=True)
    scales = torch.exp(torch.randn(n, 3, **f) - 4.0)
    opac = torch.sigmoid(2.0 * torch.randn(n, **f))
    colors = torch.cat([0.5 * torch.randn(n, 3, **f), 0.1 * torch.randn(n, color_dim - 3, **f)], 1).contiguous()
    return means, quats, scales, opac, colors

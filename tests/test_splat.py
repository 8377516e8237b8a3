"""Render evaluator: CUDA splat rasteriser of raht_b200.render_eval against its numpy restatement
oracle/splat_oracle.py.  The reference hands this step to the un-vendored gsplat (python/quality_eval.py:331-346), so
the restatement is of the published pipeline, not of reference code: parity unpinned, tolerances below."""
import math

import numpy as np
import pytest
import torch

from oracle import splat_oracle as so

DEV = "cuda:0"


def scene(n, seed, cd=48, spread=0.35):
    rng = np.random.default_rng(seed)
    means = rng.normal(0, spread, (n, 3)).astype(np.float32)
    quats = rng.normal(0, 1, (n, 4)).astype(np.float32)
    scales = np.exp(rng.normal(-3.2, 0.6, (n, 3))).astype(np.float32)
    opac = rng.uniform(0.05, 1.0, n).astype(np.float32)
    colors = rng.normal(0, 0.25, (n, cd)).astype(np.float32)
    return means, quats, scales, opac, colors


def look_at(cam_pos, center=(0, 0, 0)):
    cam_pos, center = np.asarray(cam_pos, dtype=np.float32), np.asarray(center, dtype=np.float32)
    fwd = center - cam_pos
    fwd /= np.linalg.norm(fwd)
    right = np.cross(np.array([0, 1, 0], dtype=np.float32), fwd)
    right /= np.linalg.norm(right)
    up = np.cross(fwd, right)
    vm = np.eye(4, dtype=np.float32)
    vm[0, :3], vm[1, :3], vm[2, :3] = right, up, fwd
    vm[:3, 3] = -vm[:3, :3] @ cam_pos
    return vm


def intrinsics(w, h, f=1.2):
    return np.array([[w * f, 0, w / 2], [0, w * f, h / 2], [0, 0, 1]], dtype=np.float32)


# ---- the restatement itself (CPU) -------------------------------------------------------------------------------------
def test_oracle_single_gaussian_on_the_axis():
    """One isotropic Gaussian on the optical axis whose centre falls on a pixel centre: that pixel's colour is
    alpha * (C0 * sh0 + 0.5) + (1 - alpha) * background with alpha = opacity (exp(0) = 1)."""
    means = np.array([[0, 0, 0]], dtype=np.float32)
    quats = np.array([[1, 0, 0, 0]], dtype=np.float32)
    scales = np.full((1, 3), 0.05, dtype=np.float32)
    opac = np.array([0.6], dtype=np.float32)
    colors = np.zeros((1, 48), dtype=np.float32)
    colors[0, :3] = (1.0, -0.5, 0.25)
    vm = np.eye(4, dtype=np.float32)
    vm[2, 3] = 2.0                                                        # camera 2 in front of the Gaussian
    K = np.array([[40, 0, 8.5], [0, 40, 8.5], [0, 0, 1]], dtype=np.float32)
    img = so.render(means, quats, scales, opac, colors, vm, K, 17, 17, 3)
    exp = 0.6 * np.maximum(0.28209479177387814 * colors[0, :3] + 0.5, 0) + 0.4 * 1.0
    np.testing.assert_allclose(img[8, 8], exp, rtol=0, atol=1e-6)
    # the footprint: sigma_2d^2 = (f s / z)^2 + 0.3 = 1.3 pixel^2; one pixel to the right alpha falls by exp(-0.5 / 1.3)
    a1 = 0.6 * math.exp(-0.5 / 1.3)
    np.testing.assert_allclose(img[8, 9], a1 * exp / 0.6 - a1 * 0.4 / 0.6 + (1 - a1), rtol=0, atol=1e-5)
    assert np.all(img[0, 0] == 1.0)                                       # far from the splat: background


def test_oracle_occlusion_order_and_culling():
    """Two opaque Gaussians on the axis: the nearer one wins whatever their order in the arrays; one behind the camera
    leaves the background."""
    base = dict(quats=np.array([[1, 0, 0, 0]] * 2, dtype=np.float32), scales=np.full((2, 3), 0.2, dtype=np.float32),
                opacities=np.array([0.999, 0.999], dtype=np.float32))
    colors = np.zeros((2, 3), dtype=np.float32)
    colors[0], colors[1] = (1, 0, 0), (0, 0, 1)
    vm = np.eye(4, dtype=np.float32)
    vm[2, 3] = 3.0
    K = np.array([[30, 0, 8.5], [0, 30, 8.5], [0, 0, 1]], dtype=np.float32)
    for order in ((0, 1), (1, 0)):
        means = np.array([[0, 0, -1.0], [0, 0, 1.0]], dtype=np.float32)[list(order)]         # z = 2 (near) and z = 4 (far)
        img = so.render(means, base["quats"], base["scales"], base["opacities"], colors[list(order)], vm, K, 17, 17, None)
        assert np.argmax(img[8, 8]) == np.argmax(colors[0])                                  # the red one (z = 2) is in front
    behind = so.render(np.array([[0, 0, -5.0]], dtype=np.float32), base["quats"][:1], base["scales"][:1], base["opacities"][:1],
                       colors[:1], vm, K, 17, 17, None)
    assert np.all(behind == 1.0)


# ---- CUDA against the restatement ------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("n,w,h,cd,cam", [(1500, 100, 70, 48, (0.9, 0.5, -1.6)), (800, 64, 64, 48, (-1.2, -0.4, 1.1)),
                                          (600, 50, 33, 12, (0.2, 1.5, 0.9)), (400, 40, 40, 4, (0.0, 0.3, -1.4))])
def test_cuda_rasteriser_matches_the_restatement(n, w, h, cd, cam):
    from raht_b200 import render_eval
    means, quats, scales, opac, colors = scene(n, 7 + n, cd)
    vm, K = look_at(cam), intrinsics(w, h)
    sh_degree = int(np.sqrt(cd // 3) - 1) if cd % 3 == 0 else None
    exp = so.render(means, quats, scales, opac, colors, vm, K, w, h, sh_degree)
    info = {}
    got = render_eval.render_view(*[torch.from_numpy(a).to(DEV) for a in (means, quats, scales, opac, colors)], vm, K, w, h, sh_degree,
                                  info=info).cpu().numpy()
    P = so.project(means, quats, scales, opac, colors, vm, K, w, h, sh_degree)
    assert abs(info["n_isect"] - P["n_isect"]) <= max(2, P["n_isect"] // 2000)               # a radius may round the other way
    diff = np.abs(got - exp)
    # fp32 both sides; a Gaussian whose alpha sits on the 1/255 threshold, or whose extent rounds to another tile, may
    # enter on one side only: bounded by ~1/255 of a colour at isolated pixels
    assert diff.max() < 2e-2, diff.max()
    assert np.mean(diff > 1e-4) < 2e-3, np.mean(diff > 1e-4)
    assert diff.mean() < 2e-6 + 1e-5


@pytest.mark.gpu
def test_render_gaussians_and_psnr_interface():
    from raht_b200 import render_eval
    import quality_eval                                                                     # the reference-named shim
    means, quats, scales, opac, colors = [torch.from_numpy(a).to(DEV) for a in scene(3000, 3)]
    g = torch.Generator().manual_seed(11)
    viewmats, Ks, W, H = quality_eval.generate_random_cameras(means.mean(0), 2.0, n_views=3, image_width=96, image_height=80, device=DEV,
                                                              generator=g)
    assert viewmats.shape == (3, 4, 4) and Ks.shape == (3, 3, 3) and (W, H) == (96, 80)
    R = viewmats[:, :3, :3]
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3, device=DEV).expand(3, 3, 3), atol=1e-5)   # rigid view matrices
    imgs = quality_eval.render_gaussians(means, quats, scales, opac, colors, viewmats, Ks, W, H)
    assert imgs.shape == (3, 80, 96, 3) and imgs.dtype == torch.float32
    assert float(imgs.min()) >= 0.0 and float((imgs != 1.0).float().mean()) > 0.05            # something was drawn
    for v in range(3):
        exp = so.render(*[t.cpu().numpy() for t in (means, quats, scales, opac, colors)], viewmats[v].cpu().numpy(), Ks[v].cpu().numpy(),
                        W, H, 3)
        assert np.abs(imgs[v].cpu().numpy() - exp).max() < 2e-2
    assert quality_eval.compute_psnr(imgs[0], imgs[0]) == float("inf")
    a, b = imgs[0].cpu().double().numpy(), imgs[1].cpu().double().numpy()
    mse = np.mean((a - b) ** 2)
    assert abs(quality_eval.compute_psnr(imgs[0], imgs[1]) - 20 * math.log10(1 / math.sqrt(mse))) < 1e-6


@pytest.mark.gpu
def test_try_render_comparison_of_original_and_merged_scene(tmp_path):
    """The reference's use (test_voxelize_3dgs.py:300-340): original Gaussians against the voxel-merged ones."""
    import helpers
    from raht_b200 import codec, render_eval
    n = 20000
    m, q, s, o, c = [torch.from_numpy(a).to(DEV) for a in helpers.synthetic_gaussians(n, 5)]
    s = s * 0.05                                                                             # splats a few pixels wide
    orig = dict(means=m, quats=q, scales=s, opacities=o, colors=c)
    enc = codec.encode_scene(m, q, s, o, c, J=9, step=1.0 / 64, keep_packed=True)
    P = enc.packed
    merged = dict(means=P[:, 0:3].contiguous(), quats=P[:, 3:7].contiguous(), scales=P[:, 7:10].contiguous(),
                  opacities=P[:, 10].contiguous(), colors=P[:, 11:].contiguous())
    g = torch.Generator().manual_seed(2)
    res = render_eval.try_render_comparison(orig, merged, n_views=2, image_size=128, output_dir=str(tmp_path), generator=g)
    assert set(res) == {"psnr_avg", "psnr_std", "psnr_min", "psnr_max", "psnr_per_view", "original_render_time_ms", "merged_render_time_ms"}
    assert len(res["psnr_per_view"]) == 2 and 10.0 < res["psnr_min"] <= res["psnr_max"] < 80.0
    assert (tmp_path / "view_000_original.ppm").stat().st_size == len(b"P6\n128 128\n255\n") + 128 * 128 * 3
    same = render_eval.try_render_comparison(orig, orig, n_views=1, image_size=64, generator=torch.Generator().manual_seed(2))
    assert same["psnr_avg"] == float("inf")


@pytest.mark.gpu
def test_empty_views_and_argument_errors():
    from raht_b200 import render_eval
    from raht_b200._lib import lib
    means, quats, scales, opac, colors = [torch.from_numpy(a).to(DEV) for a in scene(64, 1)]
    vm = look_at((0, 0, -2.0))
    vm[2, 3] -= 100.0                                                                        # everything behind the camera
    img = render_eval.render_view(means, quats, scales, opac, colors, vm, intrinsics(48, 32), 48, 32, 3)
    assert torch.all(img == 1.0)
    img = render_eval.render_view(means[:0], quats[:0], scales[:0], opac[:0], colors[:0], look_at((0, 0, -2.0)), intrinsics(20, 20), 20, 20, 3,
                                  background=(0.25, 0.5, 0.75))
    assert torch.equal(img[3, 4].cpu(), torch.tensor([0.25, 0.5, 0.75]))
    with pytest.raises(RuntimeError):                                                        # 12 colour entries cannot hold degree 3
        render_eval.render_view(means, quats, scales, opac, colors[:, :12].contiguous(), look_at((0, 0, -2.0)), intrinsics(20, 20), 20, 20, 3)
    assert b"color_dim" in lib.raht_last_error()

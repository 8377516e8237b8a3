"""Render evaluator: the image-space check of python/quality_eval.py on the CUDA splat rasteriser.

Same function names, arguments and return values as the reference (`generate_random_cameras` :205-280,
`render_gaussians` :283-353, `compute_psnr` :356-370, `try_render_comparison` :373-520).  Where the reference calls
the third-party `gsplat.rasterization` (:331-346: SH colours of degree sqrt(K) - 1, white background, packed=False,
gsplat's defaults otherwise), this module calls libraht_b200.so: raht_splat_project -> raht_splat_tile_keys ->
raht_radix_sort_pairs -> raht_splat_tile_ranges -> raht_splat_rasterize, one view at a time as the reference's loop
does.  gsplat is not part of the reference tree: the arithmetic restates the published 3DGS pipeline (csrc/splat.cu),
parity with gsplat is unpinned.  CUDA tensors only."""
import ctypes as C
import math
import time
from typing import Dict, Tuple

import numpy as np
import torch

from . import _lib as L
from . import ops
from ._lib import lib, check, ptr, stream

TILE = 16


def generate_random_cameras(center: torch.Tensor, radius: float, n_views: int = 5, image_width: int = 512, image_height: int = 512,
                            device: str = 'cuda', generator: torch.Generator = None) -> Tuple[torch.Tensor, torch.Tensor, int, int]:
    """quality_eval.generate_random_cameras: random positions on a sphere of `radius` around `center`, looking at it
    (camera looks down +z), pinhole intrinsics with focal = 1.2 * width.  -> (viewmats [V,4,4], Ks [V,3,3], width, height).
    `generator` (not in the reference, which draws from the global CPU generator) makes the views reproducible."""
    center = center.detach().to(device=device, dtype=torch.float32)
    viewmats = []
    for _ in range(n_views):
        theta = torch.rand(1, generator=generator).item() * 2 * math.pi                    # azimuth
        phi = (torch.rand(1, generator=generator).item() * 0.5 + 0.25) * math.pi           # elevation (avoid poles)
        offs = torch.tensor([radius * math.sin(phi) * math.cos(theta), radius * math.sin(phi) * math.sin(theta), radius * math.cos(phi)],
                            device=device, dtype=torch.float32)
        cam_pos = center + offs
        forward = center - cam_pos
        forward = forward / torch.norm(forward)
        world_up = torch.tensor([0.0, 1.0, 0.0], device=device)
        right = torch.linalg.cross(world_up, forward)
        if torch.norm(right) < 0.001:                                                       # degenerate: looking along the up vector
            world_up = torch.tensor([0.0, 0.0, 1.0], device=device)
            right = torch.linalg.cross(world_up, forward)
        right = right / torch.norm(right)
        up = torch.linalg.cross(forward, right)
        w2c = torch.eye(4, device=device)
        w2c[0, :3] = right
        w2c[1, :3] = up
        w2c[2, :3] = forward
        w2c[:3, 3] = -torch.mv(w2c[:3, :3], cam_pos)
        viewmats.append(w2c)
    viewmats = torch.stack(viewmats)
    focal = image_width * 1.2
    K = torch.tensor([[focal, 0, image_width / 2], [0, focal, image_height / 2], [0, 0, 1]], device=device)
    return viewmats, K.unsqueeze(0).repeat(n_views, 1, 1), image_width, image_height


def render_view(means, quats, scales, opacities, colors, viewmat, K, width: int, height: int, sh_degree, background=(1.0, 1.0, 1.0),
                info: dict = None) -> torch.Tensor:
    """One view -> (height, width, 3) fp32 image on the device.  viewmat (4,4), K (3,3): tensors or arrays.
    colors (N, C): SH coefficients [k][rgb] when sh_degree >= 0, else RGB in its first three entries (sh_degree None)."""
    L.require_cuda(means, "means")
    dev = means.device
    f32 = lambda t: t.detach().to(dtype=torch.float32).contiguous()
    means, quats, scales, colors = f32(means), f32(quats), f32(scales), f32(colors)
    opacities = f32(opacities).reshape(-1)
    n, cd = means.shape[0], colors.shape[1]
    vm = np.asarray(viewmat.detach().cpu() if isinstance(viewmat, torch.Tensor) else viewmat, dtype=np.float32).reshape(16)
    Kh = np.asarray(K.detach().cpu() if isinstance(K, torch.Tensor) else K, dtype=np.float32).reshape(3, 3)
    vm_c = (C.c_float * 16)(*vm.tolist())
    bg_c = (C.c_float * 3)(*[float(x) for x in background])
    tiles_x, tiles_y = (width + TILE - 1) // TILE, (height + TILE - 1) // TILE
    n_tiles = tiles_x * tiles_y
    with torch.cuda.device(dev):
        ws_bytes = lib.raht_splat_workspace_bytes(n)
        ws = L.workspace_cached("splat", ws_bytes, dev)
        n_isect = torch.zeros(1, dtype=torch.int64, device=dev)
        check(lib.raht_splat_project(ptr(means), ptr(quats), ptr(scales), ptr(opacities), ptr(colors), cd, -1 if sh_degree is None else int(sh_degree),
                                     n, vm_c, float(Kh[0, 0]), float(Kh[1, 1]), float(Kh[0, 2]), float(Kh[1, 2]), width, height,
                                     ptr(ws), ws_bytes, ptr(n_isect), stream()), "raht_splat_project")
        m = int(L.read_host(n_isect)[0][0])                               # the one synchronisation: sizes of the sort
        image = torch.empty(height, width, 3, dtype=torch.float32, device=dev)
        ranges = torch.empty(n_tiles, 2, dtype=torch.int32, device=dev)
        keys = torch.empty(max(m, 1), dtype=torch.int64, device=dev)
        vals = torch.empty(max(m, 1), dtype=torch.int32, device=dev)
        check(lib.raht_splat_tile_keys(ptr(ws), ws_bytes, n, width, height, ptr(keys), ptr(vals), m, stream()), "raht_splat_tile_keys")
        if m:
            tile_bits = max(1, (n_tiles - 1).bit_length())
            keys, vals, _ = ops.radix_sort_pairs(keys[:m], 32 + tile_bits, vals[:m])
        check(lib.raht_splat_tile_ranges(ptr(keys), m, n_tiles, ptr(ranges), stream()), "raht_splat_tile_ranges")
        check(lib.raht_splat_rasterize(ptr(ws), ws_bytes, n, ptr(vals), ptr(ranges), width, height, bg_c, ptr(image), stream()),
              "raht_splat_rasterize")
    if info is not None:
        info["n_isect"] = m
    return image


def render_gaussians(means: torch.Tensor, quats: torch.Tensor, scales: torch.Tensor, opacities: torch.Tensor, colors: torch.Tensor,
                     viewmats: torch.Tensor, Ks: torch.Tensor, width: int, height: int) -> torch.Tensor:
    """quality_eval.render_gaussians: [V, H, W, 3] RGB images over a white background.  colors [N, 48] = 16 SH
    coefficients x RGB (:319-322); a colour dimension that is not a multiple of 3 uses its first three entries as RGB."""
    if colors.shape[1] % 3 == 0:
        K = colors.shape[1] // 3
        sh_degree = int(np.sqrt(K) - 1)
    else:
        sh_degree = None
    images = [render_view(means, quats, scales, opacities, colors, viewmats[i], Ks[i], width, height, sh_degree)
              for i in range(viewmats.shape[0])]
    return torch.stack(images)


def compute_psnr(img1: torch.Tensor, img2: torch.Tensor) -> float:
    """quality_eval.compute_psnr: images in [0, 1]; inf for identical images.  The squared error is summed in fp64 on the
    device (raht_sse)."""
    a, b = img1.reshape(-1, 1).contiguous(), img2.reshape(-1, 1).contiguous()
    mse = float(ops.sse(a, b).item()) / max(a.numel(), 1)
    if mse == 0:
        return float('inf')
    return 20 * math.log10(1.0 / math.sqrt(mse))


def try_render_comparison(original_params: Dict[str, torch.Tensor], merged_params: Dict[str, torch.Tensor], n_views: int = 5,
                          image_size: int = 512, output_dir: str = None, generator: torch.Generator = None, verbose: bool = False
                          ) -> Dict[str, float]:
    """quality_eval.try_render_comparison: both sets of Gaussians (dicts with means, quats, scales, opacities, colors) from
    the same random views; PSNR per view and its statistics, render times.  output_dir: the images are written as binary
    PPM files (the reference writes PNGs through PIL)."""
    dev = original_params['means'].device
    center = original_params['means'].mean(dim=0)
    bbox_size = (original_params['means'].max(dim=0)[0] - original_params['means'].min(dim=0)[0]).max().item()
    radius = bbox_size * 1.5
    viewmats, Ks, width, height = generate_random_cameras(center, radius, n_views, image_size, image_size, device=dev, generator=generator)

    def timed(p):
        torch.cuda.synchronize(dev)
        t0 = time.time()
        imgs = render_gaussians(p['means'], p['quats'], p['scales'], p['opacities'], p['colors'], viewmats, Ks, width, height)
        torch.cuda.synchronize(dev)
        return imgs, time.time() - t0

    original_images, original_time = timed(original_params)
    merged_images, merged_time = timed(merged_params)
    psnrs = [compute_psnr(original_images[i], merged_images[i]) for i in range(n_views)]
    if verbose:
        for i, p in enumerate(psnrs):
            print(f"    View {i + 1}: PSNR={p:.2f} dB")
    if output_dir is not None:
        import os
        os.makedirs(output_dir, exist_ok=True)
        for i in range(n_views):
            for name, img in (("original", original_images[i]), ("merged", merged_images[i])):
                u8 = (img.clamp(0, 1) * 255).to(torch.uint8).cpu().numpy()
                with open(os.path.join(output_dir, f"view_{i:03d}_{name}.ppm"), "wb") as f:
                    f.write(f"P6\n{width} {height}\n255\n".encode())
                    f.write(u8.tobytes())
    with np.errstate(invalid='ignore'):                                  # identical images: inf - inf in the deviation
        std = float(np.std(psnrs))
    return {'psnr_avg': float(np.mean(psnrs)), 'psnr_std': std, 'psnr_min': float(np.min(psnrs)),
            'psnr_max': float(np.max(psnrs)), 'psnr_per_view': psnrs, 'original_render_time_ms': original_time * 1000,
            'merged_render_time_ms': merged_time * 1000}

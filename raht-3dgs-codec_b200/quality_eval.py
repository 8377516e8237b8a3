"""Drop-in for the functions of the reference's python/quality_eval.py that touch the GPU.

save_ply (quality_eval.py:18-117) writes the scene through one device buffer (raht_b200.ply_io); the render comparison
(generate_random_cameras :205-280, render_gaussians :283-353, compute_psnr :356-370, try_render_comparison :373-520) runs
on the CUDA splat rasteriser of this package (raht_b200.render_eval, csrc/splat.cu) where the reference calls the
third-party gsplat.  CUDA tensors only: this package has no CPU path."""
from raht_b200.ply_io import save_ply                                                    # noqa: F401
from raht_b200.render_eval import (generate_random_cameras, render_gaussians, compute_psnr,   # noqa: F401
                                   try_render_comparison)

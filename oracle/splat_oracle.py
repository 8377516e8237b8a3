"""TEST INFRASTRUCTURE -- numpy restatement of the splat rasterisation the render evaluator performs.

The reference (python/quality_eval.py:283-353) hands the rendering to the third-party gsplat.rasterization, which is
not vendored in the reference tree: PARITY UNPINNED.  What is restated is the published 3DGS / gsplat "classic"
pipeline with gsplat's default parameters, written pixel by pixel and Gaussian by Gaussian (no tiles sorted by key, no
batches): a Gaussian contributes to a pixel when the pixel's 16 x 16 tile lies in its tile rectangle, in order of
camera depth (ties: index).  fp32 throughout, one operation at a time, so that the CUDA kernels (built --fmad=false)
can be compared closely.  Only tests/ import this."""
import numpy as np

f32 = np.float32
NEAR, FAR, EPS2D, TILE = f32(0.01), f32(1e10), f32(0.3), 16
ALPHA_MIN, ALPHA_MAX, T_MIN = f32(1.0) / f32(255.0), f32(0.999), f32(1e-4)
C0 = f32(0.28209479177387814)
C1 = f32(0.4886025119029199)
C2 = [f32(v) for v in (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)]
C3 = [f32(v) for v in (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
                       1.445305721320277, -0.5900435899266435)]


def sh_colors(sh, dirs, deg):
    """sh (N, K, 3), unit dirs (N, 3) -> (N, 3) = SH(dir), 3DGS convention."""
    x, y, z = (dirs[:, i:i + 1] for i in range(3))
    r = C0 * sh[:, 0]
    if deg >= 1:
        r = r + (-C1 * y * sh[:, 1] + C1 * z * sh[:, 2] - C1 * x * sh[:, 3])
    if deg >= 2:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        r = r + (C2[0] * xy * sh[:, 4] + C2[1] * yz * sh[:, 5] + C2[2] * (f32(2) * zz - xx - yy) * sh[:, 6] + C2[3] * xz * sh[:, 7] +
                 C2[4] * (xx - yy) * sh[:, 8])
    if deg >= 3:
        r = r + (C3[0] * y * (f32(3) * xx - yy) * sh[:, 9] + C3[1] * xy * z * sh[:, 10] + C3[2] * y * (f32(4) * zz - xx - yy) * sh[:, 11] +
                 C3[3] * z * (f32(2) * zz - f32(3) * xx - f32(3) * yy) * sh[:, 12] + C3[4] * x * (f32(4) * zz - xx - yy) * sh[:, 13] +
                 C3[5] * z * (xx - yy) * sh[:, 14] + C3[6] * x * (xx - f32(3) * yy) * sh[:, 15])
    return r.astype(f32)


def project(means, quats, scales, opacities, colors, viewmat, K, width, height, sh_degree):
    """-> dict of per-Gaussian records (xy, conic, opacity, rgb, depth, rect, valid)."""
    means, quats, scales = means.astype(f32), quats.astype(f32), scales.astype(f32)
    R, t = viewmat[:3, :3].astype(f32), viewmat[:3, 3].astype(f32)
    fx, fy, cx, cy = (f32(K[0, 0]), f32(K[1, 1]), f32(K[0, 2]), f32(K[1, 2]))
    pc = (means @ R.T + t).astype(f32)
    x, y, z = pc[:, 0], pc[:, 1], pc[:, 2]
    n = means.shape[0]
    valid = (z >= NEAR) & (z <= FAR)
    q = quats / np.sqrt((quats * quats).sum(1, keepdims=True)).astype(f32)
    w_, x_, y_, z_ = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    Rq = np.stack([1 - 2 * (y_ * y_ + z_ * z_), 2 * (x_ * y_ - w_ * z_), 2 * (x_ * z_ + w_ * y_),
                   2 * (x_ * y_ + w_ * z_), 1 - 2 * (x_ * x_ + z_ * z_), 2 * (y_ * z_ - w_ * x_),
                   2 * (x_ * z_ - w_ * y_), 2 * (y_ * z_ + w_ * x_), 1 - 2 * (x_ * x_ + y_ * y_)], 1).reshape(n, 3, 3).astype(f32)
    M = (Rq * scales[:, None, :]).astype(f32)
    W = np.einsum('ab,nbc->nac', R, M).astype(f32)
    S = np.einsum('nab,ncb->nac', W, W).astype(f32)
    tanx, tany = f32(0.5) * f32(width) / fx, f32(0.5) * f32(height) / fy
    lxp, lxn = (f32(width) - cx) / fx + f32(0.3) * tanx, cx / fx + f32(0.3) * tanx
    lyp, lyn = (f32(height) - cy) / fy + f32(0.3) * tany, cy / fy + f32(0.3) * tany
    zs = np.where(valid, z, f32(1))
    rz = (f32(1) / zs).astype(f32)
    rz2 = rz * rz
    tx = zs * np.minimum(lxp, np.maximum(-lxn, x * rz))
    ty = zs * np.minimum(lyp, np.maximum(-lyn, y * rz))
    J = np.zeros((n, 2, 3), dtype=f32)
    J[:, 0, 0], J[:, 0, 2], J[:, 1, 1], J[:, 1, 2] = fx * rz, -fx * tx * rz2, fy * rz, -fy * ty * rz2
    cov = np.einsum('nab,nbc,ndc->nad', J, S, J).astype(f32)
    c00, c01, c11 = cov[:, 0, 0] + EPS2D, cov[:, 0, 1], cov[:, 1, 1] + EPS2D
    det = c00 * c11 - c01 * c01
    valid &= det > 0
    dets = np.where(det > 0, det, f32(1))
    px, py = fx * x * rz + cx, fy * y * rz + cy
    b = f32(0.5) * (c00 + c11)
    lam = b + np.sqrt(np.maximum(f32(0.01), b * b - dets))
    radius = np.ceil(f32(3) * np.sqrt(lam)).astype(f32)
    valid &= (px + radius > 0) & (px - radius < width) & (py + radius > 0) & (py - radius < height)
    tiles_x, tiles_y = (width + TILE - 1) // TILE, (height + TILE - 1) // TILE
    rect = np.stack([np.clip(np.floor((px - radius) / TILE), 0, tiles_x), np.clip(np.floor((py - radius) / TILE), 0, tiles_y),
                     np.clip(np.ceil((px + radius) / TILE), 0, tiles_x), np.clip(np.ceil((py + radius) / TILE), 0, tiles_y)], 1)
    rect = np.where(valid[:, None], np.nan_to_num(rect), 0).astype(np.int64)
    valid &= (rect[:, 2] > rect[:, 0]) & (rect[:, 3] > rect[:, 1])
    if sh_degree is None:
        rgb = colors[:, :3].astype(f32)
    else:
        campos = (-R.T @ t).astype(f32)
        d = (means - campos).astype(f32)
        d = d / np.sqrt(np.maximum((d * d).sum(1, keepdims=True), f32(1e-24))).astype(f32)
        rgb = np.maximum(sh_colors(colors.reshape(n, -1, 3).astype(f32), d, sh_degree) + f32(0.5), f32(0))
    conic = np.stack([c11 / dets, -c01 / dets, c00 / dets], 1).astype(f32)
    return {"xy": np.stack([px, py], 1).astype(f32), "conic": conic, "opacity": opacities.reshape(-1).astype(f32), "rgb": rgb.astype(f32),
            "depth": z.astype(f32), "rect": rect, "valid": valid, "n_isect": int(((rect[:, 2] - rect[:, 0]) * (rect[:, 3] - rect[:, 1]))[valid].sum())}


def render(means, quats, scales, opacities, colors, viewmat, K, width, height, sh_degree, background=(1.0, 1.0, 1.0)):
    """-> (height, width, 3) fp32.  Loops over pixels: small images only."""
    P = project(means, quats, scales, opacities, colors, np.asarray(viewmat, dtype=f32), np.asarray(K, dtype=f32), width, height, sh_degree)
    ids = np.nonzero(P["valid"])[0]
    ids = ids[np.argsort(P["depth"][ids].view(np.uint32), kind="stable")]
    bg = np.asarray(background, dtype=f32)
    img = np.zeros((height, width, 3), dtype=f32)
    rect = P["rect"]
    for ty in range((height + TILE - 1) // TILE):
        for tx in range((width + TILE - 1) // TILE):
            sel = ids[(rect[ids, 0] <= tx) & (tx < rect[ids, 2]) & (rect[ids, 1] <= ty) & (ty < rect[ids, 3])]
            xy, co, op, rgb = P["xy"][sel], P["conic"][sel], P["opacity"][sel], P["rgb"][sel]
            for yy in range(ty * TILE, min(height, (ty + 1) * TILE)):
                for xx in range(tx * TILE, min(width, (tx + 1) * TILE)):
                    px, py = f32(xx) + f32(0.5), f32(yy) + f32(0.5)
                    dx, dy = xy[:, 0] - px, xy[:, 1] - py
                    sigma = f32(0.5) * (co[:, 0] * dx * dx + co[:, 2] * dy * dy) + co[:, 1] * dx * dy
                    alpha = np.minimum(ALPHA_MAX, op * np.exp(-sigma).astype(f32))
                    T, acc = f32(1), np.zeros(3, dtype=f32)
                    for j in np.nonzero((sigma >= 0) & (alpha >= ALPHA_MIN))[0]:
                        nT = T * (f32(1) - alpha[j])
                        if nT <= T_MIN:
                            break
                        acc = acc + rgb[j] * (alpha[j] * T)
                        T = nT
                    img[yy, xx] = acc + T * bg
    return img

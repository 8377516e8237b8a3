// splat.cu -- render evaluator: 3D Gaussian splatting as CUDA compute, no graphics API.
// The reference renders the original and the merged / decoded Gaussians from random views and compares the images
// (python/quality_eval.py:283-353 render_gaussians, :356-370 compute_psnr, :373-520 try_render_comparison).  Its
// rasteriser is the third-party `gsplat.rasterization` (not vendored in the reference tree, not installed here): the
// arithmetic below restates the published 3DGS / gsplat "classic" pipeline with gsplat's default parameters --
// near plane 0.01, eps2d 0.3, 3-sigma extent, 16 x 16 tiles, alpha threshold 1/255, alpha cap 0.999, transmittance
// cut 1e-4, colours = max(SH(dir) + 0.5, 0) -- and is checked against a numpy restatement in the test tree, not against
// gsplat ("parity unpinned").
//
//   project     a thread per Gaussian: camera transform, near cull, covariance from quaternion + scale, EWA
//               projection, conic, extent, tile rectangle, SH colour of the view direction; tile counts per block
//   tile_keys   (tile << 32 | depth bits, Gaussian) for every tile a Gaussian touches  -> raht_radix_sort_pairs
//   tile_ranges first / last sorted entry of every tile
//   rasterize   a CTA per 16 x 16 tile: the tile's Gaussians, front to back, staged through shared memory 256 at a
//               time; a thread per pixel blends until its transmittance is spent
#include "common.cuh"

namespace {

constexpr int SP_THREADS = 256;
constexpr int SP_TILE = 16;
constexpr float SP_NEAR = 0.01f, SP_FAR = 1e10f, SP_EPS2D = 0.3f;
constexpr float SP_ALPHA_MIN = 1.0f / 255.0f, SP_ALPHA_MAX = 0.999f, SP_T_MIN = 1e-4f;

struct Camera {
    float R[9], t[3];            // world -> camera (rows of the view matrix)
    float pos[3];                // camera position in the world
    float fx, fy, cx, cy;
    int width, height, tiles_x, tiles_y;
};

struct SplatBufs {
    float2 *xy;                  // [n] pixel coordinates of the centre
    float4 *conic_opac;          // [n] conic a, b, c and opacity
    float4 *rgb_depth;           // [n] colour and camera depth
    int4 *rect;                  // [n] tile rectangle x0, y0, x1, y1 (x1 <= x0: culled)
    unsigned long long *block_counts;   // [blocks + 1] tiles touched per block of SP_THREADS Gaussians, then its scan
    size_t total;
};
SplatBufs splat_layout(void *ws, int64_t n)
{
    SplatBufs B;
    char *p = (char *)ws;
    size_t o = 0;
    const size_t nn = (size_t)(n > 0 ? n : 0);
    B.xy = (float2 *)(p + o); o += align_up(nn * 8, 256);
    B.conic_opac = (float4 *)(p + o); o += align_up(nn * 16, 256);
    B.rgb_depth = (float4 *)(p + o); o += align_up(nn * 16, 256);
    B.rect = (int4 *)(p + o); o += align_up(nn * 16, 256);
    B.block_counts = (unsigned long long *)(p + o); o += align_up(((nn + SP_THREADS - 1) / SP_THREADS + 2) * 8, 256);
    B.total = o + 256;
    return B;
}

// real spherical harmonics up to degree 3 in the 3DGS convention (16 coefficients per colour channel, coefficient-major)
__device__ __forceinline__ float sh_channel(const float *__restrict__ sh, int stride, int deg, float x, float y, float z)
{
    float r = 0.28209479177387814f * sh[0];
    if (deg >= 1) {
        r += -0.4886025119029199f * y * sh[1 * stride] + 0.4886025119029199f * z * sh[2 * stride] - 0.4886025119029199f * x * sh[3 * stride];
        if (deg >= 2) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            r += 1.0925484305920792f * xy * sh[4 * stride] - 1.0925484305920792f * yz * sh[5 * stride] +
                 0.31539156525252005f * (2.0f * zz - xx - yy) * sh[6 * stride] - 1.0925484305920792f * xz * sh[7 * stride] +
                 0.5462742152960396f * (xx - yy) * sh[8 * stride];
            if (deg >= 3) {
                r += -0.5900435899266435f * y * (3.0f * xx - yy) * sh[9 * stride] + 2.890611442640554f * xy * z * sh[10 * stride] -
                     0.4570457994644658f * y * (4.0f * zz - xx - yy) * sh[11 * stride] +
                     0.3731763325901154f * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12 * stride] -
                     0.4570457994644658f * x * (4.0f * zz - xx - yy) * sh[13 * stride] + 1.445305721320277f * z * (xx - yy) * sh[14 * stride] -
                     0.5900435899266435f * x * (xx - 3.0f * yy) * sh[15 * stride];
            }
        }
    }
    return r;
}

__global__ void __launch_bounds__(SP_THREADS) splat_project_kernel(const float *__restrict__ means, const float *__restrict__ quats,
                                                                  const float *__restrict__ scales, const float *__restrict__ opac,
                                                                  const float *__restrict__ colors, int cd, int sh_degree, int64_t n,
                                                                  Camera C, SplatBufs B)
{
    __shared__ unsigned int s_cnt[SP_THREADS / 32];
    const int64_t i = (int64_t)blockIdx.x * SP_THREADS + threadIdx.x;
    unsigned int ntiles = 0;
    if (i < n) {
        const float mx = means[i * 3], my = means[i * 3 + 1], mz = means[i * 3 + 2];
        const float x = C.R[0] * mx + C.R[1] * my + C.R[2] * mz + C.t[0];
        const float y = C.R[3] * mx + C.R[4] * my + C.R[5] * mz + C.t[1];
        const float z = C.R[6] * mx + C.R[7] * my + C.R[8] * mz + C.t[2];
        int4 rect = make_int4(0, 0, 0, 0);
        if (z >= SP_NEAR && z <= SP_FAR) {
            // world covariance M M^T, M = R(q) diag(s); the quaternion (w, x, y, z) is normalised first (quality_eval.py:335)
            float qw = quats[i * 4], qx = quats[i * 4 + 1], qy = quats[i * 4 + 2], qz = quats[i * 4 + 3];
            const float qn = rsqrtf(qw * qw + qx * qx + qy * qy + qz * qz);
            qw *= qn; qx *= qn; qy *= qn; qz *= qn;
            const float sx = scales[i * 3], sy = scales[i * 3 + 1], sz = scales[i * 3 + 2];
            const float r00 = 1.f - 2.f * (qy * qy + qz * qz), r01 = 2.f * (qx * qy - qw * qz), r02 = 2.f * (qx * qz + qw * qy);
            const float r10 = 2.f * (qx * qy + qw * qz), r11 = 1.f - 2.f * (qx * qx + qz * qz), r12 = 2.f * (qy * qz - qw * qx);
            const float r20 = 2.f * (qx * qz - qw * qy), r21 = 2.f * (qy * qz + qw * qx), r22 = 1.f - 2.f * (qx * qx + qy * qy);
            // W = R_cam R(q) diag(s): camera-space square root of the covariance
            float W[9];
            const float M[9] = {r00 * sx, r01 * sy, r02 * sz, r10 * sx, r11 * sy, r12 * sz, r20 * sx, r21 * sy, r22 * sz};
            for (int a = 0; a < 3; ++a)
                for (int b = 0; b < 3; ++b) W[a * 3 + b] = C.R[a * 3] * M[b] + C.R[a * 3 + 1] * M[3 + b] + C.R[a * 3 + 2] * M[6 + b];
            float S[6];           // camera covariance W W^T: 00 01 02 11 12 22
            S[0] = W[0] * W[0] + W[1] * W[1] + W[2] * W[2];
            S[1] = W[0] * W[3] + W[1] * W[4] + W[2] * W[5];
            S[2] = W[0] * W[6] + W[1] * W[7] + W[2] * W[8];
            S[3] = W[3] * W[3] + W[4] * W[4] + W[5] * W[5];
            S[4] = W[3] * W[6] + W[4] * W[7] + W[5] * W[8];
            S[5] = W[6] * W[6] + W[7] * W[7] + W[8] * W[8];
            // perspective Jacobian with the centre clamped to 1.3 x the field of view
            const float tanx = 0.5f * (float)C.width / C.fx, tany = 0.5f * (float)C.height / C.fy;
            const float lxp = ((float)C.width - C.cx) / C.fx + 0.3f * tanx, lxn = C.cx / C.fx + 0.3f * tanx;
            const float lyp = ((float)C.height - C.cy) / C.fy + 0.3f * tany, lyn = C.cy / C.fy + 0.3f * tany;
            const float rz = 1.0f / z, rz2 = rz * rz;
            const float tx = z * fminf(lxp, fmaxf(-lxn, x * rz)), ty = z * fminf(lyp, fmaxf(-lyn, y * rz));
            const float j00 = C.fx * rz, j02 = -C.fx * tx * rz2, j11 = C.fy * rz, j12 = -C.fy * ty * rz2;
            // cov2d = J S J^T
            const float a0 = j00 * S[0] + j02 * S[2], a1 = j00 * S[1] + j02 * S[4], a2 = j00 * S[2] + j02 * S[5];
            const float b1 = j11 * S[3] + j12 * S[4], b2 = j11 * S[4] + j12 * S[5];
            float c00 = a0 * j00 + a2 * j02, c01 = a1 * j11 + a2 * j12, c11 = b1 * j11 + b2 * j12;
            c00 += SP_EPS2D;
            c11 += SP_EPS2D;
            const float det = c00 * c11 - c01 * c01;
            if (det > 0.0f) {
                const float px = C.fx * x * rz + C.cx, py = C.fy * y * rz + C.cy;
                const float bmid = 0.5f * (c00 + c11);
                const float lam = bmid + sqrtf(fmaxf(0.01f, bmid * bmid - det));
                const float radius = ceilf(3.0f * sqrtf(lam));
                if (px + radius > 0.0f && px - radius < (float)C.width && py + radius > 0.0f && py - radius < (float)C.height) {
                    rect.x = min(C.tiles_x, max(0, (int)floorf((px - radius) / (float)SP_TILE)));
                    rect.y = min(C.tiles_y, max(0, (int)floorf((py - radius) / (float)SP_TILE)));
                    rect.z = min(C.tiles_x, max(0, (int)ceilf((px + radius) / (float)SP_TILE)));
                    rect.w = min(C.tiles_y, max(0, (int)ceilf((py + radius) / (float)SP_TILE)));
                    if (rect.z > rect.x && rect.w > rect.y) {
                        ntiles = (unsigned int)((rect.z - rect.x) * (rect.w - rect.y));
                        const float inv = 1.0f / det;
                        B.xy[i] = make_float2(px, py);
                        B.conic_opac[i] = make_float4(c11 * inv, -c01 * inv, c00 * inv, opac[i]);
                        float3 col;
                        if (sh_degree >= 0) {
                            float dx = mx - C.pos[0], dy = my - C.pos[1], dz = mz - C.pos[2];
                            const float dn = rsqrtf(fmaxf(dx * dx + dy * dy + dz * dz, 1e-24f));
                            dx *= dn; dy *= dn; dz *= dn;
                            const float *sh = colors + (size_t)i * cd;          // [coefficient][rgb]
                            col.x = fmaxf(sh_channel(sh, 3, sh_degree, dx, dy, dz) + 0.5f, 0.0f);
                            col.y = fmaxf(sh_channel(sh + 1, 3, sh_degree, dx, dy, dz) + 0.5f, 0.0f);
                            col.z = fmaxf(sh_channel(sh + 2, 3, sh_degree, dx, dy, dz) + 0.5f, 0.0f);
                        } else {
                            col = make_float3(colors[(size_t)i * cd], colors[(size_t)i * cd + 1], colors[(size_t)i * cd + 2]);
                        }
                        B.rgb_depth[i] = make_float4(col.x, col.y, col.z, z);
                    } else {
                        rect = make_int4(0, 0, 0, 0);
                    }
                }
            }
        }
        B.rect[i] = rect;
    }
    unsigned int v = ntiles;
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(~0u, v, o);
    if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < SP_THREADS / 32; ++w) t += s_cnt[w];
        B.block_counts[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(SP_THREADS) splat_keys_kernel(SplatBufs B, int64_t n, int tiles_x, int64_t m,
                                                               unsigned long long *__restrict__ keys, int32_t *__restrict__ vals)
{
    __shared__ unsigned int s_w[SP_THREADS / 32];
    const int64_t i = (int64_t)blockIdx.x * SP_THREADS + threadIdx.x;
    int4 rect = make_int4(0, 0, 0, 0);
    if (i < n) rect = B.rect[i];
    const unsigned int mine = (unsigned int)(max(rect.z - rect.x, 0) * max(rect.w - rect.y, 0));
    unsigned int x = mine;                                   // inclusive scan over the block
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int y = __shfl_up_sync(~0u, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) s_w[warp] = x;
    __syncthreads();
    unsigned int before = 0;
    for (int w = 0; w < warp; ++w) before += s_w[w];
    unsigned long long at = B.block_counts[blockIdx.x] + before + (x - mine);
    if (!mine) return;
    const unsigned int depth_bits = __float_as_uint(B.rgb_depth[i].w);     // positive floats order like their bits
    for (int ty = rect.y; ty < rect.w; ++ty)
        for (int tx = rect.x; tx < rect.z; ++tx) {
            if ((int64_t)at < m) {
                keys[at] = ((unsigned long long)(unsigned int)(ty * tiles_x + tx) << 32) | depth_bits;
                vals[at] = (int32_t)i;
            }
            ++at;
        }
}

__global__ void __launch_bounds__(SP_THREADS) splat_ranges_kernel(const unsigned long long *__restrict__ keys, int64_t m, int n_tiles,
                                                                 int32_t *__restrict__ ranges)
{
    const int64_t i = (int64_t)blockIdx.x * SP_THREADS + threadIdx.x;
    if (i >= m) return;
    const int t = (int)(keys[i] >> 32);
    if (t < 0 || t >= n_tiles) return;
    const int prev = i ? (int)(keys[i - 1] >> 32) : -1;
    if (t != prev) {
        ranges[2 * t] = (int32_t)i;
        if (prev >= 0 && prev < n_tiles) ranges[2 * prev + 1] = (int32_t)i;
    }
    if (i == m - 1) ranges[2 * t + 1] = (int32_t)m;
}

__global__ void __launch_bounds__(SP_TILE *SP_TILE) splat_raster_kernel(SplatBufs B, const int32_t *__restrict__ ids,
                                                                       const int32_t *__restrict__ ranges, int width, int height,
                                                                       int tiles_x, float3 bg, float *__restrict__ image)
{
    __shared__ float2 s_xy[SP_THREADS];
    __shared__ float4 s_co[SP_THREADS];
    __shared__ float4 s_rgb[SP_THREADS];
    const int tile = blockIdx.y * tiles_x + blockIdx.x;
    const int px_i = blockIdx.x * SP_TILE + (threadIdx.x % SP_TILE), py_i = blockIdx.y * SP_TILE + (threadIdx.x / SP_TILE);
    const bool inside = px_i < width && py_i < height;
    const float px = (float)px_i + 0.5f, py = (float)py_i + 0.5f;
    const int lo = ranges[2 * tile], hi = ranges[2 * tile + 1];
    float T = 1.0f, r = 0.0f, g = 0.0f, b = 0.0f;
    bool done = !inside;
    for (int base = lo; base < hi; base += SP_THREADS) {
        if (__syncthreads_count(!done) == 0) break;
        const int k = base + (int)threadIdx.x;
        if (k < hi) {
            const int id = ids[k];
            s_xy[threadIdx.x] = B.xy[id];
            s_co[threadIdx.x] = B.conic_opac[id];
            s_rgb[threadIdx.x] = B.rgb_depth[id];
        }
        __syncthreads();
        const int cnt = min(SP_THREADS, hi - base);
        for (int j = 0; j < cnt && !done; ++j) {
            const float2 c = s_xy[j];
            const float4 co = s_co[j];
            const float dx = c.x - px, dy = c.y - py;
            const float sigma = 0.5f * (co.x * dx * dx + co.z * dy * dy) + co.y * dx * dy;
            const float alpha = fminf(SP_ALPHA_MAX, co.w * expf(-sigma));
            if (sigma < 0.0f || alpha < SP_ALPHA_MIN) continue;
            const float nT = T * (1.0f - alpha);
            if (nT <= SP_T_MIN) { done = true; break; }
            const float4 col = s_rgb[j];
            const float w = alpha * T;
            r += col.x * w; g += col.y * w; b += col.z * w;
            T = nT;
        }
    }
    if (inside) {
        float *o = image + ((size_t)py_i * width + px_i) * 3;
        o[0] = r + T * bg.x;
        o[1] = g + T * bg.y;
        o[2] = b + T * bg.z;
    }
}

Camera make_camera(const float *viewmat16, float fx, float fy, float cx, float cy, int width, int height)
{
    Camera C;
    for (int a = 0; a < 3; ++a) {
        for (int b = 0; b < 3; ++b) C.R[a * 3 + b] = viewmat16[a * 4 + b];
        C.t[a] = viewmat16[a * 4 + 3];
    }
    for (int a = 0; a < 3; ++a) C.pos[a] = -(C.R[a] * C.t[0] + C.R[3 + a] * C.t[1] + C.R[6 + a] * C.t[2]);    // -R^T t
    C.fx = fx; C.fy = fy; C.cx = cx; C.cy = cy;
    C.width = width; C.height = height;
    C.tiles_x = (width + SP_TILE - 1) / SP_TILE;
    C.tiles_y = (height + SP_TILE - 1) / SP_TILE;
    return C;
}

}  // namespace

RAHT_API size_t raht_splat_workspace_bytes(int64_t n) { return splat_layout(nullptr, n).total; }

RAHT_API int raht_splat_project(const float *means, const float *quats, const float *scales, const float *opacities, const float *colors,
                                int color_dim, int sh_degree, int64_t n, const float *viewmat16_host, float fx, float fy, float cx, float cy,
                                int width, int height, void *workspace, size_t workspace_bytes, int64_t *n_isect, void *stream)
{
    RAHT_REQUIRE(n >= 0 && n < ((int64_t)1 << 31) && width > 0 && height > 0 && width <= 16384 && height <= 16384 && viewmat16_host &&
                     n_isect && workspace && fx > 0 && fy > 0,
                 "raht_splat_project: bad arguments");
    RAHT_REQUIRE(sh_degree >= -1 && sh_degree <= 3 && color_dim >= (sh_degree < 0 ? 3 : 3 * (sh_degree + 1) * (sh_degree + 1)),
                 "raht_splat_project: color_dim does not hold the coefficients of sh_degree");
    RAHT_REQUIRE(n == 0 || (means && quats && scales && opacities && colors), "raht_splat_project: null pointer");
    if (workspace_bytes < raht_splat_workspace_bytes(n) || ((uintptr_t)workspace & 255)) {
        raht_set_error("raht_splat_project: workspace too small or not 256-byte aligned");
        return RAHT_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    SplatBufs B = splat_layout(workspace, n);
    const Camera C = make_camera(viewmat16_host, fx, fy, cx, cy, width, height);
    const int64_t nb = (n + SP_THREADS - 1) / SP_THREADS;
    raht_count_launches(n ? 2 : 1);
    if (n) splat_project_kernel<<<(unsigned)nb, SP_THREADS, 0, st>>>(means, quats, scales, opacities, colors, color_dim, sh_degree, n, C, B);
    scan_counts_kernel<<<1, 1024, 0, st>>>(B.block_counts, nb, n_isect);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_splat_tile_keys(void *workspace, size_t workspace_bytes, int64_t n, int width, int height, uint64_t *keys, int32_t *vals,
                                  int64_t m, void *stream)
{
    RAHT_REQUIRE(n >= 0 && m >= 0 && workspace && width > 0 && height > 0, "raht_splat_tile_keys: bad arguments");
    RAHT_REQUIRE(workspace_bytes >= raht_splat_workspace_bytes(n), "raht_splat_tile_keys: workspace too small");
    if (n == 0 || m == 0) return RAHT_OK;
    RAHT_REQUIRE(keys && vals, "raht_splat_tile_keys: null pointer");
    SplatBufs B = splat_layout(workspace, n);
    raht_count_launches(1);
    splat_keys_kernel<<<(unsigned)((n + SP_THREADS - 1) / SP_THREADS), SP_THREADS, 0, (cudaStream_t)stream>>>(
        B, n, (width + SP_TILE - 1) / SP_TILE, m, (unsigned long long *)keys, vals);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_splat_tile_ranges(const uint64_t *sorted_keys, int64_t m, int n_tiles, int32_t *ranges, void *stream)
{
    RAHT_REQUIRE(m >= 0 && n_tiles > 0 && ranges, "raht_splat_tile_ranges: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    RAHT_CUDA(cudaMemsetAsync(ranges, 0, (size_t)n_tiles * 8, st));
    if (m == 0) return RAHT_OK;
    RAHT_REQUIRE(sorted_keys, "raht_splat_tile_ranges: null pointer");
    raht_count_launches(1);
    splat_ranges_kernel<<<(unsigned)((m + SP_THREADS - 1) / SP_THREADS), SP_THREADS, 0, st>>>((const unsigned long long *)sorted_keys, m, n_tiles,
                                                                                           ranges);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_splat_rasterize(void *workspace, size_t workspace_bytes, int64_t n, const int32_t *sorted_vals, const int32_t *ranges,
                                  int width, int height, const float *background3_host, float *image, void *stream)
{
    RAHT_REQUIRE(n >= 0 && workspace && ranges && image && width > 0 && height > 0 && background3_host, "raht_splat_rasterize: bad arguments");
    RAHT_REQUIRE(workspace_bytes >= raht_splat_workspace_bytes(n), "raht_splat_rasterize: workspace too small");
    SplatBufs B = splat_layout(workspace, n);
    const int tx = (width + SP_TILE - 1) / SP_TILE, ty = (height + SP_TILE - 1) / SP_TILE;
    raht_count_launches(1);
    splat_raster_kernel<<<dim3((unsigned)tx, (unsigned)ty), SP_TILE * SP_TILE, 0, (cudaStream_t)stream>>>(
        B, sorted_vals, ranges, width, height, tx, make_float3(background3_host[0], background3_host[1], background3_host[2]), image);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

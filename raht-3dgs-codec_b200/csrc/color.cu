// color.cu -- RGB -> YUV of the point-cloud path (SURVEY 8 row g: python/encode_ply.py:108 calls utils.rgb_to_yuv before
// RAHT).  Replaces python/utils.py:4-33: yuv = clamp([r g b 1] / 255 @ Q, 0, 1) * 255 with the BT.709 full-range matrix of
// utils.py:17-22, in fp64.  The reference evaluates the 4-term products as a torch matmul (BLAS: summation order and FMA
// contraction unspecified), so this is a floating-point row: tolerance 1e-10 absolute on values in [0, 255]
// (tests/golden/rgb_to_yuv.npz), not bit-exactness.
#include "common.cuh"

namespace {

template <typename Tin>
__global__ void __launch_bounds__(256) rgb_to_yuv_kernel(const Tin *__restrict__ rgb, int64_t n, double *__restrict__ yuv)
{
    const double Q[4][3] = {{0.21260000, -0.114572, 0.5},                   // utils.py:17-22
                            {0.71520000, -0.385428, -0.454153},
                            {0.07220000, 0.5, -0.045847},
                            {0.0, 0.50196078, 0.50196078}};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double r = (double)rgb[3 * i] / 255.0, g = (double)rgb[3 * i + 1] / 255.0, b = (double)rgb[3 * i + 2] / 255.0;   // :14
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            double v = r * Q[0][k];
            v = fma(g, Q[1][k], v);
            v = fma(b, Q[2][k], v);
            v = v + Q[3][k];
            v = fmin(fmax(v, 0.0), 1.0) * 255.0;                            // :32
            yuv[3 * i + k] = v;
        }
    }
}

}  // namespace

RAHT_API int raht_rgb_to_yuv(const void *rgb, int dtype, int64_t n, double *yuv, void *stream)
{
    RAHT_REQUIRE(n >= 0, "raht_rgb_to_yuv: bad arguments");
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(rgb && yuv, "raht_rgb_to_yuv: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    raht_count_launches(1);
    if (dtype == RAHT_F64) rgb_to_yuv_kernel<double><<<grid_for(n, 256, 8), 256, 0, st>>>((const double *)rgb, n, yuv);
    else if (dtype == RAHT_F32) rgb_to_yuv_kernel<float><<<grid_for(n, 256, 8), 256, 0, st>>>((const float *)rgb, n, yuv);
    else {
        raht_set_error("raht_rgb_to_yuv: bad dtype");
        return RAHT_E_INVALID;
    }
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// ply.cu -- vertex records of the 3DGS PLY files on the device (scene I/O).
// The reference writes a scene vertex by vertex in a Python loop (python/quality_eval.py:18-117: six small
// f.write calls per Gaussian) and reads it back through a numpy structured dtype with one np.stack per property group
// (python/data_util.py:272-382).  Here the file body is ONE buffer: a kernel interleaves the five attribute arrays into
// the vertex records (or splits the records back), the host does a single write / read of that buffer.
//
// Record layout, little-endian fp32 (quality_eval.py:71-110, data_util.py:331-340), cd = colour channels (48):
//   x y z | nx ny nz (zeros) | colours[cd] (f_dc_0..2, f_rest_0..cd-4) | opacity | scale_0..2 | rot_0..3      = 14 + cd floats
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256) ply_pack_kernel(const float *__restrict__ means, const float *__restrict__ quats,
                                                       const float *__restrict__ scales, const float *__restrict__ opac,
                                                       const float *__restrict__ colors, int cd, int64_t n, float *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const int W = 14 + cd;
    for (int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); i < n; i += (int64_t)gridDim.x * 8) {
        float *dst = out + (size_t)i * W;
        for (int c = lane; c < W; c += 32) {
            float v;
            if (c < 3) v = means[(size_t)i * 3 + c];
            else if (c < 6) v = 0.0f;                                           // normals: placeholder (quality_eval.py:96)
            else if (c < 6 + cd) v = colors[(size_t)i * cd + c - 6];
            else if (c == 6 + cd) v = opac[i];
            else if (c < 10 + cd) v = scales[(size_t)i * 3 + c - 7 - cd];
            else v = quats[(size_t)i * 4 + c - 10 - cd];
            dst[c] = v;
        }
    }
}

// records -> V_int (n, 3) int64 (positions truncated like torch's .long(), data_util.py:346) and attributes
// (n, 8 + cd) fp32 in the order [rot 4 | scale 3 | opacity 1 | colours cd] (data_util.py:348-361); means_f32 optional.
__global__ void __launch_bounds__(256) ply_unpack_kernel(const float *__restrict__ in, int cd, int64_t n, int64_t *__restrict__ vint,
                                                         float *__restrict__ means, float *__restrict__ attrs)
{
    const int lane = threadIdx.x & 31;
    const int W = 14 + cd, A = 8 + cd;
    for (int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); i < n; i += (int64_t)gridDim.x * 8) {
        const float *src = in + (size_t)i * W;
        for (int c = lane; c < W; c += 32) {
            const float v = src[c];
            if (c < 3) {
                if (vint) vint[(size_t)i * 3 + c] = (int64_t)v;
                if (means) means[(size_t)i * 3 + c] = v;
            } else if (c < 6) {
            } else if (c < 6 + cd) attrs[(size_t)i * A + 8 + (c - 6)] = v;
            else if (c == 6 + cd) attrs[(size_t)i * A + 7] = v;
            else if (c < 10 + cd) attrs[(size_t)i * A + 4 + (c - 7 - cd)] = v;
            else attrs[(size_t)i * A + (c - 10 - cd)] = v;
        }
    }
}

}  // namespace

RAHT_API int raht_ply_pack_vertices(const float *means, const float *quats, const float *scales, const float *opacities,
                                    const float *colors, int color_dim, int64_t n, float *records, void *stream)
{
    RAHT_REQUIRE(n >= 0 && color_dim >= 0, "raht_ply_pack_vertices: bad arguments");
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(means && quats && scales && opacities && (colors || color_dim == 0) && records, "raht_ply_pack_vertices: null pointer");
    raht_count_launches(1);
    ply_pack_kernel<<<grid_for(n, 8, 8), 256, 0, (cudaStream_t)stream>>>(means, quats, scales, opacities, colors, color_dim, n, records);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_ply_unpack_vertices(const float *records, int color_dim, int64_t n, int64_t *v_int, float *means, float *attributes,
                                      void *stream)
{
    RAHT_REQUIRE(n >= 0 && color_dim >= 0, "raht_ply_unpack_vertices: bad arguments");
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(records && attributes, "raht_ply_unpack_vertices: null pointer");
    raht_count_launches(1);
    ply_unpack_kernel<<<grid_for(n, 8, 8), 256, 0, (cudaStream_t)stream>>>(records, color_dim, n, v_int, means, attributes);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

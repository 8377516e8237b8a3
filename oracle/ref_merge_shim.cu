// ref_merge_shim.cu -- TEST INFRASTRUCTURE ONLY (see oracle/raht_oracle.c header).
//
// A C-ABI launcher around the reference's OWN segment-merge kernel, compiled from the source where it
// lies under /root/reference (cuda/merge_cluster.cu:2-111, a bare __global__ with no includes) by
// oracle/Makefile into oracle/_ref/libmerge_ref_{ieee,default,fast}.so.  No reference source is copied
// into this repo: the kernel file is #included through -I at build time and launched the way the
// reference's torch wrapper launches it (cuda/merge_cluster_wrapper.cu:86-107: 256 threads per block,
// one thread per cluster, outputs pre-zeroed by the caller as torch::zeros does at :64-69).
//
// Three builds of the same source:
//   ieee     --fmad=false            every product and sum rounded separately (what oracle/raht_oracle.c
//                                    restates and what csrc/merge.cu computes): the bit-exact comparison
//   default  nvcc defaults           mul+add contracted to FMA (a build without any arithmetic flag)
//   fast     --use_fast_math         the flags the reference ships (cuda/setup.py:27-30, there for sm_86)
#include "merge_cluster.cu"
#include <cuda_runtime.h>

extern "C" __attribute__((visibility("default")))
int ref_merge_launch(const int *cluster_indices, const int *cluster_offsets, int num_clusters,
                     const float *means, const float *quats, const float *scales, const float *opacities,
                     const float *colors, int color_dim, int weight_by_opacity,
                     float *merged_means, float *merged_quats, float *merged_scales, float *merged_opacities,
                     float *merged_colors, void *stream)
{
    if (num_clusters <= 0) return 0;
    const int threads = 256;                                           // merge_cluster_wrapper.cu:87
    const int blocks = (num_clusters + threads - 1) / threads;         // :88
    merge_weighted_mean_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(
        cluster_indices, cluster_offsets, num_clusters, means, quats, scales, opacities, colors, color_dim,
        weight_by_opacity != 0, merged_means, merged_quats, merged_scales, merged_opacities, merged_colors);
    return (int)cudaGetLastError();
}

// Shared pieces of the fused qkv + attention kernel for 65-token images (qkv_attn_tc.cuh): parameters, the weight
// regrouping done once at pack time, and the optional clock-stamp trace.
#pragma once

#include "attention.cuh"
#include "gemm2.cuh"

namespace ddc {

constexpr int QA_NB = 192;            // q | k | v rows of one head pair
constexpr int QA_L = 65;

struct QkvAttnParams {
    int B, C, n_hg, stages, num_tp;   // images, embed_dim, head pairs, W ring stages, tile pairs (4 images each)
    int n_split;                      // work units per tile pair (tcgen05 version): the head pairs are cut into this many runs
    const float* bias_hp;             // [n_hg][192] bias in the rearranged order
    const bf16* qkv_cls;              // [B, 3C] q | k | v of the class rows (standard column order)
    bf16* out;                        // [B * 65, C]
    float scale_log2e;
    long long* trace;                 // optional [6 roles][32 head-pair iterations][8 events] clock stamps of CTA 0 (tuning aid)
};

__device__ __forceinline__ void qa_stamp(const QkvAttnParams& p, int role, uint32_t it, int ev) {
    if (p.trace != nullptr && blockIdx.x == 0 && it < 32u && (threadIdx.x & 31) == 0) p.trace[(role * 32 + it) * 8 + ev] = clock64();
}

// W_hp[hg*192 + s*64 + j][:] = W_qkv[s*C + hg*64 + j][:]  (16 bit), bias likewise: the q | k | v rows of a head pair as one tile
__global__ void repack_qkv_hp_kernel(const bf16* __restrict__ w, const float* __restrict__ b, bf16* __restrict__ w_hp,
                                     float* __restrict__ b_hp, int C) {
    const int n_rows = 3 * C;
    for (int r = blockIdx.x; r < n_rows; r += gridDim.x) {
        const int hg = r / QA_NB, rem = r % QA_NB, s = rem / 64, j = rem % 64;
        const int src = s * C + hg * 64 + j;
        for (int k = threadIdx.x; k < C / 8; k += blockDim.x)
            reinterpret_cast<uint4*>(w_hp + static_cast<size_t>(r) * C)[k] = reinterpret_cast<const uint4*>(w + static_cast<size_t>(src) * C)[k];
        if (threadIdx.x == 0) b_hp[r] = b[src];
    }
}

}  // namespace ddc

// Short-sequence fused attention (ViT.py:107-114): softmax(q k^T * d^-0.5) v per (image, head).
//
// Token counts here are 65 / 257 / 626 with head_dim 32 or 64, so K and V of one head fit in
// shared memory and the [L, L] score matrix never exists in HBM.  One CTA per (head, image);
// each warp owns 16-query tiles, keeps scores / probabilities in registers (flash-style
// running max and sum in fp32, exp2 with the log2e-folded scale), and feeds the probability
// fragments straight back in as the A operand of P*V.  Tiles this small (K = head_dim = 32,
// M = 65) cannot fill a 128-row tcgen05 tile, so the contraction runs on warp-level
// mma.sync m16n8k16 bf16 with fp32 accumulation.
//
// Layouts: qkv [rows, 3C] bf16 with column s*C + head*D + e (s = q,k,v) exactly as the fused
// qkv Linear emits it (ViT.py:107); out [rows, C] bf16 with column head*D + e (ViT.py:114).
#pragma once

#include "common.cuh"

namespace ddc {

constexpr int ATT_MAX_WARPS = 8;

template <int D>
__host__ __device__ constexpr int att_row_bytes() { return (D + 8) * 2; }   // +16 B pad: conflict-free ldmatrix

inline size_t att_smem_bytes(int L, int D, int KC) {
    const int lk = (L + KC - 1) / KC * KC;
    return static_cast<size_t>(2) * lk * (D + 8) * 2;
}

template <int D, int KC>
__global__ void __launch_bounds__(ATT_MAX_WARPS * 32)
attention_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ out, int L, int C, float scale_log2e) {
    extern __shared__ __align__(16) uint8_t att_smem[];
    constexpr int RS = att_row_bytes<D>();
    constexpr int NT = KC / 8;    // score n-tiles per key chunk
    constexpr int DT = D / 8;     // output n-tiles
    constexpr int DK = D / 16;    // k-steps of q k^T
    const int head = blockIdx.x;
    const int img = blockIdx.y;
    const int lk_pad = (L + KC - 1) / KC * KC;
    uint8_t* Ks = att_smem;
    uint8_t* Vs = att_smem + static_cast<size_t>(lk_pad) * RS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nwarps = blockDim.x >> 5;
    const size_t ld = static_cast<size_t>(3) * C;
    const bf16* base = qkv + static_cast<size_t>(img) * L * ld + head * D;

    // ---- stage K and V of this head (zero the padded key rows: 0 * garbage must stay 0)
    constexpr int CPR = D / 8;    // 16-byte chunks per row
    for (int i = threadIdx.x; i < lk_pad * CPR; i += blockDim.x) {
        const int r = i / CPR, c = i - r * CPR;
        const uint32_t kd = smem_u32(Ks + r * RS + c * 16);
        const uint32_t vd = smem_u32(Vs + r * RS + c * 16);
        if (r < L) {
            cp_async16(kd, base + r * ld + C + c * 8);
            cp_async16(vd, base + r * ld + 2 * C + c * 8);
        } else {
            *reinterpret_cast<uint4*>(Ks + r * RS + c * 16) = make_uint4(0, 0, 0, 0);
            *reinterpret_cast<uint4*>(Vs + r * RS + c * 16) = make_uint4(0, 0, 0, 0);
        }
    }
    cp_async_wait_all();
    __syncthreads();

    const int g = lane >> 2, t = lane & 3;
    const int qtiles = (L + 15) / 16;
    for (int qt = warp; qt < qtiles; qt += nwarps) {
        const int q0 = qt * 16;
        // ---- Q fragments straight from global (each element is used by exactly one warp)
        uint32_t qa[DK][4];
        {
            const int r0 = q0 + g, r1 = q0 + g + 8;
            const uint32_t* p0 = reinterpret_cast<const uint32_t*>(base + static_cast<size_t>(r0) * ld);
            const uint32_t* p1 = reinterpret_cast<const uint32_t*>(base + static_cast<size_t>(r1) * ld);
#pragma unroll
            for (int ks = 0; ks < DK; ++ks) {
                qa[ks][0] = r0 < L ? __ldg(p0 + ks * 8 + t) : 0u;
                qa[ks][1] = r1 < L ? __ldg(p1 + ks * 8 + t) : 0u;
                qa[ks][2] = r0 < L ? __ldg(p0 + ks * 8 + 4 + t) : 0u;
                qa[ks][3] = r1 < L ? __ldg(p1 + ks * 8 + 4 + t) : 0u;
            }
        }
        float o[DT][4];
#pragma unroll
        for (int i = 0; i < DT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
        float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

        for (int kc0 = 0; kc0 < lk_pad; kc0 += KC) {
            float s[NT][4];
#pragma unroll
            for (int i = 0; i < NT; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
            // ---- S = Q K^T
#pragma unroll
            for (int nt = 0; nt < NT; nt += 2) {
#pragma unroll
                for (int ks = 0; ks < DK; ++ks) {
                    uint32_t b0, b1, b2, b3;
                    const int krow = kc0 + nt * 8 + (lane & 7) + ((lane >> 4) << 3);
                    const int kcol = ks * 16 + (((lane >> 3) & 1) << 3);
                    ldmatrix_x4(b0, b1, b2, b3, smem_u32(Ks + krow * RS + kcol * 2));
                    mma_bf16_16816(s[nt], qa[ks], b0, b1);
                    mma_bf16_16816(s[nt + 1], qa[ks], b2, b3);
                }
            }
            // ---- mask padded keys, running max
            float mx0 = m0, mx1 = m1;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const int key = kc0 + nt * 8 + 2 * t;
                s[nt][0] = key < L ? s[nt][0] * scale_log2e : -INFINITY;
                s[nt][1] = key + 1 < L ? s[nt][1] * scale_log2e : -INFINITY;
                s[nt][2] = key < L ? s[nt][2] * scale_log2e : -INFINITY;
                s[nt][3] = key + 1 < L ? s[nt][3] * scale_log2e : -INFINITY;
                mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
                mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
            }
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
            const float c0 = exp2f(m0 - mx0), c1 = exp2f(m1 - mx1);
            m0 = mx0; m1 = mx1;
            l0 *= c0; l1 *= c1;
#pragma unroll
            for (int i = 0; i < DT; ++i) { o[i][0] *= c0; o[i][1] *= c0; o[i][2] *= c1; o[i][3] *= c1; }
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                s[nt][0] = exp2f(s[nt][0] - m0); s[nt][1] = exp2f(s[nt][1] - m0);
                s[nt][2] = exp2f(s[nt][2] - m1); s[nt][3] = exp2f(s[nt][3] - m1);
                l0 += s[nt][0] + s[nt][1];
                l1 += s[nt][2] + s[nt][3];
            }
            // ---- O += P V  (P fragments come directly from the score accumulators)
#pragma unroll
            for (int kk = 0; kk < KC / 16; ++kk) {
                uint32_t pa[4];
                pa[0] = pack_bf16(s[2 * kk][0], s[2 * kk][1]);
                pa[1] = pack_bf16(s[2 * kk][2], s[2 * kk][3]);
                pa[2] = pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
                pa[3] = pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
                for (int dt = 0; dt < DT; dt += 2) {
                    uint32_t b0, b1, b2, b3;
                    const int vrow = kc0 + kk * 16 + (lane & 7) + (((lane >> 3) & 1) << 3);
                    const int vcol = dt * 8 + ((lane >> 4) << 3);
                    ldmatrix_x4_trans(b0, b1, b2, b3, smem_u32(Vs + vrow * RS + vcol * 2));
                    mma_bf16_16816(o[dt], pa, b0, b1);
                    mma_bf16_16816(o[dt + 1], pa, b2, b3);
                }
            }
        }
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
        l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        const float i0 = 1.f / l0, i1 = 1.f / l1;
        const int r0 = q0 + g, r1 = q0 + g + 8;
        bf16* ob = out + static_cast<size_t>(img) * L * C + head * D;
#pragma unroll
        for (int dt = 0; dt < DT; ++dt) {
            if (r0 < L)
                *reinterpret_cast<uint32_t*>(ob + static_cast<size_t>(r0) * C + dt * 8 + 2 * t) =
                    pack_bf16(o[dt][0] * i0, o[dt][1] * i0);
            if (r1 < L)
                *reinterpret_cast<uint32_t*>(ob + static_cast<size_t>(r1) * C + dt * 8 + 2 * t) =
                    pack_bf16(o[dt][2] * i1, o[dt][3] * i1);
        }
    }
}

}  // namespace ddc

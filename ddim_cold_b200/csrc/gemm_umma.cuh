// Token GEMM on the 5th-gen tensor cores:  OUT[M, N] = epilogue( A'[M, K] * W[N, K]^T ).
//
// One CTA owns a 128-row M tile and walks every N tile of it, so the A operand (the whole
// K extent, K <= 512) is staged in shared memory exactly once:
//   * A_TMA : A is a bf16 activation matrix, fetched by TMA (SWIZZLE_128B boxes of 64 x 128).
//   * A_LN  : A is LayerNorm(X) of the fp32 residual stream; seven warps normalise the rows
//             (two-pass fp32 statistics, eps 1e-5, affine) and write bf16 straight into the
//             same swizzled layout, so no normalised copy ever touches HBM (ViT.py:133,137,212).
// W tiles (bn x 64) stream through a TMA ring.  A single thread issues tcgen05.mma
// (M=128, N=bn, K=16) into one of two TMEM accumulator stages; four epilogue warps drain the
// other stage with tcgen05.ld and apply the fused epilogue:
//   EPI_BF16      out = bf16(acc + bias)                       qkv      (ViT.py:107)
//   EPI_BF16_GELU out = bf16(gelu_erf(acc + bias))             fc1      (ViT.py:85-86)
//   EPI_RESID     X  += acc + bias          (fp32, in place)   proj/fc2 (ViT.py:115,88,136-137)
//   EPI_F32       out = acc + bias          (fp32)             head     (ViT.py:213)
//   EPI_EMBED     X   = acc + posb[l] + time[t]                patch embed + cls/pos/time
//                                                              (ViT.py:154,202-205)
#pragma once

#include "common.cuh"

namespace ddc {

enum { A_TMA = 0, A_LN = 1 };
enum { EPI_BF16 = 0, EPI_BF16_GELU = 1, EPI_RESID = 2, EPI_F32 = 3, EPI_EMBED = 4 };

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;                       // one 128-byte swizzle row of bf16
constexpr int GEMM_THREADS = 256;                 // warp0 TMA, warp1 MMA, warp2 TMEM, warp3 spare, warps4-7 epilogue
constexpr int GEMM_A_KBLOCK_BYTES = GEMM_BM * 128;

struct GemmParams {
    int M, N, K, bn, stages;
    // A_LN prologue
    const float* x;
    int ldx;
    const float* ln_g;
    const float* ln_b;
    float ln_eps;
    // epilogue
    const float* bias;
    bf16* out_bf16;
    int ld_bf16;
    float* out_f32;
    int ld_f32;
    // EPI_EMBED
    const float* posb;       // [L, N]: row 0 = cls + pos[0]; row l>0 = pos[l] + conv bias
    int L;
    const float* time_tab;   // [total_steps, N] fp32
    const int* t_img;        // per-image timestep (forward()), or null -> tbl[*cur].t
    const StepEntry* tbl;
    const int* cur;
};

__host__ __device__ inline uint32_t pow2_cols(uint32_t c) {
    uint32_t p = 32;
    while (p < c) p <<= 1;
    return p;
}

inline size_t gemm_smem_bytes(int K, int bn, int stages) {
    return 1024 /*align slack*/ + static_cast<size_t>(K / GEMM_BK) * GEMM_A_KBLOCK_BYTES +
           static_cast<size_t>(stages) * bn * 128 + 256;
}

template <int EPI>
__device__ __forceinline__ void epilogue_chunk16(const GemmParams& p, int row, int n0, float* v, const float* embed_pos,
                                                 const float* embed_time) {
    if (EPI == EPI_EMBED) {
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
            float4 a = __ldg(reinterpret_cast<const float4*>(embed_pos + n0 + i));
            float4 b = __ldg(reinterpret_cast<const float4*>(embed_time + n0 + i));
            v[i] += a.x + b.x; v[i + 1] += a.y + b.y; v[i + 2] += a.z + b.z; v[i + 3] += a.w + b.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
            float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + i));
            v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
        }
    }
    if (EPI == EPI_BF16 || EPI == EPI_BF16_GELU) {
        if (EPI == EPI_BF16_GELU) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = gelu_erf(v[i]);
        }
        uint4 lo, hi;
        lo.x = pack_bf16(v[0], v[1]);   lo.y = pack_bf16(v[2], v[3]);
        lo.z = pack_bf16(v[4], v[5]);   lo.w = pack_bf16(v[6], v[7]);
        hi.x = pack_bf16(v[8], v[9]);   hi.y = pack_bf16(v[10], v[11]);
        hi.z = pack_bf16(v[12], v[13]); hi.w = pack_bf16(v[14], v[15]);
        uint4* dst = reinterpret_cast<uint4*>(p.out_bf16 + static_cast<size_t>(row) * p.ld_bf16 + n0);
        dst[0] = lo;
        dst[1] = hi;
    } else {
        float4* dst = reinterpret_cast<float4*>(p.out_f32 + static_cast<size_t>(row) * p.ld_f32 + n0);
        if (EPI == EPI_RESID) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float4 r = dst[i];
                v[4 * i] += r.x; v[4 * i + 1] += r.y; v[4 * i + 2] += r.z; v[4 * i + 3] += r.w;
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
}

template <int AMODE, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

    const int kblocks = p.K / GEMM_BK;
    const int num_nt = p.N / p.bn;
    const uint32_t b_stage_bytes = static_cast<uint32_t>(p.bn) * 128u;
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + static_cast<size_t>(kblocks) * GEMM_A_KBLOCK_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + static_cast<size_t>(p.stages) * b_stage_bytes);
    uint64_t* full_bar = bars;                     // [stages]  TMA -> MMA
    uint64_t* empty_bar = bars + p.stages;         // [stages]  MMA -> TMA
    uint64_t* tfull_bar = bars + 2 * p.stages;     // [2]       MMA -> epilogue
    uint64_t* tempty_bar = tfull_bar + 2;          // [2]       epilogue -> MMA
    uint64_t* a_bar = tempty_bar + 2;              // A operand resident in smem
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(a_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * GEMM_BM;
    const uint32_t acc_stride = pow2_cols(p.bn);
    const uint32_t tmem_cols = pow2_cols(2 * acc_stride);

    if (warp == 0 && lane == 0) {
        if (AMODE == A_TMA) tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&tfull_bar[0], 1);
        mbar_init(&tfull_bar[1], 1);
        mbar_init(&tempty_bar[0], 4);
        mbar_init(&tempty_bar[1], 4);
        mbar_init(a_bar, AMODE == A_TMA ? 1 : 7);
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc(tmem_slot, tmem_cols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (AMODE == A_LN && warp != 0) {
        // ---- LayerNorm prologue: warps 1..7 normalise the tile's 128 rows into swizzled smem
        const int chunks = p.K >> 3;                       // 16-byte (8 x bf16) chunks per row
        const float inv_k = 1.0f / static_cast<float>(p.K);
        for (int r = warp - 1; r < GEMM_BM; r += 7) {
            const float* xr = p.x + static_cast<size_t>(m0 + r) * p.ldx;
            float v[2][8];
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int ci = lane + 32 * j;
                if (ci < chunks) {
                    const float4 a = *reinterpret_cast<const float4*>(xr + ci * 8);
                    const float4 b = *reinterpret_cast<const float4*>(xr + ci * 8 + 4);
                    v[j][0] = a.x; v[j][1] = a.y; v[j][2] = a.z; v[j][3] = a.w;
                    v[j][4] = b.x; v[j][5] = b.y; v[j][6] = b.z; v[j][7] = b.w;
#pragma unroll
                    for (int i = 0; i < 8; ++i) s += v[j][i];
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            const float mean = s * inv_k;
            float q = 0.f;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (lane + 32 * j < chunks) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float d = v[j][i] - mean;
                        q += d * d;
                    }
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
            const float rstd = rsqrtf(q * inv_k + p.ln_eps);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int ci = lane + 32 * j;
                if (ci < chunks) {
                    const float4 g0 = __ldg(reinterpret_cast<const float4*>(p.ln_g + ci * 8));
                    const float4 g1 = __ldg(reinterpret_cast<const float4*>(p.ln_g + ci * 8 + 4));
                    const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.ln_b + ci * 8));
                    const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.ln_b + ci * 8 + 4));
                    uint4 o;
                    o.x = pack_bf16((v[j][0] - mean) * rstd * g0.x + b0.x, (v[j][1] - mean) * rstd * g0.y + b0.y);
                    o.y = pack_bf16((v[j][2] - mean) * rstd * g0.z + b0.z, (v[j][3] - mean) * rstd * g0.w + b0.w);
                    o.z = pack_bf16((v[j][4] - mean) * rstd * g1.x + b1.x, (v[j][5] - mean) * rstd * g1.y + b1.y);
                    o.w = pack_bf16((v[j][6] - mean) * rstd * g1.z + b1.z, (v[j][7] - mean) * rstd * g1.w + b1.w);
                    const int kb = ci >> 3;
                    const int c = ci & 7;
                    *reinterpret_cast<uint4*>(smem_a + static_cast<size_t>(kb) * GEMM_A_KBLOCK_BYTES + r * 128 +
                                              ((c ^ (r & 7)) << 4)) = o;
                }
            }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(a_bar);
    }

    if (warp == 0) {
        // ---- TMA producer
        if (lane == 0) {
            if (AMODE == A_TMA) {
                mbar_arrive_expect_tx(a_bar, static_cast<uint32_t>(kblocks) * GEMM_A_KBLOCK_BYTES);
                for (int kb = 0; kb < kblocks; ++kb)
                    tma_load_2d(smem_a + static_cast<size_t>(kb) * GEMM_A_KBLOCK_BYTES, &tmA, a_bar, kb * GEMM_BK, m0);
            }
            uint32_t stage = 0, phase = 0;
            for (int nt = 0; nt < num_nt; ++nt) {
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    mbar_arrive_expect_tx(&full_bar[stage], b_stage_bytes);
                    tma_load_2d(smem_b + static_cast<size_t>(stage) * b_stage_bytes, &tmB, &full_bar[stage], kb * GEMM_BK,
                                nt * p.bn);
                    if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ---- MMA issuer (lane 0 issues and commits)
        const uint32_t idesc = umma_idesc_bf16(GEMM_BM, static_cast<uint32_t>(p.bn));
        const uint32_t a_base = smem_u32(smem_a);
        const uint32_t b_base = smem_u32(smem_b);
        mbar_wait(a_bar, 0);
        tc_fence_after();
        uint32_t stage = 0, phase = 0;
        for (int nt = 0; nt < num_nt; ++nt) {
            const uint32_t acc = nt & 1;
            mbar_wait(&tempty_bar[acc], ((nt >> 1) & 1) ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + acc * acc_stride;
            for (int kb = 0; kb < kblocks; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                if (lane == 0) {
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; ++k) {
                        const uint64_t ad = umma_desc_sw128(a_base + kb * GEMM_A_KBLOCK_BYTES + k * 32);
                        const uint64_t bd = umma_desc_sw128(b_base + stage * b_stage_bytes + k * 32);
                        umma_bf16_ss(d_tmem, ad, bd, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty_bar[stage]);
                    if (kb == kblocks - 1) umma_commit(&tfull_bar[acc]);
                }
                __syncwarp();
                if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ---- epilogue: warp w drains TMEM lanes [32*(w-4), +32); thread = one output row
        const int q = warp - 4;
        const int row = m0 + q * 32 + lane;
        const bool row_ok = row < p.M;
        const float* embed_pos = nullptr;
        const float* embed_time = nullptr;
        if (EPI == EPI_EMBED && row_ok) {
            const int b = row / p.L;
            const int l = row - b * p.L;
            const int t = p.t_img ? p.t_img[b] : p.tbl[*p.cur].t;
            embed_pos = p.posb + static_cast<size_t>(l) * p.N;
            embed_time = p.time_tab + static_cast<size_t>(t) * p.N;
        }
        for (int nt = 0; nt < num_nt; ++nt) {
            const uint32_t acc = nt & 1;
            mbar_wait(&tfull_bar[acc], (nt >> 1) & 1);
            tc_fence_after();
            const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * acc_stride;
            for (int c = 0; c < p.bn; c += 16) {
                float v[16];
                tmem_ld16(t_row + c, v);           // warp-collective: every lane takes part
                if (row_ok) epilogue_chunk16<EPI>(p, row, nt * p.bn + c, v, embed_pos, embed_time);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, tmem_cols);
}

}  // namespace ddc

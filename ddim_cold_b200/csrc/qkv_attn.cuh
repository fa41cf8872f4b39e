// LayerNorm-fed qkv GEMM and attention of a transformer block in ONE kernel (ViT.py:105-114), for the 65-token shapes
// (64 patch tokens + the class token, head_dim 32: S1).  Separate kernels write Q/K/V to HBM (153 MB per block at 1024
// images) and read them straight back; here they never leave the SM:
//
//   per CTA pair (cta_group::2, M = 256): the 64 patch rows of FOUR images (two per CTA) as the A tile, resident for the tile;
//   per 64 qkv columns (a pair of heads) "hg":
//     tensor core   [256 x 192] = XN_tile * W_hg^T      (W rearranged at pack time so that the q | k | v rows of a head pair
//                                                        are one 192-row tile; each CTA stages 96 of them)
//     epilogue      TMEM -> + bias -> 16 bit -> the head pair's Q / K / V tiles of the CTA's two images in shared memory
//                   (the class token's q / k / v row comes from a small GEMM over the class rows that ran just before)
//     attention     one warp per (image, 16 patch query rows, head): mma.sync flash-style softmax(q k^T) v from shared
//                   memory (attend_qtile of attention.cuh), output staged over the warp's own Q columns and written as
//                   64-byte row segments; the class token's query row (one per image and head: 65 dot products and a
//                   65-term sum) is done on the CUDA cores by the epilogue warps while the attention warps work
//   the accumulator is double-buffered in TMEM, so the tensor core computes head pair hg + 1 while the warps work on hg.
//
// Why the class token is separate: 65 rows per image do not tile a 128-row MMA; 64 do.  Its q / k / v (one row per image)
// are a [B x 3C] GEMM of their own (same kernel as the plain qkv GEMM, with a strided A map), read here from L2.
#pragma once

#include "attention.cuh"
#include "gemm2.cuh"

namespace ddc {

constexpr int QA_THREADS = 768;       // warps 0,3 TMA, 1 MMA, 2 TMEM | 4-7 epilogue + class rows | 8-23 attention
constexpr int QA_NB = 192;            // q | k | v rows of one head pair
constexpr int QA_REG = 80 * 128;      // one image's Q (or K, or V) tile: 80 rows x 128 B (64 patches, class row, padding)
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

inline size_t qkv_attn_smem_bytes(int C, int stages) {
    return 1024 + static_cast<size_t>(C / 64) * G2_A_KB_BYTES + static_cast<size_t>(stages) * 96 * 128 + 3 * 2 * QA_REG +
           static_cast<size_t>(C / 64) * QA_NB * 4 + 1024;
}

template <bool F16>
__global__ void __launch_bounds__(QA_THREADS, 1)
qkv_attn_kernel(const __grid_constant__ CUtensorMap tmA,      // XN [B*65, C] 16 bit, box 64 cols x 64 rows
                const __grid_constant__ CUtensorMap tmW,      // W_hp [n_hg*192, C] 16 bit, box 64 cols x 96 rows
                const QkvAttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
    const int kblocks = p.C / 64;
    constexpr uint32_t kWStage = 96 * 128;
    uint8_t* smem_a = smem;
    uint8_t* smem_w = smem_a + static_cast<size_t>(kblocks) * G2_A_KB_BYTES;
    uint8_t* smem_q = smem_w + static_cast<size_t>(p.stages) * kWStage;      // [image 0/1][80 x 128 B]
    uint8_t* smem_k = smem_q + 2 * QA_REG;
    uint8_t* smem_v = smem_k + 2 * QA_REG;
    float* smem_bias = reinterpret_cast<float*>(smem_v + 2 * QA_REG);         // [n_hg][192]
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(smem_bias) + static_cast<size_t>(p.n_hg) * QA_NB * 4);
    uint64_t* full_bar = bars;            // [8] W ring: TMA -> MMA (leader's copy collects both CTAs' bytes)
    uint64_t* empty_bar = bars + 8;       // [8] W ring: MMA -> TMA (pair commit)
    uint64_t* afull_bar = bars + 16;      // [8] A K block landed (leader)
    uint64_t* aempty_bar = bars + 24;     // [8] A K block no longer needed by this tile (pair commit)
    uint64_t* tfull_bar = bars + 32;      // [2] accumulator stage complete (pair commit)
    uint64_t* tempty_bar = bars + 34;     // [2] accumulator stage drained (leader: 4 epilogue warps x 2 CTAs)
    uint64_t* ready_bar = bars + 36;      // [1] Q/K/V tiles of a head pair are in shared memory (4 epilogue warps)
    uint64_t* free_bar = bars + 37;       // [1] everyone is done with them (16 attention warps + 4 class-query warps)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 40);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    const bool leader = crank == 0;
    const int first_tp = static_cast<int>(blockIdx.x) / 2, tp_step = static_cast<int>(gridDim.x) / 2;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmW);
        for (int i = 0; i < 8; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
            mbar_init(&afull_bar[i], 1);
            mbar_init(&aempty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull_bar[i], 1);
            mbar_init(&tempty_bar[i], 8);
        }
        mbar_init(ready_bar, 4);
        mbar_init(free_bar, 20);
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc<2>(tmem_slot, 512);
        tmem_relinquish<2>();
    }
    if (warp >= 4) {
        for (int i = threadIdx.x - 128; i < p.n_hg * QA_NB; i += QA_THREADS - 128) smem_bias[i] = p.bias_hp[i];
        // rows 64..79 of every Q / K / V region: row 64 is rewritten per head pair, 65..79 stay zero (masked keys must be
        // finite: their probability is 0, and 0 * NaN would poison P*V)
        for (int i = threadIdx.x - 128; i < 6 * 16 * 8; i += QA_THREADS - 128) {
            const int reg = i / 128, rem = i % 128;
            *reinterpret_cast<uint4*>(smem_q + reg * QA_REG + (64 + rem / 8) * 128 + (rem % 8) * 16) = make_uint4(0, 0, 0, 0);
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    // image 0 / 1 of this CTA in tile pair tp
    auto image_of = [&](int tp, int j) { return tp * 4 + static_cast<int>(crank) * 2 + j; };

    if (warp == 0 || warp == 3) {
        // ================================================================= TMA producers (alternate loads)
        if (lane == 0) {
            const uint32_t me = warp == 0 ? 0u : 1u;
            const uint64_t pol_a = l2_policy_evict_first(), pol_w = l2_policy_evict_last();
            uint32_t bit = 0, tc = 0;
            for (int tp = first_tp; tp < p.num_tp; tp += tp_step, ++tc) {
                for (int kb = static_cast<int>(me); kb < kblocks; kb += 2) {
                    mbar_wait(&aempty_bar[kb], (tc & 1) ^ 1);
                    if (leader) mbar_arrive_expect_tx(&afull_bar[kb], 2 * G2_A_KB_BYTES);
                    uint8_t* dst = smem_a + static_cast<size_t>(kb) * G2_A_KB_BYTES;
                    // the 64 patch rows of the CTA's two images (token 0 of an image is its class row)
                    tma_load_2d_pair_hint(dst, &tmA, &afull_bar[kb], kb * 64, image_of(tp, 0) * QA_L + 1, pol_a);
                    tma_load_2d_pair_hint(dst + 64 * 128, &tmA, &afull_bar[kb], kb * 64, image_of(tp, 1) * QA_L + 1, pol_a);
                }
                for (int hg = 0; hg < p.n_hg; ++hg) {
                    for (int kb = 0; kb < kblocks; ++kb, ++bit) {
                        if ((bit & 1u) != me) continue;
                        const uint32_t stage = bit % static_cast<uint32_t>(p.stages);
                        const uint32_t phase = (bit / static_cast<uint32_t>(p.stages)) & 1u;
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * kWStage);
                        tma_load_2d_pair_hint(smem_w + static_cast<size_t>(stage) * kWStage, &tmW, &full_bar[stage], kb * 64,
                                              hg * QA_NB + static_cast<int>(crank) * 96, pol_w);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1 && leader) {
        // ================================================================= MMA issuer (pair leader)
        const uint32_t idesc = umma_idesc_16bit(256, QA_NB, F16);
        const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint64_t ad0 = umma_desc_sw128(smem_u32(smem_a));
        const uint64_t bd0 = umma_desc_sw128(smem_u32(smem_w));
        uint32_t stage = 0, phase = 0, tc = 0, acc_it = 0;
        for (int tp = first_tp; tp < p.num_tp; tp += tp_step, ++tc) {
            for (int hg = 0; hg < p.n_hg; ++hg, ++acc_it) {
                const uint32_t acc = acc_it & 1u;
                mbar_wait(&tempty_bar[acc], ((acc_it >> 1) & 1u) ^ 1u);
                tc_fence_after();
                qa_stamp(p, 0, acc_it, 0);
                for (int kb = 0; kb < kblocks; ++kb) {
                    if (hg == 0) mbar_wait(&afull_bar[kb], tc & 1);
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    if (kb == 0) qa_stamp(p, 0, acc_it, 1);
                    if (elect_one()) {
                        umma_kblock<2>(tb + acc * 256u, ad0 + static_cast<uint64_t>(kb * (G2_A_KB_BYTES >> 4)),
                                       bd0 + static_cast<uint64_t>(stage * (kWStage >> 4)), idesc, kb != 0 ? 1u : 0u);
                        umma_commit_pair(&empty_bar[stage]);
                        if (hg == p.n_hg - 1) umma_commit_pair(&aempty_bar[kb]);
                        if (kb == kblocks - 1) umma_commit_pair(&tfull_bar[acc]);
                    }
                    __syncwarp();
                    if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
                }
                qa_stamp(p, 0, acc_it, 2);
            }
        }
    } else if (warp >= 4 && warp < 8) {
        // ================================================================= epilogue: accumulator -> Q / K / V tiles
        const int q = warp & 3;
        const int trow = q * 32 + lane;                   // row of the CTA's 128-row tile: image trow / 64, patch trow % 64
        const int reg = trow >> 6, lrow = trow & 63;
        const uint32_t lane_taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        uint8_t* dst[3] = {smem_q + reg * QA_REG + lrow * 128, smem_k + reg * QA_REG + lrow * 128, smem_v + reg * QA_REG + lrow * 128};
        const int cj = q >> 1, chd = q & 1;               // this warp's class query: image cj of the CTA, head chd of the pair
        const int vch = chd * 4 + (lane >> 3), voff = (lane & 7) * 2;
        uint32_t acc_it = 0, hg_it = 0;
        for (int tp = first_tp; tp < p.num_tp; tp += tp_step) {
            const int cimg = image_of(tp, cj);
            for (int hg = 0; hg < p.n_hg; ++hg, ++acc_it, ++hg_it) {
                const uint32_t acc = acc_it & 1u;
                mbar_wait(&tfull_bar[acc], (acc_it >> 1) & 1u);
                tc_fence_after();
                if (q == 0) qa_stamp(p, 1, hg_it, 0);
                if (hg_it > 0) mbar_wait(free_bar, (hg_it - 1) & 1u);       // everyone has left the previous head pair's tiles
                if (q == 0) qa_stamp(p, 1, hg_it, 1);
                const float* bias = smem_bias + hg * QA_NB;
#pragma unroll
                for (int c = 0; c < 6; ++c) {                               // 32 columns at a time: q 0-1, k 2-3, v 4-5
                    float v[32];
                    tmem_ld32(lane_taddr + acc * 256u + 32 * c, v);
                    uint8_t* row = dst[c >> 1];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float4 b0 = *reinterpret_cast<const float4*>(bias + 32 * c + 8 * j);
                        const float4 b1 = *reinterpret_cast<const float4*>(bias + 32 * c + 8 * j + 4);
                        uint4 u;
                        u.x = pack2<F16>(v[8 * j] + b0.x, v[8 * j + 1] + b0.y);
                        u.y = pack2<F16>(v[8 * j + 2] + b0.z, v[8 * j + 3] + b0.w);
                        u.z = pack2<F16>(v[8 * j + 4] + b1.x, v[8 * j + 5] + b1.y);
                        u.w = pack2<F16>(v[8 * j + 6] + b1.z, v[8 * j + 7] + b1.w);
                        *reinterpret_cast<uint4*>(row + ((((c & 1) * 4 + j) ^ (lrow & 7)) << 4)) = u;
                    }
                }
                // class token's k | v of image q (warps 0, 1) from the class-row GEMM -> key row 64 of that image's tiles
                if (q < 2 && lane < 16 && image_of(tp, q) < p.B) {
                    const int sk = lane >> 3, ch = lane & 7;
                    const uint4 val = __ldg(reinterpret_cast<const uint4*>(p.qkv_cls + static_cast<size_t>(image_of(tp, q)) * 3 * p.C +
                                                                           (1 + sk) * p.C + hg * 64) + ch);
                    *reinterpret_cast<uint4*>((sk == 0 ? smem_k : smem_v) + q * QA_REG + 64 * 128 + (ch << 4)) = val;   // (64 & 7) = 0
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive_leader(&tempty_bar[acc]);                   // the accumulator stage may be overwritten
                    mbar_arrive(ready_bar);
                }
                if (q == 0) qa_stamp(p, 1, hg_it, 2);
                // ---- class token's query row of (cimg, head chd) on the CUDA cores: lane = key for the scores, lane = dim for P*V
                uint4 qv[4];                                                // the class token's q row for (cimg, head): 32 values
                if (cimg < p.B) {
                    const uint4* qg = reinterpret_cast<const uint4*>(p.qkv_cls + static_cast<size_t>(cimg) * 3 * p.C + hg * 64 + chd * 32);
#pragma unroll
                    for (int c = 0; c < 4; ++c) qv[c] = __ldg(qg + c);
                }
                mbar_wait(ready_bar, hg_it & 1u);
                if (q == 0) qa_stamp(p, 1, hg_it, 3);
                if (cimg < p.B) {
                    const uint8_t* Kt = smem_k + cj * QA_REG;
                    const uint8_t* Vt = smem_v + cj * QA_REG;
                    float sc[3];
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        const int key = lane + 32 * i;
                        float a = -INFINITY;
                        if (key < QA_L) {
                            a = 0.f;
                            const uint8_t* row = Kt + key * 128;
#pragma unroll
                            for (int c = 0; c < 4; ++c) {
                                const uint4 kv = *reinterpret_cast<const uint4*>(row + (((chd * 4 + c) ^ (key & 7)) << 4));
                                const uint32_t kw[4] = {kv.x, kv.y, kv.z, kv.w};
                                const uint32_t qw[4] = {qv[c].x, qv[c].y, qv[c].z, qv[c].w};
#pragma unroll
                                for (int w = 0; w < 4; ++w) {
                                    const float2 a2 = unpack2<F16>(qw[w]), b2 = unpack2<F16>(kw[w]);
                                    a = fmaf(a2.x, b2.x, a);
                                    a = fmaf(a2.y, b2.y, a);
                                }
                            }
                            a *= p.scale_log2e;
                        }
                        sc[i] = a;
                    }
                    float m = fmaxf(sc[0], fmaxf(sc[1], sc[2]));
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                    float e[3], sum = 0.f;
#pragma unroll
                    for (int i = 0; i < 3; ++i) { e[i] = ex2_approx(sc[i] - m); sum += e[i]; }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                    float ov = 0.f;
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
#pragma unroll
                        for (int src = 0; src < 32; ++src) {
                            const int key = src + 32 * i;
                            if (key < QA_L) {
                                const float pk = __shfl_sync(0xffffffffu, e[i], src);
                                const uint16_t hv = *reinterpret_cast<const uint16_t*>(Vt + key * 128 + ((vch ^ (key & 7)) << 4) + voff);
                                ov = fmaf(pk, unpack1<F16>(hv), ov);
                            }
                        }
                    }
                    ov *= __frcp_rn(sum);
                    reinterpret_cast<uint16_t*>(p.out)[static_cast<size_t>(cimg) * QA_L * p.C + hg * 64 + chd * 32 + lane] =
                        static_cast<uint16_t>(pack2<F16>(ov, 0.f) & 0xffffu);
                }
                __syncwarp();
                if (q == 0) qa_stamp(p, 1, hg_it, 4);
                if (lane == 0) mbar_arrive(free_bar);
            }
        }
    } else if (warp >= 8) {
        // ================================================================= attention: task = (image j, 16 patch rows, head)
        const int aw = warp - 8;
        const int j = aw >> 3, q0 = ((aw >> 1) & 3) * 16, cbase = (aw & 1) * 4;
        const int g = lane >> 2, t = lane & 3;
        uint8_t* Qt = smem_q + j * QA_REG;
        const uint8_t* Kt = smem_k + j * QA_REG;
        const uint8_t* Vt = smem_v + j * QA_REG;
        uint32_t hg_it = 0;
        for (int tp = first_tp; tp < p.num_tp; tp += tp_step) {
            const int img = image_of(tp, j);
            for (int hg = 0; hg < p.n_hg; ++hg, ++hg_it) {
                mbar_wait(ready_bar, hg_it & 1u);
                if (aw == 0) qa_stamp(p, 2, hg_it, 0);
                if (img < p.B) {
                    uint32_t qa[2][4];
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks) {
                        const int qrow = q0 + (lane & 7) + (((lane >> 3) & 1) << 3);
                        ldmatrix_x4(qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], att_addr<true>(Qt, qrow, cbase + ks * 2 + (lane >> 4), 0));
                    }
                    float o[4][4];
                    attend_qtile<32, 80, F16, true, 128, true>(qa, Kt, Vt, QA_L, 80, p.scale_log2e, lane, o, cbase);
                    if (aw == 0) qa_stamp(p, 2, hg_it, 1);
                    __syncwarp();                                 // all lanes have their Q fragments: reuse those columns as stage
#pragma unroll
                    for (int dt = 0; dt < 4; ++dt) {
                        const int r0 = q0 + g, r1 = q0 + g + 8;
                        *reinterpret_cast<uint32_t*>(Qt + r0 * 128 + (((cbase + dt) ^ (r0 & 7)) << 4) + 4 * t) = pack2<F16>(o[dt][0], o[dt][1]);
                        *reinterpret_cast<uint32_t*>(Qt + r1 * 128 + (((cbase + dt) ^ (r1 & 7)) << 4) + 4 * t) = pack2<F16>(o[dt][2], o[dt][3]);
                    }
                    __syncwarp();
                    // this warp's 16 rows x 64 B -> global; local row r is token 1 + r
                    bf16* ob = p.out + static_cast<size_t>(img) * QA_L * p.C + hg * 64 + (aw & 1) * 32;
#pragma unroll
                    for (int i = lane; i < 16 * 4; i += 32) {
                        const int r = q0 + (i >> 2), ch = i & 3;
                        *reinterpret_cast<uint4*>(ob + static_cast<size_t>(r + 1) * p.C + ch * 8) =
                            *reinterpret_cast<const uint4*>(Qt + r * 128 + (((cbase + ch) ^ (r & 7)) << 4));
                    }
                }
                __syncwarp();
                if (aw == 0) qa_stamp(p, 2, hg_it, 2);
                if (lane == 0) mbar_arrive(free_bar);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 2) tmem_dealloc<2>(tmem_base, 512);
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

// LayerNorm-fed qkv GEMM and attention of a transformer block in ONE kernel, every contraction on tcgen05 (ViT.py:105-114),
// for the 65-token shapes (64 patch tokens + the class token, heads of 32: S1).  q / k / v never reach HBM.
//
//   per CTA pair: the 64 patch rows of FOUR images (two per CTA) are the resident A tile of a cta_group::2 GEMM (M = 256);
//   per 64 qkv columns (a pair of heads) "hg":
//     pair MMA      acc[256 x 192] = XN_tile * W_hg^T            (W regrouped at pack time: q | k | v rows of a head pair)
//     epilogue      TMEM -> + bias -> 16 bit -> this CTA's Q, K, V tiles [128 rows x 64 dims] in shared memory, in the
//                   128-byte-swizzled K-major layout that both the K operand of Q K^T and the MN-major V operand of P V use
//     per head h    S = Q_h K_h^T      one cta_group::1 MMA, M = 128 (both images' queries) x N = 144: both images' 64 patch
//                                      keys, then the two class keys (rows 128, 129 of the K tile, copied there from the
//                                      class-row GEMM's output).  A row uses the 64 columns of its own image + 1 class column
//                   softmax            one thread per query row: 65 scores from TMEM, exponentials, P (16 bit) to shared
//                                      memory (head 0's P overwrites the Q tile, which no MMA needs once both S are done)
//                   O = P V_h          per image j one MMA group: D[:, 32 j : 32 j + 32] = P[128 x 64] V_j[64 keys x 32],
//                                      written over the S columns (a row's valid result is in its own image's 32 columns)
//                   output             O from TMEM, + p_class * v_class, / row sum, 16 bit, straight to global memory
//     class query   one warp per (image, head): 65 dot products, softmax, 65-term weighted sum on the CUDA cores
//   The class token's q | k | v row of every image comes from a small GEMM over the class rows that ran just before
//   (the 65th row does not tile a 128-row MMA; 64 do).
//
// Both kinds of MMA live in one kernel: the pair GEMM is cta_group::2 (each CTA stages half of every W tile), the
// attention MMAs are cta_group::1 inside each CTA (checked on hardware by scripts/probes/mixed_cg_probe.cu: exact results
// with both kinds in flight into tensor memory allocated pair-wide).
//
// Work unit = (4-image tile, half of the head pairs): twice as many, half as long units as whole tiles, so the last
// round of the persistent grid wastes half as much (256 tiles on 74 pairs: 3.46 rounds of work cost 3.5, not 4).
//
// Tensor memory (512 columns): [0,192) qkv accumulator | [192,336) S / O head 0 | [336,480) S / O head 1.
// Five roles run five different instruction streams on one SM; the loop bodies are kept small (rolled loops where the
// trip count allows) so that together they fit the 32 KB instruction cache -- unrolled, the epilogue ran 1.7x slower.
#pragma once

#include "qkv_attn.cuh"

namespace ddc {

constexpr int QT_THREADS = 640;       // warps 0,3 TMA | 1 pair MMA | 2 TMEM + attention MMA | 4-7 epilogue | 8-11, 12-15 softmax | 16-19 class rows
constexpr int QT_TILE = 128 * 128;    // one Q / K / V / P tile: 128 rows x 128 B
constexpr int QT_KTILE = 144 * 128;   // K tile: 128 patch keys + 16 rows for the two class keys
constexpr uint32_t QT_S = 192, QT_SW = 144;

inline size_t qkv_attn_tc_smem_bytes(int C, int stages) {
    return 1024 + static_cast<size_t>(C / 64) * G2_A_KB_BYTES + static_cast<size_t>(stages) * 96 * 128 + 3 * QT_TILE + QT_KTILE +
           static_cast<size_t>(C / 64) * QA_NB * 4 + 1024 + 1536 + 1024 + 8192;
}

__device__ __forceinline__ void tmem_ld1_nowait(uint32_t taddr, float& v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=f"(v) : "r"(taddr) : "memory");
}

// O[:, 0:32] (fresh) = P[128 x 64 keys] V[64 keys x 32 dims]: four K = 16 steps; P K-major (+32 B per step), V MN-major
// (16-key slices 2048 B apart)
__device__ __forceinline__ void umma_pv64(uint32_t tmem_o, uint64_t pdesc, uint64_t vdesc, uint32_t idesc) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t.reg .b64 a1, b1;\n\t"
        "setp.ne.b32 p, 0, 0;\n\t"
        "setp.eq.b32 q, 0, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "add.u64 a1, %1, 2;\n\tadd.u64 b1, %2, 128;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, q;\n\t"
        "add.u64 a1, %1, 4;\n\tadd.u64 b1, %2, 256;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, q;\n\t"
        "add.u64 a1, %1, 6;\n\tadd.u64 b1, %2, 384;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, q;\n\t}"
        ::"r"(tmem_o), "l"(pdesc), "l"(vdesc), "r"(idesc)
        : "memory");
}

template <bool F16, bool TRACE>
__global__ void __launch_bounds__(QT_THREADS, 1)
qkv_attn_tc_kernel(const __grid_constant__ CUtensorMap tmA,      // XN [B*65, C] 16 bit, box 64 cols x 64 rows
                   const __grid_constant__ CUtensorMap tmW,      // W_hp [n_hg*192, C] 16 bit, box 64 cols x 96 rows
                   const QkvAttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
    const int kblocks = p.C / 64;
    constexpr uint32_t kWStage = 96 * 128;
    uint8_t* smem_a = smem;
    uint8_t* smem_q = smem_a + static_cast<size_t>(kblocks) * G2_A_KB_BYTES;   // tiles first: they need 1024-byte alignment
    uint8_t* smem_k = smem_q + QT_TILE;                                        // Q tile doubles as head 0's P tile
    uint8_t* smem_v = smem_k + QT_KTILE;
    uint8_t* smem_p1 = smem_v + QT_TILE;                                       // head 1's P tile
    uint8_t* smem_w = smem_p1 + QT_TILE;
    float* smem_bias = reinterpret_cast<float*>(smem_w + static_cast<size_t>(p.stages) * kWStage);   // [n_hg][192]
    float* vcls = smem_bias + static_cast<size_t>(p.n_hg) * QA_NB;             // [parity][image][64] class value of the head pair
    float* cls_scratch = vcls + 256;                                           // [class warp][32 q | 64 p]
    uint64_t* bars = reinterpret_cast<uint64_t*>(cls_scratch + 4 * 96);
    uint64_t* full_bar = bars;            // [8] W ring: TMA -> MMA (leader's copy collects both CTAs' bytes)
    uint64_t* empty_bar = bars + 8;       // [8] W ring: MMA -> TMA (pair commit)
    uint64_t* afull_bar = bars + 16;      // [8] A K block landed (leader)
    uint64_t* aempty_bar = bars + 24;     // [8] A K block no longer needed by this unit (pair commit)
    uint64_t* tfull_bar = bars + 32;      // qkv accumulator complete (pair commit)
    uint64_t* tempty_bar = bars + 33;     // qkv accumulator drained (leader: 8 draining warps x 2 CTAs)
    uint64_t* qkv_free = bars + 35;       // Q / K / V tiles no longer needed (one commit per head after its P V + 4 class-row warps)
    uint64_t* cls_ready = bars + 36;      // class values in vcls (4 class-row warps), for the output pass
    uint64_t* s_full = bars + 37;         // [2] S of head h complete (commit)
    uint64_t* o_full = bars + 41;         // [2] O of head h complete (commit)
    uint64_t* o_free = bars + 43;         // [2] O of head h read out: its columns may take the next S (4 softmax warps)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 46);
    uint8_t* out_stage = reinterpret_cast<uint8_t*>(bars + 128);      // [8 softmax warps][16 rows x 64 B]: output rows on their way to global memory

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    const bool leader = crank == 0;
    // work units: u -> tile pair u / n_split, head pairs [(u % n_split) * hg_per, + hg_per)
    const int first_u = static_cast<int>(blockIdx.x) / 2, u_step = static_cast<int>(gridDim.x) / 2;
    const int n_units = p.num_tp * p.n_split, hg_per = p.n_hg / p.n_split;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmW);
        for (int i = 0; i < 8; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
            mbar_init(&afull_bar[i], 1);
            mbar_init(&aempty_bar[i], 1);
        }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, 16);
        mbar_init(qkv_free, 6);
        mbar_init(cls_ready, 4);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s_full[i], 1);
            mbar_init(&o_full[i], 1);
            mbar_init(&o_free[i], 4);
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc<2>(tmem_slot, 512);
        tmem_relinquish<2>();
    }
    if (warp >= 4)
        for (int i = threadIdx.x - 128; i < p.n_hg * QA_NB; i += QT_THREADS - 128) smem_bias[i] = p.bias_hp[i];
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    // image 0 / 1 of this CTA in tile pair tp
    auto image_of = [&](int tp, int j) { return tp * 4 + static_cast<int>(crank) * 2 + j; };

    // Accumulator -> Q / K / V tiles: three 32-column chunks starting at chunk c0 (q = chunks 0-1, k = 2-3, v = 4-5) of
    // the calling warp's TMEM lane quarter.  The epilogue warps take chunks 0-2, the class-row warps (idle at that
    // time, same lane quarters) chunks 3-5.  One rolled loop: straight-line code runs at instruction-fetch speed here.
    auto drain3 = [&](int c0, int hg, uint32_t trace_it) {
        const int q = warp & 3, r = q * 32 + lane;
        const uint32_t lane_taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        const float* bias = smem_bias + hg * QA_NB;
#pragma unroll 1
        for (int c = c0; c < c0 + 3; ++c) {
            // the chunk's bias first: the TMEM read below is an asm statement with a memory clobber, nothing is scheduled
            // across it, and shared-memory loads issued after it would add their latency to every chunk
            float4 bb[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) bb[j] = *reinterpret_cast<const float4*>(bias + 32 * c + 4 * j);
            float v[32];
            tmem_ld32(lane_taddr + 32 * c, v);
            uint8_t* row = (c < 2 ? smem_q : (c < 4 ? smem_k : smem_v)) + r * 128;
            const int half = (c & 1) * 4;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 b0 = bb[2 * j], b1 = bb[2 * j + 1];
                uint4 w;
                w.x = pack2<F16>(v[8 * j] + b0.x, v[8 * j + 1] + b0.y);
                w.y = pack2<F16>(v[8 * j + 2] + b0.z, v[8 * j + 3] + b0.w);
                w.z = pack2<F16>(v[8 * j + 4] + b1.x, v[8 * j + 5] + b1.y);
                w.w = pack2<F16>(v[8 * j + 6] + b1.z, v[8 * j + 7] + b1.w);
                *reinterpret_cast<uint4*>(row + (((half + j) ^ (r & 7)) << 4)) = w;
            }
            if constexpr (TRACE) { if (q == 0) qa_stamp(p, c0 == 0 ? 1 : 4, trace_it, 3 + c - c0); }
        }
    };
    auto drain_finish = [&]() {
        fence_proxy_async_smem();                                   // the tiles (and the class keys) are MMA operands
        tc_fence_before();
        named_bar_sync(1, 256);                                     // all eight draining warps of this CTA are done
        if (lane == 0) mbar_arrive_leader(tempty_bar);              // the accumulator may be overwritten
    };
    // S_h = Q_h K_h^T for both heads of the pair, issued by one of the draining warps right behind their barrier (a
    // separate issuer warp behind an mbarrier cost ~0.9k cycles per head pair in wake-up and instruction fetch)
    auto issue_s = [&](uint32_t it) {
        tc_fence_after();
        if (it > 0) {                                               // S of head h goes where its previous O was
            mbar_wait_parked(&o_free[0], (it - 1) & 1u);
            mbar_wait_parked(&o_free[1], (it - 1) & 1u);
            tc_fence_after();
        }
        if (elect_one()) {
            const uint32_t idesc_s = umma_idesc_16bit(128, QT_SW, F16);
            const uint64_t qd = umma_desc_sw128(smem_u32(smem_q));
            const uint64_t kd = umma_desc_sw128(smem_u32(smem_k));
            umma_k32(tmem_base + QT_S, qd, kd, idesc_s);                    // head 0: bytes [0, 64) of every row
            umma_commit(&s_full[0]);
            umma_k32(tmem_base + QT_S + QT_SW, qd + 4, kd + 4, idesc_s);    // head 1: bytes [64, 128)
            umma_commit(&s_full[1]);
        }
        __syncwarp();
    };

    if (warp == 0 || warp == 3) {
        // ================================================================= TMA producers (alternate loads)
        if (lane == 0) {
            const uint32_t me = warp == 0 ? 0u : 1u;
            const uint64_t pol_a = l2_policy_evict_first(), pol_w = l2_policy_evict_last();
            uint32_t bit = 0, uc = 0;
            for (int u = first_u; u < n_units; u += u_step, ++uc) {
                const int tp = u / p.n_split, hg0 = (u % p.n_split) * hg_per;
                for (int kb = static_cast<int>(me); kb < kblocks; kb += 2) {
                    mbar_wait_parked(&aempty_bar[kb], (uc & 1) ^ 1);
                    if (leader) mbar_arrive_expect_tx(&afull_bar[kb], 2 * G2_A_KB_BYTES);
                    uint8_t* dst = smem_a + static_cast<size_t>(kb) * G2_A_KB_BYTES;
                    // the 64 patch rows of the CTA's two images (token 0 of an image is its class row)
                    tma_load_2d_pair_hint(dst, &tmA, &afull_bar[kb], kb * 64, image_of(tp, 0) * QA_L + 1, pol_a);
                    tma_load_2d_pair_hint(dst + 64 * 128, &tmA, &afull_bar[kb], kb * 64, image_of(tp, 1) * QA_L + 1, pol_a);
                }
                for (int hg = hg0; hg < hg0 + hg_per; ++hg) {
                    for (int kb = 0; kb < kblocks; ++kb, ++bit) {
                        if ((bit & 1u) != me) continue;
                        const uint32_t stage = bit % static_cast<uint32_t>(p.stages);
                        const uint32_t phase = (bit / static_cast<uint32_t>(p.stages)) & 1u;
                        mbar_wait_parked(&empty_bar[stage], phase ^ 1);
                        if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * kWStage);
                        tma_load_2d_pair_hint(smem_w + static_cast<size_t>(stage) * kWStage, &tmW, &full_bar[stage], kb * 64,
                                              hg * QA_NB + static_cast<int>(crank) * 96, pol_w);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ================================================================= pair MMA issuer (leader CTA): the qkv GEMM
        if (leader) {
            const uint32_t idesc = umma_idesc_16bit(256, QA_NB, F16);
            const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
            const uint64_t ad0 = umma_desc_sw128(smem_u32(smem_a));
            const uint64_t bd0 = umma_desc_sw128(smem_u32(smem_w));
            uint32_t stage = 0, phase = 0, uc = 0, it = 0;
            for (int u = first_u; u < n_units; u += u_step, ++uc) {
                for (int hl = 0; hl < hg_per; ++hl, ++it) {
                    mbar_wait_parked(tempty_bar, (it & 1u) ^ 1u);
                    tc_fence_after();
                    if constexpr (TRACE) { qa_stamp(p, 0, it, 0); }
                    for (int kb = 0; kb < kblocks; ++kb) {
                        if (hl == 0) mbar_wait_parked(&afull_bar[kb], uc & 1);
                        mbar_wait_parked(&full_bar[stage], phase);
                        tc_fence_after();
                        if constexpr (TRACE) { if (kb == 0) qa_stamp(p, 0, it, 1); }
                        if (elect_one()) {
                            umma_kblock<2>(tb, ad0 + static_cast<uint64_t>(kb * (G2_A_KB_BYTES >> 4)),
                                           bd0 + static_cast<uint64_t>(stage * (kWStage >> 4)), idesc, kb != 0 ? 1u : 0u);
                            umma_commit_pair(&empty_bar[stage]);
                            if (hl == hg_per - 1) umma_commit_pair(&aempty_bar[kb]);
                            if (kb == kblocks - 1) umma_commit_pair(tfull_bar);
                        }
                        __syncwarp();
                        if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
                    }
                    if constexpr (TRACE) { qa_stamp(p, 0, it, 2); }
                }
            }
        }
    } else if (warp == 2) {
        // TMEM allocator only
    } else if (warp < 8) {
        // ================================================================= epilogue: accumulator -> Q / K / V tiles
        const int q = warp & 3;
        uint32_t it = 0;
        for (int u = first_u; u < n_units; u += u_step) {
            const int hg0 = (u % p.n_split) * hg_per;
            for (int hg = hg0; hg < hg0 + hg_per; ++hg, ++it) {
                mbar_wait_parked(tfull_bar, it & 1u);
                tc_fence_after();
                if constexpr (TRACE) { if (q == 0) qa_stamp(p, 1, it, 0); }
                if (it > 0) mbar_wait_parked(qkv_free, (it - 1) & 1u);
                if constexpr (TRACE) { if (q == 0) qa_stamp(p, 1, it, 1); }
                drain3(0, hg, it);
                drain_finish();
                if (q == 0) {
                    issue_s(it);
                    if constexpr (TRACE) { qa_stamp(p, 3, it, 1); }
                }
                if constexpr (TRACE) { if (q == 0) qa_stamp(p, 1, it, 2); }
            }
        }
    } else if (warp < 16) {
        // ================================================================= softmax + output: one thread per query row
        const int h = (warp - 8) >> 2;
        const int q = warp & 3;
        const int r = q * 32 + lane, j = r >> 6;
        const uint32_t row_taddr_base = tmem_base + QT_S + QT_SW * h;
        const uint32_t row_taddr = row_taddr_base + (static_cast<uint32_t>(q * 32) << 16);
        uint8_t* prow = (h == 0 ? smem_q : smem_p1) + r * 128;
        const bool st0 = warp == 8;
        uint32_t it = 0;
        for (int u = first_u; u < n_units; u += u_step) {
            const int tp = u / p.n_split, hg0 = (u % p.n_split) * hg_per;
            const int img = image_of(tp, j);
            for (int hg = hg0; hg < hg0 + hg_per; ++hg, ++it) {
                const uint32_t par = it & 1u;
                mbar_wait_parked(cls_ready, par);                                  // vcls (and the class keys behind S) are in place
                mbar_wait_parked(&s_full[h], par);
                tc_fence_after();
                if constexpr (TRACE) { if (st0) qa_stamp(p, 2, it, 0); }
                // the row's 64 scores in two halves: the second half's TMEM read runs under the first half's exponentials
                // (taken against the first half's maximum and corrected afterwards if the second half holds a larger one)
                float s0[32], s1[32], sc;
                tmem_ld32_nowait(row_taddr + 64 * j, s0);
                tmem_ld1_nowait(row_taddr + 128 + j, sc);                   // score against this image's class key
                tmem_ld_wait32(s0);
                asm volatile("" : "+f"(sc));
                tmem_ld32_nowait(row_taddr + 64 * j + 32, s1);
                float m1 = sc;
#pragma unroll
                for (int i = 0; i < 32; ++i) m1 = fmaxf(m1, s0[i]);
                const float ms1 = m1 * p.scale_log2e;
                float sum = 0.f;
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    s0[i] = ex2_approx(fmaf(s0[i], p.scale_log2e, -ms1));
                    sum += s0[i];
                }
                tmem_ld_wait32(s1);
                float m = m1;
#pragma unroll
                for (int i = 0; i < 32; ++i) m = fmaxf(m, s1[i]);
                const float ms = m * p.scale_log2e;
                const float corr = ex2_approx(ms1 - ms);                    // 1 unless the second half raised the maximum
                const float pc = ex2_approx(fmaf(sc, p.scale_log2e, -ms));
                sum = fmaf(sum, corr, pc);
                if (h == 0) mbar_wait_parked(&s_full[1], par);              // both S are done: the Q tile may become head 0's P
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint4 w;
                    w.x = pack2<F16>(s0[8 * c] * corr, s0[8 * c + 1] * corr); w.y = pack2<F16>(s0[8 * c + 2] * corr, s0[8 * c + 3] * corr);
                    w.z = pack2<F16>(s0[8 * c + 4] * corr, s0[8 * c + 5] * corr); w.w = pack2<F16>(s0[8 * c + 6] * corr, s0[8 * c + 7] * corr);
                    *reinterpret_cast<uint4*>(prow + ((c ^ (r & 7)) << 4)) = w;
                }
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float e[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        e[i] = ex2_approx(fmaf(s1[8 * c + i], p.scale_log2e, -ms));
                        sum += e[i];
                    }
                    uint4 w;
                    w.x = pack2<F16>(e[0], e[1]); w.y = pack2<F16>(e[2], e[3]); w.z = pack2<F16>(e[4], e[5]); w.w = pack2<F16>(e[6], e[7]);
                    *reinterpret_cast<uint4*>(prow + (((4 + c) ^ (r & 7)) << 4)) = w;
                }
                fence_proxy_async_smem();                                   // P is an MMA operand
                tc_fence_before();                                          // S has been read: its columns may take O
                named_bar_sync(2 + h, 128);                                 // the head's four softmax warps
                if (q == 0) {                                               // O_h = P_h V_h, issued right here
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t idesc_o = umma_idesc_16bit(128, 32, F16) | (1u << 16);      // B operand (V) is MN-major
                        const uint64_t ph = umma_desc_sw128(smem_u32(h == 0 ? smem_q : smem_p1));
                        const uint64_t vd = umma_desc_sw128(smem_u32(smem_v)) + 4 * h;
                        umma_pv64(row_taddr_base, ph, vd, idesc_o);                               // image 0: keys = rows 0..63
                        umma_pv64(row_taddr_base + 32, ph, vd + (64 * 128 >> 4), idesc_o);        // image 1: rows 64..127
                        umma_commit(&o_full[h]);
                        umma_commit(qkv_free);
                    }
                    __syncwarp();
                }
                if constexpr (TRACE) { if (st0) qa_stamp(p, 2, it, 1); }
                const float inv = __frcp_rn(sum);
                mbar_wait_parked(&o_full[h], par);
                tc_fence_after();
                if constexpr (TRACE) { if (st0) qa_stamp(p, 2, it, 2); }
                const float* vc = vcls + par * 128 + j * 64 + h * 32;
                float4 vv[8];                                               // before the TMEM read (memory clobber), see the drain
#pragma unroll
                for (int c = 0; c < 8; ++c) vv[c] = *reinterpret_cast<const float4*>(vc + 4 * c);
                float o[32];
                tmem_ld32(row_taddr + 32 * j, o);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&o_free[h]);
                uint4 w[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float4 v0 = vv[2 * c], v1 = vv[2 * c + 1];
                    w[c].x = pack2<F16>(fmaf(pc, v0.x, o[8 * c]) * inv, fmaf(pc, v0.y, o[8 * c + 1]) * inv);
                    w[c].y = pack2<F16>(fmaf(pc, v0.z, o[8 * c + 2]) * inv, fmaf(pc, v0.w, o[8 * c + 3]) * inv);
                    w[c].z = pack2<F16>(fmaf(pc, v1.x, o[8 * c + 4]) * inv, fmaf(pc, v1.y, o[8 * c + 5]) * inv);
                    w[c].w = pack2<F16>(fmaf(pc, v1.z, o[8 * c + 6]) * inv, fmaf(pc, v1.w, o[8 * c + 7]) * inv);
                }
                // A thread owns a row (64 B of this head).  Stored straight from here, one instruction would touch 32 rows
                // 16 B at a time -- half sectors at a 768-byte stride -- and the SM's load/store path stays clogged for ~2k
                // cycles, stalling the drain warps' shared-memory traffic with it (measured).  Through 1 KB of staging per
                // warp, an instruction writes eight rows' full 64-byte segments instead.
                uint8_t* stg = out_stage + (warp - 8) * 1024;
                bf16* ob = p.out + (static_cast<size_t>(img) * QA_L + 1 + ((q * 32) & 63)) * p.C + hg * 64 + h * 32;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    if ((lane >> 4) == half) {
                        const int rr = lane & 15;
#pragma unroll
                        for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(stg + rr * 64 + ((c ^ ((rr >> 1) & 3)) << 4)) = w[c];
                    }
                    __syncwarp();
                    if (img < p.B) {
#pragma unroll
                        for (int k = 0; k < 2; ++k) {
                            const int rr = (lane >> 2) + 8 * k, c = lane & 3;
                            const uint4 val = *reinterpret_cast<const uint4*>(stg + rr * 64 + ((c ^ ((rr >> 1) & 3)) << 4));
                            *reinterpret_cast<uint4*>(ob + static_cast<size_t>(16 * half + rr) * p.C + c * 8) = val;
                        }
                    }
                    __syncwarp();
                }
                if constexpr (TRACE) { if (st0) qa_stamp(p, 2, it, 3); }
            }
        }
    } else {
        // ================================================================= class rows: warp = (image j, head h) of the pair
        const int cw = warp - 16;
        const int j = cw >> 1, h = cw & 1;
        const uint8_t* Kt = smem_k + j * 64 * 128;
        const uint8_t* Vt = smem_v + j * 64 * 128;
        float* qs = cls_scratch + cw * 96;                // the class query (32 dims, fp32)
        float* ps = qs + 32;                              // its 64 patch-key probabilities
        const int kg = lane >> 3, dg = lane & 7;          // P*V: lane = (16-key group, 4-dim group)
        uint32_t it = 0;
        for (int u = first_u; u < n_units; u += u_step) {
            const int tp = u / p.n_split, hg0 = (u % p.n_split) * hg_per;
            const int img = image_of(tp, j);
            for (int hg = hg0; hg < hg0 + hg_per; ++hg, ++it) {
                const uint32_t par = it & 1u;
                // this (image, head)'s class-token q | k | v: 32 values each, one per lane, from the class-row GEMM
                float qf = 0.f, kf = 0.f, vf = 0.f;
                uint4 kraw = make_uint4(0, 0, 0, 0);
                if (img < p.B) {
                    const uint16_t* g = reinterpret_cast<const uint16_t*>(p.qkv_cls) + static_cast<size_t>(img) * 3 * p.C + hg * 64 + h * 32;
                    qf = unpack1<F16>(__ldg(g + lane));
                    kf = unpack1<F16>(__ldg(g + p.C + lane));
                    vf = unpack1<F16>(__ldg(g + 2 * p.C + lane));
                    if (lane < 4) kraw = __ldg(reinterpret_cast<const uint4*>(g + p.C) + lane);
                }
                if (it > 0) mbar_wait_parked(qkv_free, (it - 1) & 1u);             // also: the output pass of it - 2 has read this parity
                if constexpr (TRACE) { if (cw == 0) qa_stamp(p, 4, it, 0); }
                mbar_wait_parked(tfull_bar, par);                           // second half of the accumulator drain first: it is
                tc_fence_after();                                           // what the next S and the next GEMM wait for
                drain3(3, hg, it);
                // class key -> row 128 + j of the K tile (swizzle term (128 + j) & 7 = j), so S gets a class-score column
                if (lane < 4) *reinterpret_cast<uint4*>(smem_k + (128 + j) * 128 + (((h * 4 + lane) ^ j) << 4)) = kraw;
                vcls[par * 128 + j * 64 + h * 32 + lane] = vf;
                qs[lane] = qf;
                drain_finish();                                             // (its fence covers the class key too)
                if (lane == 0) mbar_arrive(cls_ready);
                float scl = qf * kf;                                        // class query x class key
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) scl += __shfl_xor_sync(0xffffffffu, scl, o);
                if constexpr (TRACE) { if (cw == 0) qa_stamp(p, 4, it, 1); }
                // class query x patch keys: lane owns keys lane and lane + 32
                float s0 = 0.f, s1 = 0.f;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float4 qa = *reinterpret_cast<const float4*>(qs + 8 * c), qb = *reinterpret_cast<const float4*>(qs + 8 * c + 4);
                    const uint4 k0 = *reinterpret_cast<const uint4*>(Kt + lane * 128 + (((h * 4 + c) ^ (lane & 7)) << 4));
                    const uint4 k1 = *reinterpret_cast<const uint4*>(Kt + (lane + 32) * 128 + (((h * 4 + c) ^ (lane & 7)) << 4));
                    float2 a;
                    a = unpack2<F16>(k0.x); s0 = fmaf(qa.x, a.x, s0); s0 = fmaf(qa.y, a.y, s0);
                    a = unpack2<F16>(k0.y); s0 = fmaf(qa.z, a.x, s0); s0 = fmaf(qa.w, a.y, s0);
                    a = unpack2<F16>(k0.z); s0 = fmaf(qb.x, a.x, s0); s0 = fmaf(qb.y, a.y, s0);
                    a = unpack2<F16>(k0.w); s0 = fmaf(qb.z, a.x, s0); s0 = fmaf(qb.w, a.y, s0);
                    a = unpack2<F16>(k1.x); s1 = fmaf(qa.x, a.x, s1); s1 = fmaf(qa.y, a.y, s1);
                    a = unpack2<F16>(k1.y); s1 = fmaf(qa.z, a.x, s1); s1 = fmaf(qa.w, a.y, s1);
                    a = unpack2<F16>(k1.z); s1 = fmaf(qb.x, a.x, s1); s1 = fmaf(qb.y, a.y, s1);
                    a = unpack2<F16>(k1.w); s1 = fmaf(qb.z, a.x, s1); s1 = fmaf(qb.w, a.y, s1);
                }
                float m = fmaxf(scl, fmaxf(s0, s1));
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                const float ms = m * p.scale_log2e;
                const float e0 = ex2_approx(fmaf(s0, p.scale_log2e, -ms)), e1 = ex2_approx(fmaf(s1, p.scale_log2e, -ms));
                const float ec = ex2_approx(fmaf(scl, p.scale_log2e, -ms));
                float sum = e0 + e1;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                sum += ec;
                ps[lane] = e0;
                ps[lane + 32] = e1;
                __syncwarp();
                // P*V: this lane's 16 keys x 4 dims, then a sum over the four key groups
                float o0 = 0.f, o1 = 0.f, o2 = 0.f, o3 = 0.f;
#pragma unroll 4
                for (int i = 0; i < 16; ++i) {
                    const int key = kg * 16 + i;
                    const float pk = ps[key];
                    const uint2 vv = *reinterpret_cast<const uint2*>(Vt + key * 128 + (((h * 4 + (dg >> 1)) ^ (key & 7)) << 4) + (dg & 1) * 8);
                    const float2 f = unpack2<F16>(vv.x), g2 = unpack2<F16>(vv.y);
                    o0 = fmaf(pk, f.x, o0); o1 = fmaf(pk, f.y, o1); o2 = fmaf(pk, g2.x, o2); o3 = fmaf(pk, g2.y, o3);
                }
#pragma unroll
                for (int o = 8; o <= 16; o <<= 1) {
                    o0 += __shfl_xor_sync(0xffffffffu, o0, o); o1 += __shfl_xor_sync(0xffffffffu, o1, o);
                    o2 += __shfl_xor_sync(0xffffffffu, o2, o); o3 += __shfl_xor_sync(0xffffffffu, o3, o);
                }
                if (kg == 0 && img < p.B) {
                    const float4 vc = *reinterpret_cast<const float4*>(vcls + par * 128 + j * 64 + h * 32 + 4 * dg);
                    const float inv = __frcp_rn(sum);
                    uint2 w;
                    w.x = pack2<F16>(fmaf(ec, vc.x, o0) * inv, fmaf(ec, vc.y, o1) * inv);
                    w.y = pack2<F16>(fmaf(ec, vc.z, o2) * inv, fmaf(ec, vc.w, o3) * inv);
                    *reinterpret_cast<uint2*>(p.out + static_cast<size_t>(img) * QA_L * p.C + hg * 64 + h * 32 + 4 * dg) = w;
                }
                __syncwarp();
                if constexpr (TRACE) { if (cw == 0) qa_stamp(p, 4, it, 2); }
                if (lane == 0) mbar_arrive(qkv_free);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 2) tmem_dealloc<2>(tmem_base, 512);
}

}  // namespace ddc

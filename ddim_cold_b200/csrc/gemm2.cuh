// Persistent token GEMM on the 5th-gen tensor cores, everything moved by TMA.
//
//   OUT[M, N] = epilogue( A[M, K] * W[N, K]^T ),  A and W bf16 K-major, K <= 512.
//
// With CS = 2 two CTAs (one per SM of a TPC) form a pair: each stages its own 128-row A tile and HALF of
// every W tile, and the pair's leader issues cta_group::2 MMAs (M = 256) that read both halves, so every
// SM ingests only half of the weight stream -- the L2->SM fabric (~6 TB/s chip-wide), not the tensor
// pipe, is what bounds these GEMMs.
// A CTA (one per SM) walks 128-row M tiles round-robin.  Roles: warp 0 = TMA producer (the A
// tile, K/64 boxes of 64x128, then the W tiles bn x 64 through a ring), warp 1 = tcgen05.mma
// issuer (M=128, N=bn, K=16 per instruction, fp32 accumulators in TMEM), warp 2 = TMEM
// allocator, warps 4-11 = epilogue (thread = one output row, two warps per TMEM lane quarter).  A is released K-block by
// K-block while the last N tile is being multiplied, so the next tile's A loads overlap it.
//
// Two flavours:
//  * stream  (S_*): N tiles alternate between two TMEM accumulator stages; the epilogue
//    converts to bf16 in 32x128-byte staging slots and hands them to TMA stores.
//      S_BF16       out = bf16(acc + bias)                        qkv   (ViT.py:107)
//      S_BF16_GELU  out = bf16(gelu_erf(acc + bias))              fc1   (ViT.py:85-86)
//      S_F32        out = acc + bias  (fp32, direct stores)       head  (ViT.py:213)
//  * row-block (R_*): N == embed_dim, the whole 128 x C fp32 row block lives in TMEM.
//      pass 1  v = acc + bias + X   (R_RESID, X prefetched by TMA; ViT.py:136-137)
//              v = acc + posb[l] + time[t]  (R_EMBED; ViT.py:154,202-205: posb rows come by TMA from a table
//                                            extended to L + 127 rows so a tile's 128 rows are contiguous in it)
//              v -> X (fp32, TMA store), v -> back into TMEM, shifted row sums (one-pass statistics)
//      pass 2  XN = bf16(LayerNorm(v) * g + b) -> TMA store.  XN is the *next* GEMM's A
//              operand, so LayerNorm (ViT.py:133,137,212) never costs its own pass over HBM.
#pragma once

#include "common.cuh"

namespace ddc {

enum { S_BF16 = 0, S_BF16_GELU = 1, S_F32 = 2, R_RESID = 3, R_EMBED = 4 };

constexpr int G2_BM = 128;
constexpr int G2_BK = 64;
constexpr int G2_THREADS = 384;             // warps 0 and 3 TMA producers, warp1 MMA, warp2 TMEM, warps4-11 epilogue
constexpr int G2_EPI_WARPS = 8;             // two warps per TMEM lane quarter, alternating column chunks
constexpr int G2_A_KB_BYTES = G2_BM * 128;
constexpr int G2_GSLOT_BYTES = 128 * 128;   // staging unit of an epilogue group (4 warps): 128 rows x 128 B
constexpr int G2_ES_STREAM = 1;             // staging slots per epilogue group (a second slot costs three ring stages: measured slower)
constexpr int G2_ES_ROW = 2;                // row-block: residual chunk in flight + chunk being processed
constexpr int G2_TAIL_BYTES = 3072;         // barriers, TMEM slot, row-statistics exchange

struct Gemm2Params {
    int M, N, K, bn, stages, num_tiles;
    const float* bias;
    float* out_f32;          // S_F32
    int ld_f32;
    int f32_tma;             // S_F32: stage 32-column chunks and store them by TMA (tmO = fp32 map) instead of per-row stores
    const float* ln_g;       // R_*: parameters of the LayerNorm that FOLLOWS this GEMM
    const float* ln_b;
    float ln_eps;
    const float* posb;       // R_EMBED: [L + 127, N] with row i = position row i % L (only read through tmR)
    int L;
    const float* time_tab;
    const int* t_img;
    const StepEntry* tbl;
    const int* cur;
    int B;
    int T;                   // R_EMBED: rows of time_tab; a per-image t outside [0, T) is clamped into it (0 = unchecked)
    int balanced;            // streaming flavours: split (M tile, N tile) units evenly over the CTA pairs
    int dbg_flags;           // probes only: 1 = no X store, 2 = no residual load, 4 = no tmem_st, 8 = skip LN passes
    long long* dbg;          // optional timeline buffer (probes only): [cta < 2][role 4][64] SM clock stamps
};

__device__ __forceinline__ void g2_stamp(const Gemm2Params& p, int role, int idx) {
    if (p.dbg != nullptr && blockIdx.x < 2 && idx < 64) p.dbg[(blockIdx.x * 4 + role) * 64 + idx] = clock64();
}

__host__ __device__ inline uint32_t g2_pow2_cols(uint32_t c) {
    uint32_t p = 32;
    while (p < c) p <<= 1;
    return p;
}

// bias (and, row-block, LayerNorm gamma/beta) are staged in shared memory: with the full carve-out
// given to shared memory there is no L1 left, and a per-chunk __ldg would pay an L2 round trip.
__host__ __device__ inline size_t gemm2_param_bytes(int N, bool rowblock) { return static_cast<size_t>(rowblock ? 3 : 1) * N * 4; }

inline size_t gemm2_smem_bytes(int K, int N, int bn, int stages, bool rowblock) {
    return 1024 + static_cast<size_t>(K / G2_BK) * G2_A_KB_BYTES + static_cast<size_t>(stages) * bn * 128 +
           static_cast<size_t>(2) * (rowblock ? G2_ES_ROW : G2_ES_STREAM) * G2_GSLOT_BYTES +
           ((gemm2_param_bytes(N, rowblock) + 127) & ~static_cast<size_t>(127)) + G2_TAIL_BYTES;
}

// byte offset of 16-byte chunk `j` of row `r` inside a 128-byte-swizzled 32-row slot
__device__ __forceinline__ uint32_t slot_off(int r, int j) { return r * 128 + ((j ^ (r & 7)) << 4); }

// Probes only (scripts/probes define DDC_PROBE_TIMING): per-section cycle accounting of one epilogue warp.
#ifdef DDC_PROBE_TIMING
#define PT_DECL long long pt_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, pt_last = clock64()
#define PT(k) do { const long long pt_now = clock64(); pt_acc[k] += pt_now - pt_last; pt_last = pt_now; } while (0)
#define PT_FLUSH(p) do { if ((p).dbg != nullptr && blockIdx.x == 0 && warp == 4 && lane == 0) for (int i = 0; i < 8; ++i) (p).dbg[3 * 64 + i] = pt_acc[i]; } while (0)
#else
#define PT_DECL
#define PT(k)
#define PT_FLUSH(p)
#endif

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

template <int EPI, bool F16, int CS>
__global__ void __launch_bounds__(G2_THREADS, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
             const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmR,
             const __grid_constant__ CUtensorMap tmN, const Gemm2Params p) {
    constexpr bool ROWBLOCK = (EPI == R_RESID || EPI == R_EMBED);
    constexpr int ES = ROWBLOCK ? G2_ES_ROW : G2_ES_STREAM;
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

    const int kblocks = p.K / G2_BK;
    const int num_nt = p.N / p.bn;
    const uint32_t b_stage_bytes = static_cast<uint32_t>(p.bn / CS) * 128u;   // this CTA's share of a W tile
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem_a + static_cast<size_t>(kblocks) * G2_A_KB_BYTES;
    uint8_t* smem_e = smem_b + static_cast<size_t>(p.stages) * b_stage_bytes;
    float* smem_p = reinterpret_cast<float*>(smem_e + 2 * ES * G2_GSLOT_BYTES);   // bias | gamma | beta
    uint64_t* bars = reinterpret_cast<uint64_t*>(
        reinterpret_cast<uint8_t*>(smem_p) + ((gemm2_param_bytes(p.N, ROWBLOCK) + 127) & ~static_cast<size_t>(127)));
    uint64_t* full_bar = bars;             // [8]  W ring: TMA -> MMA
    uint64_t* empty_bar = bars + 8;        // [8]  W ring: MMA -> TMA
    uint64_t* afull_bar = bars + 16;       // [8]  A K-block landed
    uint64_t* aempty_bar = bars + 24;      // [8]  A K-block no longer needed by this tile
    uint64_t* tfull_bar = bars + 32;       // [4]  accumulator stage / N tile complete
    uint64_t* tempty_bar = bars + 36;      // [4]  accumulator stage drained
    uint64_t* res_bar = bars + 40;         // [16] residual chunk landed (per epilogue group x slot)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 56);
    float* red = reinterpret_cast<float*>(bars + 64);     // [2 stats][2 halves][128 rows] partial row statistics

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    // The CTAs of a pair take adjacent M tiles; p.num_tiles is a multiple of CS, so both run the same number of
    // iterations (a surplus tile stores nothing).
    const uint32_t crank = CS > 1 ? cluster_ctarank() : 0u;
    const int first_tile = static_cast<int>((blockIdx.x / CS) * CS + crank);
    const int tile_step = gridDim.x;
    // Work split.  Row-block flavours need whole rows, so a CTA (pair) takes whole M tiles, strided by the grid.
    // The streaming flavours split the (M tile, N tile) units of the whole problem into one contiguous, equally
    // long run per CTA (pair): with 260 pair tiles on 74 pairs, whole tiles would cost four rounds where 3.5 are
    // needed.  A run covers segments (M tile, [nt0, nt1)); a pair loads the A rows of a split tile a second time.
    const int seg_pairs = static_cast<int>(gridDim.x) / CS, seg_pidx = static_cast<int>(blockIdx.x) / CS;
    const long long seg_units = static_cast<long long>(p.num_tiles / CS) * num_nt;
    const int seg_u0 = static_cast<int>(seg_units * seg_pidx / seg_pairs);
    const int seg_u1 = static_cast<int>(seg_units * (seg_pidx + 1) / seg_pairs);
    auto segment = [&](uint32_t k, int& tile, int& nt0, int& nt1) -> bool {
        if (ROWBLOCK || !p.balanced) {
            tile = first_tile + static_cast<int>(k) * tile_step;
            nt0 = 0;
            nt1 = num_nt;
            return tile < p.num_tiles;
        }
        const int ptile = seg_u0 / num_nt + static_cast<int>(k);
        tile = ptile * CS + static_cast<int>(crank);
        nt0 = k == 0 ? seg_u0 % num_nt : 0;
        nt1 = seg_u1 - ptile * num_nt < num_nt ? seg_u1 - ptile * num_nt : num_nt;
        return nt1 > nt0;
    };
    const bool leader = crank == 0;
    // stream flavour: as many accumulator stages as the 512 TMEM columns hold (max 4), so the MMA issuer can run
    // several N tiles ahead of the epilogue and the drain handshake stays off the critical path
    const uint32_t acc_stride = g2_pow2_cols(p.bn);
    const uint32_t nacc = (512u / acc_stride) < 4u ? (512u / acc_stride) : 4u;
    const uint32_t tmem_cols = ROWBLOCK ? g2_pow2_cols(p.N) : g2_pow2_cols(nacc * acc_stride);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        if (EPI != S_F32 || p.f32_tma) tma_prefetch_desc(&tmO);
        if (ROWBLOCK) tma_prefetch_desc(&tmR);
        if (ROWBLOCK) tma_prefetch_desc(&tmN);
        for (int i = 0; i < 8; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
            mbar_init(&afull_bar[i], 1);
            mbar_init(&aempty_bar[i], 1);
        }
        for (int i = 0; i < 4; ++i) mbar_init(&tfull_bar[i], 1);
        for (int i = 0; i < 4; ++i) mbar_init(&tempty_bar[i], G2_EPI_WARPS * CS);   // both CTAs' epilogue warps
        for (int i = 0; i < 16; ++i) mbar_init(&res_bar[i], 1);
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc<CS>(tmem_slot, tmem_cols);
        tmem_relinquish<CS>();
    }
    if (warp >= 4) {
        for (int i = threadIdx.x - 128; i < p.N; i += G2_EPI_WARPS * 32) {
            if (EPI != R_EMBED) smem_p[i] = p.bias[i];      // R_EMBED: the time row, staged after pdl_wait below
            if (ROWBLOCK) {
                smem_p[p.N + i] = p.ln_g[i];
                smem_p[2 * p.N + i] = p.ln_b[i];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();              // peers' barriers are initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();        // set-up above touched only weights; activations and the step counter come from earlier kernels
    if (EPI == R_EMBED && warp >= 4) {
        // "bias" of the patch GEMM = the timestep embedding row (ViT.py:205).  In the sampling loop every image shares
        // t (read from the device step table); forward() with per-image t adds the rows per thread instead.
        const float* trow = p.t_img == nullptr ? p.time_tab + static_cast<size_t>(p.tbl[*p.cur].t) * p.N : nullptr;
        for (int i = threadIdx.x - 128; i < p.N; i += G2_EPI_WARPS * 32) smem_p[i] = trow != nullptr ? trow[i] : 0.f;
        named_bar_sync(7, G2_EPI_WARPS * 32);
    }

    if (warp == 0 || warp == 3) {
        // ================================================================= TMA producers
        // One thread needs several hundred cycles per TMA instruction (wait, expect_tx, issue), more than a
        // K block's MMAs take, so two warps issue alternate loads: producer 0 the even ones, producer 1 the odd.
        if (lane == 0) {
            const uint32_t me = warp == 0 ? 0u : 1u;
            uint32_t bit = 0, tcount = 0;          // bit: running index of W loads (stage = bit % stages)
            // L2 priorities: activations are read once (evict_first), the weights by every CTA all the time (evict_last)
            const bool nohint = (p.dbg_flags & 16) != 0;
            const uint64_t pol_a = nohint ? l2_policy_evict_normal() : l2_policy_evict_first();
            const uint64_t pol_w = nohint ? l2_policy_evict_normal() : l2_policy_evict_last();
            for (int tile, nt0, nt1; segment(tcount, tile, nt0, nt1); ++tcount) {
                const int m0 = tile * G2_BM;
                for (int kb = static_cast<int>(me); kb < kblocks; kb += 2) {
                    mbar_wait(&aempty_bar[kb], (tcount & 1) ^ 1);
                    if (CS == 1) {
                        mbar_arrive_expect_tx(&afull_bar[kb], G2_A_KB_BYTES);
                        tma_load_2d_hint(smem_a + static_cast<size_t>(kb) * G2_A_KB_BYTES, &tmA, &afull_bar[kb], kb * G2_BK, m0, pol_a);
                    } else {      // both CTAs' bytes are credited to the leader's barrier
                        if (leader) mbar_arrive_expect_tx(&afull_bar[kb], CS * G2_A_KB_BYTES);
                        tma_load_2d_pair_hint(smem_a + static_cast<size_t>(kb) * G2_A_KB_BYTES, &tmA, &afull_bar[kb], kb * G2_BK, m0, pol_a);
                    }
                }
                int ntile, na, nb;
                if (me == 0 && segment(tcount + 1, ntile, na, nb))      // next tile's A rows -> L2 while this one runs
                    for (int kb = 0; kb < kblocks; ++kb) tma_prefetch_l2_2d(&tmA, kb * G2_BK, ntile * G2_BM);
                for (int nt = nt0; nt < nt1; ++nt) {
                    for (int kb = 0; kb < kblocks; ++kb, ++bit) {
                        if ((bit & 1u) != me) continue;
                        const uint32_t stage = bit % static_cast<uint32_t>(p.stages);
                        const uint32_t phase = (bit / static_cast<uint32_t>(p.stages)) & 1u;
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        if (CS == 1) {
                            mbar_arrive_expect_tx(&full_bar[stage], b_stage_bytes);
                            tma_load_2d_hint(smem_b + static_cast<size_t>(stage) * b_stage_bytes, &tmB, &full_bar[stage],
                                             kb * G2_BK, nt * p.bn, pol_w);
                        } else {      // this CTA stages rows [crank*bn/2, +bn/2) of the W tile
                            if (leader) mbar_arrive_expect_tx(&full_bar[stage], CS * b_stage_bytes);
                            tma_load_2d_pair_hint(smem_b + static_cast<size_t>(stage) * b_stage_bytes, &tmB, &full_bar[stage],
                                                  kb * G2_BK, nt * p.bn + static_cast<int>(crank) * (p.bn / CS), pol_w);
                        }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1 && leader) {
        // ================================================================= MMA issuer (pair leader only)
        const uint32_t idesc = umma_idesc_16bit(G2_BM * CS, static_cast<uint32_t>(p.bn), F16);
        const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);     // warp-uniform for ptxas
        const uint64_t ad0 = umma_desc_sw128(smem_u32(smem_a));
        const uint64_t bd0 = umma_desc_sw128(smem_u32(smem_b));
        const uint32_t b_stage_units = b_stage_bytes >> 4;
        uint32_t stage = 0, phase = 0, tcount = 0, acc_it = 0;
        for (int tile, nt0, nt1; segment(tcount, tile, nt0, nt1); ++tcount) {
            for (int nt = nt0; nt < nt1; ++nt) {
                uint32_t d_tmem;
                uint64_t* done_bar;
                if (ROWBLOCK) {
                    if (nt == 0) mbar_wait(&tempty_bar[0], (tcount & 1) ^ 1);
                    d_tmem = tb + static_cast<uint32_t>(nt * p.bn);
                    done_bar = &tfull_bar[nt];
                } else {
                    const uint32_t acc = acc_it % nacc;
                    mbar_wait(&tempty_bar[acc], ((acc_it / nacc) & 1) ^ 1);
                    d_tmem = tb + acc * acc_stride;
                    done_bar = &tfull_bar[acc];
                    ++acc_it;
                }
                tc_fence_after();
                if (lane == 0) g2_stamp(p, 1, (tcount * 8 + nt) * 2);
                for (int kb = 0; kb < kblocks; ++kb) {
                    if (nt == nt0) mbar_wait(&afull_bar[kb], tcount & 1);
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    if (elect_one()) {
                        umma_kblock<CS>(d_tmem, ad0 + static_cast<uint64_t>(kb * (G2_A_KB_BYTES >> 4)),
                                        bd0 + static_cast<uint64_t>(stage * b_stage_units), idesc, kb != 0 ? 1u : 0u);
                        if (CS == 1) {
                            umma_commit(&empty_bar[stage]);
                            if (nt == nt1 - 1) umma_commit(&aempty_bar[kb]);
                            if (kb == kblocks - 1) umma_commit(done_bar);
                        } else {      // both CTAs' producers / epilogues are released
                            umma_commit_pair(&empty_bar[stage]);
                            if (nt == nt1 - 1) umma_commit_pair(&aempty_bar[kb]);
                            if (kb == kblocks - 1) umma_commit_pair(done_bar);
                        }
                    }
                    __syncwarp();
                    if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
                }
                if (lane == 0) g2_stamp(p, 1, (tcount * 8 + nt) * 2 + 1);
            }
        }
    } else if (warp >= 4) {
        // ================================================================= epilogue
        // warp w drains TMEM lanes [32*(w%4), +32) (thread = one output row).  The eight warps form two groups
        // (hf = 0/1) of four that take alternating column chunks; a group shares one [128 rows x 128 B] staging
        // slot per chunk and its leader issues ONE 16 KB TMA operation for it (a TMA instruction costs the SM's
        // TMA unit and the issuing thread a few hundred cycles whatever its size).
        const int q = warp & 3;
        const int hf = (warp - 4) >> 2;
        const bool gleader = q == 0 && lane == 0;
        const int gbar = 1 + hf;                                  // named barrier of this group (128 threads)
        uint8_t* grp_slots = smem_e + static_cast<size_t>(hf) * ES * G2_GSLOT_BYTES;
        const int trow = q * 32 + lane;                           // this thread's row inside the 128-row tile
        const uint32_t lane_taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        const float* s_bias = smem_p;
        uint32_t tcount = 0;
        if (!ROWBLOCK) {
            uint32_t acc_it = 0, si = 0;            // si: staging-slot counter of this group
            PT_DECL;
            const int cpn = p.bn / 64;              // 64-column chunks per N tile (bf16 flavours)
            for (int tile, nt0, nt1; segment(tcount, tile, nt0, nt1); ++tcount) {
                const int m0 = tile * G2_BM;
                const int row = m0 + trow;
                for (int nt = nt0; nt < nt1; ++nt, ++acc_it) {
                    const uint32_t acc = acc_it % nacc;
                    PT(0);
                    mbar_wait(&tfull_bar[acc], (acc_it / nacc) & 1);
                    tc_fence_after();
                    PT(1);
                    if (warp == 4 && lane == 0) g2_stamp(p, 2, (tcount * 8 + nt) * 2);
                    const uint32_t t_row = lane_taddr + acc * acc_stride;
                    if (EPI == S_F32 && p.f32_tma) {
                        // fp32 output in 32-column chunks: [128 rows x 128 B] slot per group, one TMA store each
                        const int cpn32 = p.bn / 32;
                        for (int c = 0; c < cpn32; ++c) {
                            if (((nt * cpn32 + c) & 1) != hf) continue;
                            uint8_t* slot = grp_slots + (si++ % ES) * G2_GSLOT_BYTES;
                            if (gleader) bulk_wait_read<ES - 1>();
                            named_bar_sync(gbar, 128);
                            const int n0 = nt * p.bn + c * 32;
                            float v[32];
                            tmem_ld32(t_row + c * 32, v);
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const float4 b = *reinterpret_cast<const float4*>(s_bias + n0 + 4 * j);
                                *reinterpret_cast<float4*>(slot + slot_off(trow, j)) =
                                    make_float4(v[4 * j] + b.x, v[4 * j + 1] + b.y, v[4 * j + 2] + b.z, v[4 * j + 3] + b.w);
                            }
                            fence_proxy_async_smem();
                            named_bar_sync(gbar, 128);
                            if (gleader && m0 < p.M) {
                                tma_store_2d(&tmO, slot, n0, m0);
                                bulk_commit();
                            }
                        }
                    } else if (EPI == S_F32) {
                        for (int c = 16 * hf; c < p.bn; c += 32) {
                            float v[16];
                            tmem_ld16(t_row + c, v);
                            if (row < p.M) {
                                const int n0 = nt * p.bn + c;
                                float4* dst = reinterpret_cast<float4*>(p.out_f32 + static_cast<size_t>(row) * p.ld_f32 + n0);
#pragma unroll
                                for (int i = 0; i < 4; ++i) {
                                    const float4 b = *reinterpret_cast<const float4*>(s_bias + n0 + 4 * i);
                                    dst[i] = make_float4(v[4 * i] + b.x, v[4 * i + 1] + b.y, v[4 * i + 2] + b.z,
                                                         v[4 * i + 3] + b.w);
                                }
                            }
                        }
                    } else {
                        for (int c = 0; c < cpn; ++c) {
                            if (((nt * cpn + c) & 1) != hf) continue;
                            uint8_t* slot = grp_slots + (si++ % ES) * G2_GSLOT_BYTES;
                            PT(0);
                            if (gleader) bulk_wait_read<ES - 1>();  // the slot's previous store has read it
                            PT(2);
                            named_bar_sync(gbar, 128);
                            PT(3);
                            const int n0 = nt * p.bn + c * 64;
#pragma unroll
                            for (int h = 0; h < 2; ++h) {
                                float v[32];
                                tmem_ld32(t_row + c * 64 + 32 * h, v);
                                PT(4);
#pragma unroll
                                for (int i = 0; i < 32; i += 4) {
                                    const float4 b = *reinterpret_cast<const float4*>(s_bias + n0 + 32 * h + i);
                                    v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
                                }
                                if (EPI == S_BF16_GELU) {
#pragma unroll
                                    for (int i = 0; i < 32; ++i) v[i] = gelu_erf(v[i]);
                                }
#pragma unroll
                                for (int j = 0; j < 4; ++j) {
                                    uint4 o;
                                    o.x = pack2<F16>(v[8 * j], v[8 * j + 1]);
                                    o.y = pack2<F16>(v[8 * j + 2], v[8 * j + 3]);
                                    o.z = pack2<F16>(v[8 * j + 4], v[8 * j + 5]);
                                    o.w = pack2<F16>(v[8 * j + 6], v[8 * j + 7]);
                                    *reinterpret_cast<uint4*>(slot + slot_off(trow, 4 * h + j)) = o;
                                }
                                PT(5);
                            }
                            fence_proxy_async_smem();
                            named_bar_sync(gbar, 128);
                            PT(6);
                            if (gleader && m0 < p.M) {
                                tma_store_2d(&tmO, slot, n0, m0);
                                bulk_commit();
                            }
                            PT(7);
                        }
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) { if (CS == 1) mbar_arrive(&tempty_bar[acc]); else mbar_arrive_leader(&tempty_bar[acc]); }
                    if (warp == 4 && lane == 0) g2_stamp(p, 2, (tcount * 8 + nt) * 2 + 1);
                }
            }
            PT_FLUSH(p);
        } else {
            // ---------------------------------------------------------- row-block flavour
            const int nchunk = p.N / 32;            // fp32 chunks of 128 B per row; this group takes c % 2 == hf
            const float inv_n = 1.0f / static_cast<float>(p.N);
            const float* s_gamma = smem_p + p.N;
            const float* s_beta = smem_p + 2 * p.N;
            float* red_sum = red + hf * 128 + trow;
            float* red_sq = red + 256 + hf * 128 + trow;
            const float* oth_sum = red + (hf ^ 1) * 128 + trow;
            const float* oth_sq = red + 256 + (hf ^ 1) * 128 + trow;
            uint64_t* my_res = res_bar + hf * ES;
            uint32_t ri = 0;                        // staging-slot counter (this group)
            uint32_t res_phase = 0;                 // per-slot parity of the residual barriers (bit s)
            // The residual X rows come from DRAM; with one chunk of look-ahead per group every chunk would pay a
            // full DRAM latency.  Their addresses are known a whole tile early, so the group leader pulls the NEXT
            // tile's chunks into L2 while this tile is processed (and this tile's right away for the first one).
            if (EPI == R_RESID && gleader && first_tile < p.num_tiles)
                for (int c = hf; c < nchunk; c += 2) tma_prefetch_l2_2d(&tmR, c * 32, first_tile * G2_BM);
            for (int tile = first_tile; tile < p.num_tiles; tile += tile_step, ++tcount) {
                const int m0 = tile * G2_BM;
                const int row = m0 + trow;
                if (EPI == R_RESID && gleader && tile + tile_step < p.num_tiles)
                    for (int c = hf; c < nchunk; c += 2) tma_prefetch_l2_2d(&tmR, c * 32, (tile + tile_step) * G2_BM);
                const float* embed_time = nullptr;      // per-image t only (forward())
                if (EPI == R_EMBED && p.t_img != nullptr) {
                    int b = row / p.L;
                    if (b > p.B - 1) b = p.B - 1;
                    int tt = p.t_img[b];             // the host class raises IndexError first; a raw C-ABI caller gets a clamp
                    if (p.T > 0) tt = tt < 0 ? 0 : (tt >= p.T ? p.T - 1 : tt);
                    embed_time = p.time_tab + static_cast<size_t>(tt) * p.N;
                }
                // residual rows: X[m0 ..] (R_RESID) or the position rows of these tokens, (m0 % L) + r in the table
                const int r_row = EPI == R_EMBED ? m0 % p.L : m0;
                if (hf < nchunk && gleader && !(p.dbg_flags & 2)) {
                    bulk_wait_read<0>();            // slots may still feed the previous tile's XN stores
                    const uint32_t s = ri % ES;
                    mbar_arrive_expect_tx(&my_res[s], G2_GSLOT_BYTES);
                    tma_load_2d(grp_slots + s * G2_GSLOT_BYTES, &tmR, &my_res[s], hf * 32, r_row);
                }
                if (warp == 4 && lane == 0) g2_stamp(p, 2, tcount * 4);
                // one-pass statistics: sums around a shift (the row's first value in this group's columns, so nothing
                // cancels), merged exactly with the partner group's half (same scheme as block_tail.cuh)
                float shift = 0.f, s1 = 0.f, s2 = 0.f;
                int waited_nt = -1;
                for (int c = hf; c < nchunk; c += 2, ++ri) {
                    const uint32_t s = ri % ES;
                    uint8_t* slot = grp_slots + s * G2_GSLOT_BYTES;
                    if (gleader) {
                        bulk_wait_read<0>();        // the other slot's store (previous chunk) has released it
                        if (c + 2 < nchunk && !(p.dbg_flags & 2)) {
                            const uint32_t s2 = (ri + 1) % ES;
                            mbar_arrive_expect_tx(&my_res[s2], G2_GSLOT_BYTES);
                            tma_load_2d(grp_slots + s2 * G2_GSLOT_BYTES, &tmR, &my_res[s2], (c + 2) * 32, r_row);
                        }
                    }
                    const int n0 = c * 32;
                    const int nt = n0 / p.bn;
                    if (nt != waited_nt) {
                        mbar_wait(&tfull_bar[nt], tcount & 1);
                        tc_fence_after();
                        waited_nt = nt;
                    }
                    float v[32];
                    tmem_ld32(lane_taddr + n0, v);
                    if (!(p.dbg_flags & 2)) {
                        mbar_wait(&my_res[s], (res_phase >> s) & 1);
                        res_phase ^= (1u << s);
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 r = *reinterpret_cast<const float4*>(slot + slot_off(trow, j));
                        const float4 b = *reinterpret_cast<const float4*>(s_bias + n0 + 4 * j);
                        v[4 * j] += r.x + b.x; v[4 * j + 1] += r.y + b.y;
                        v[4 * j + 2] += r.z + b.z; v[4 * j + 3] += r.w + b.w;
                    }
                    if (EPI == R_EMBED && embed_time != nullptr) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 b = __ldg(reinterpret_cast<const float4*>(embed_time + n0 + 4 * j));
                            v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
                        }
                    }
                    if (c == hf) shift = v[0];
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float d = v[i] - shift;
                        s1 += d;
                        s2 = fmaf(d, d, s2);
                    }
                    if (!(p.dbg_flags & 4)) tmem_st32(lane_taddr + n0, v);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<float4*>(slot + slot_off(trow, j)) =
                            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    fence_proxy_async_smem();
                    named_bar_sync(gbar, 128);
                    if (gleader && m0 < p.M && !(p.dbg_flags & 1)) {
                        tma_store_2d(&tmO, slot, n0, m0);
                        bulk_commit();
                    }
                }
                // ---- row statistics: each group turns its sums into (mean, M2) of its half of the row; halves of equal
                // count n/2 merge as mean = (m0 + m1)/2, M2 = M2_0 + M2_1 + n/4 (m0 - m1)^2  (Chan et al.)
                if (warp == 4 && lane == 0) g2_stamp(p, 2, tcount * 4 + 1);
                const float half_n = 0.5f * static_cast<float>(p.N);
                const float mg = shift + s1 / half_n;
                const float m2g = s2 - s1 * s1 / half_n;
                *red_sum = mg;
                *red_sq = m2g;
                tc_fence_before();                  // the partner's pass 3 reads TMEM columns this warp stored
                named_bar_sync(3 + q, 64);
                tc_fence_after();
                const float mo = *oth_sum, m2o = *oth_sq;
                const float mean = 0.5f * (mg + mo);
                const float dm = mg - mo;
                const float rstd = rsqrtf((m2g + m2o + 0.5f * half_n * dm * dm) * inv_n + p.ln_eps);
                const float nmr = -mean * rstd;
                named_bar_sync(3 + q, 64);          // both partners have read the exchange slots before they are reused
                if (warp == 4 && lane == 0) g2_stamp(p, 2, tcount * 4 + 2);
                for (int c = hf; c < ((p.dbg_flags & 8) ? 0 : p.N / 64); c += 2, ++ri) {
                    uint8_t* slot = grp_slots + (ri % ES) * G2_GSLOT_BYTES;
                    if (gleader) bulk_wait_read<ES - 1>();
                    named_bar_sync(gbar, 128);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        float v[32];
                        const int n0 = c * 64 + 32 * h;
                        tmem_ld32(lane_taddr + n0, v);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float4 g0 = *reinterpret_cast<const float4*>(s_gamma + n0 + 8 * j);
                            const float4 g1 = *reinterpret_cast<const float4*>(s_gamma + n0 + 8 * j + 4);
                            const float4 b0 = *reinterpret_cast<const float4*>(s_beta + n0 + 8 * j);
                            const float4 b1 = *reinterpret_cast<const float4*>(s_beta + n0 + 8 * j + 4);
                            uint4 o;
                            o.x = pack2<F16>(fmaf(fmaf(v[8 * j], rstd, nmr), g0.x, b0.x), fmaf(fmaf(v[8 * j + 1], rstd, nmr), g0.y, b0.y));
                            o.y = pack2<F16>(fmaf(fmaf(v[8 * j + 2], rstd, nmr), g0.z, b0.z), fmaf(fmaf(v[8 * j + 3], rstd, nmr), g0.w, b0.w));
                            o.z = pack2<F16>(fmaf(fmaf(v[8 * j + 4], rstd, nmr), g1.x, b1.x), fmaf(fmaf(v[8 * j + 5], rstd, nmr), g1.y, b1.y));
                            o.w = pack2<F16>(fmaf(fmaf(v[8 * j + 6], rstd, nmr), g1.z, b1.z), fmaf(fmaf(v[8 * j + 7], rstd, nmr), g1.w, b1.w));
                            *reinterpret_cast<uint4*>(slot + slot_off(trow, 4 * h + j)) = o;
                        }
                    }
                    fence_proxy_async_smem();
                    named_bar_sync(gbar, 128);
                    if (gleader && m0 < p.M) {
                        tma_store_2d(&tmN, slot, c * 64, m0);
                        bulk_commit();
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (warp == 4 && lane == 0) g2_stamp(p, 2, tcount * 4 + 3);
                if (lane == 0) { if (CS == 1) mbar_arrive(&tempty_bar[0]); else mbar_arrive_leader(&tempty_bar[0]); }
            }
        }
        if (gleader) bulk_wait_all();        // smem must outlive the last TMA stores
        __syncwarp();
    }

    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();              // no CTA leaves while a peer may still write its smem / barriers
    if (warp == 2) tmem_dealloc<CS>(tmem_base, tmem_cols);
}

}  // namespace ddc

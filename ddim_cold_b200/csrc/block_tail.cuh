// Fused tail of a transformer block (ViT.py:136-137 with Mlp.forward :84-90 and the next LayerNorm):
//
//     X1 = X + O * Wproj^T + b_proj                       (attention projection + residual)
//     H  = gelu_erf(LayerNorm2(X1) * Wfc1^T + b_fc1)
//          (the affine part of every LayerNorm is folded into the weight and bias of the Linear that consumes it at pack
//           time -- W' = W o gamma, b' = b + W beta -- so the kernels only normalise: no per-column vectors to load)
//     X' = X1 + H * Wfc2^T + b_fc2                        (MLP + residual)        -> X  (fp32)
//     XN = LayerNorm_next(X')                                                      -> XN (16 bit)
//
// in ONE persistent kernel: three GEMMs share a tile's fixed costs, X is read and written once per block, and
// the 16-bit intermediates (LayerNorm2 output, H) never leave the SM.  Per 128-row tile of a CTA pair
// (cta_group::2, M = 256), with C = embed_dim (multiple of 128, <= 384), NB = C/128, KB = C/64:
//
//   TMEM   cols [0, C)      the fp32 row block: proj accumulator -> X1 -> fc2 accumulates on top of it
//          cols [C, C+128)  fc1 accumulator for one 128-wide hidden block
//   smem   Abuf  KB x 16 KB  O tile (TMA), later overwritten by LayerNorm2(X1) in the same swizzled K-major layout
//          Hbuf  2 x 32 KB   one hidden block of H each (two K blocks); doubles as staging for residual loads
//                            (phase 1) and for the X / XN stores (phase 3) while no hidden block is live
//          W ring            64-row halves of 128 x 64 weight tiles, gk K blocks per stage, streamed in MMA order
//
//   MMA order   proj(n = 0..NB-1) | for j: fc1(j) ; fc2 partial(j): X1[:, n-tile] += H_j * Wfc2[n-tile, j-th K range]
//   epilogue    epi1: + residual + bias -> TMEM, LayerNorm2 (shifted one-pass statistics, pairwise merge) -> Abuf
//               epi2(j): + bias, GELU -> Hbuf[j & 1]
//               epi3: + bias -> X (TMA store), LayerNorm_next -> XN (TMA store)
#pragma once

#include "common.cuh"
#include "gemm2.cuh"

namespace ddc {

// warps 0,3 TMA, warp 1 MMA (pair leader), warp 2 TMEM, then NG epilogue groups of four warps.  NG = 2 (384 threads) is
// the default.  NG = 4 (640 threads) exists as a measured alternative: 640 threads start at 96 registers each and
// setmaxnreg can only move registers inside what the CTA got at launch (640 x 96 = 61440; asking for more blocks
// forever), so the four control warps shrink to 64 and the epilogue warps grow to 104.  It is slower (see the launcher).
__host__ __device__ constexpr int bt_threads(int ng) { return 128 + 128 * ng; }
constexpr int BT_TAIL_BYTES = 6144;          // barriers, TMEM slot, row-statistics exchange [2][4][128], residual barriers
constexpr int BT_BN = 128;                   // N tile of every GEMM in here
constexpr int BT_HSLOT_BYTES = 2 * G2_A_KB_BYTES;   // one hidden block: 128 rows x 128 cols, two K blocks

struct BlockTailParams {
    int M, C, stages, num_tiles;
    int gk;                  // K blocks (of 64) per W ring stage: 1 or 2.  One handshake moves gk * 8 KB per CTA.
    const float* b_proj;
    const float* b_fc1;      // with LayerNorm2's beta folded in (and Wfc1 carries its gamma)
    const float* b_fc2;
    float ln_eps;
    long long* dbg;          // probes only: [role 2][64] SM clock stamps of CTA 0, [128..] wait accounting
    int flags;               // tuning: bit 0 = L2-prefetch the next tile's O / residual rows (measured: no gain);
                             //         bit 1 = L2 eviction hints (weights evict_last; O, residual, X evict_first)
    int stagger;             // SM cycles by which CTA pairs with a spare round start late (0 = off), see the kernel
};

// Timeline stamps exist only in probe builds (scripts/probes define DDC_PROBE_TIMING): in the product kernel they would
// cost registers the 16-warp epilogue does not have.
__device__ __forceinline__ void bt_stamp(const BlockTailParams& p, int role, int idx) {
#ifdef DDC_PROBE_TIMING
    if (p.dbg != nullptr && blockIdx.x == 0 && idx < 64) p.dbg[role * 64 + idx] = clock64();
#endif
}

inline size_t block_tail_smem_bytes(int C, int stages, int gk) {
    return 1024 + static_cast<size_t>(C / 64) * G2_A_KB_BYTES + 2 * BT_HSLOT_BYTES + static_cast<size_t>(stages) * gk * 64 * 128 +
           ((static_cast<size_t>(3) * C * 4 + 127) & ~static_cast<size_t>(127)) + BT_TAIL_BYTES;
}

template <bool F16, int NG>
__global__ void __launch_bounds__(bt_threads(NG), 1)
block_tail_kernel(const __grid_constant__ CUtensorMap tmO,      // O      [Mpad, C] 16 bit, box 64 x 128 (A operand)
                  const __grid_constant__ CUtensorMap tmWp,     // Wproj  [C, C]    16 bit, box 64 x 64
                  const __grid_constant__ CUtensorMap tmW1,     // Wfc1
                  const __grid_constant__ CUtensorMap tmW2,     // Wfc2
                  const __grid_constant__ CUtensorMap tmXld,    // X      [Mpad, C] fp32,  box 32 x 128 (residual)
                  const __grid_constant__ CUtensorMap tmXst,    // X      [M, C]    fp32,  box 32 x 128 (store)
                  const __grid_constant__ CUtensorMap tmXN,     // XN     [M, C]    16 bit, box 64 x 128 (store)
                  const BlockTailParams p) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

    const int C = p.C;
    const int KB = C / 64;                       // K blocks of every GEMM (K == C)
    const int NB = C / BT_BN;                    // N tiles == hidden blocks
    constexpr uint32_t kWBytes = 64 * 128;       // this CTA's half (64 rows) of a 128 x 64 weight tile
    const uint32_t gk = static_cast<uint32_t>(p.gk);
    const uint32_t kStageBytes = gk * kWBytes;   // a ring stage: gk consecutive K blocks of one N tile
    uint8_t* abuf = smem;
    uint8_t* hbuf = abuf + static_cast<size_t>(KB) * G2_A_KB_BYTES;
    uint8_t* ring = hbuf + 2 * BT_HSLOT_BYTES;
    float* sp = reinterpret_cast<float*>(ring + static_cast<size_t>(p.stages) * kStageBytes);
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sp) +
                                                 ((static_cast<size_t>(3) * C * 4 + 127) & ~static_cast<size_t>(127)));
    uint64_t* full_bar = bars;            // [8]  W ring: TMA -> MMA            (leader's copy collects both CTAs' bytes)
    uint64_t* empty_bar = bars + 8;       // [8]  W ring: MMA -> TMA            (pair commit)
    uint64_t* afull_bar = bars + 16;      // [8]  O K block landed              (leader)
    uint64_t* aempty_bar = bars + 24;     // [1]  Abuf no longer read by fc1    (pair commit)
    uint64_t* proj_bar = bars + 25;       // [4]  proj N tile complete          (pair commit)
    uint64_t* ln2_bar = bars + 56;        // [8]  LayerNorm2 K block in Abuf    (leader, 4 warps x 2 CTAs)
    uint64_t* t1full_bar = bars + 30;     // [1]  fc1 accumulator complete      (pair commit)
    uint64_t* t1free_bar = bars + 31;     // [1]  fc1 accumulator drained       (leader, 16 arrivals)
    uint64_t* hready_bar = bars + 32;     // [2]  hidden block written          (leader, 16 arrivals)
    uint64_t* hfree_bar = bars + 34;      // [2]  hidden block consumed by fc2  (pair commit)
    uint64_t* final_bar = bars + 36;      // [1]  row block holds X1 + H Wfc2^T (pair commit)
    uint64_t* tfree_bar = bars + 42;      // [4]  N tile of the row block drained by epi3 (leader, 16 arrivals)
    uint64_t* res_bar = bars + 576;       // [4]  residual chunk landed (group x slot), behind the statistics area
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 48);
    float* red = reinterpret_cast<float*>(bars + 64);     // [2 stats][NG groups][128 rows]

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    const bool leader = crank == 0;
    const int first_tile = static_cast<int>(blockIdx.x);       // the pair takes tiles (2c, 2c+1), (2c + grid, ...), ...
    const int tile_step = gridDim.x;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmO); tma_prefetch_desc(&tmWp); tma_prefetch_desc(&tmW1); tma_prefetch_desc(&tmW2);
        tma_prefetch_desc(&tmXld); tma_prefetch_desc(&tmXst); tma_prefetch_desc(&tmXN);
        for (int i = 0; i < 8; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
            mbar_init(&afull_bar[i], 1);
        }
        mbar_init(aempty_bar, 1);
        for (int i = 0; i < 4; ++i) mbar_init(&proj_bar[i], 1);
        for (int i = 0; i < 8; ++i) mbar_init(&ln2_bar[i], 8);
        mbar_init(t1full_bar, 1);
        mbar_init(t1free_bar, 8 * NG);              // every epilogue warp of both CTAs
        mbar_init(&hready_bar[0], 8 * NG);
        mbar_init(&hready_bar[1], 8 * NG);
        mbar_init(&hfree_bar[0], 1);
        mbar_init(&hfree_bar[1], 1);
        mbar_init(final_bar, 1);
        for (int i = 0; i < 4; ++i) mbar_init(&tfree_bar[i], 16);
        for (int i = 0; i < 4; ++i) mbar_init(&res_bar[i], 1);
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc<2>(tmem_slot, 512);
        tmem_relinquish<2>();
    }
    if (warp >= 4) {
        const float* src[3] = {p.b_proj, p.b_fc1, p.b_fc2};
        for (int i = threadIdx.x - 128; i < 3 * C; i += 128 * NG) sp[i] = src[i / C][i % C];
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();        // set-up above touched only weights; O and X come from earlier kernels
    // Phase stagger.  Every pair walks the same phases per tile -- HBM reads (O, residual), an MLP stretch that moves no
    // activation bytes, HBM writes (X, XN) -- and with all pairs in lockstep the memory system sees bursts it cannot
    // absorb followed by idle time.  When the tile count does not divide evenly, the pairs that run one round fewer
    // have a whole tile of slack: they start `stagger` cycles late (a good part of a tile), so their bursts fall into the
    // other pairs' quiet stretch.  Costs nothing: they still finish before the pairs with the extra round.
    if (p.stagger > 0) {
        const int my_rounds = (p.num_tiles - first_tile + tile_step - 1) / tile_step;
        const int max_rounds = (p.num_tiles + tile_step - 1) / tile_step;
        if (my_rounds < max_rounds) {
            const long long t0 = clock64();
            while (clock64() - t0 < p.stagger) __nanosleep(256);
        }
    }
    if (NG == 4) {     // warp-group aligned: warps 0-3 shrink, the sixteen epilogue warps grow (see bt_threads)
        if (warp < 4) asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
        else asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
    }
    const float* s_bproj = sp;
    const float* s_bfc1 = sp + C;
    const float* s_bfc2 = sp + 2 * C;

    if (warp == 0 || warp == 3) {
        // ================================================================= TMA producers (alternate loads)
        if (lane == 0) {
            const uint32_t me = warp == 0 ? 0u : 1u;
            uint32_t bit = 0, tc = 0;
            const bool hint = (p.flags & 2) != 0;
            const uint64_t pol_w = l2_policy_evict_last(), pol_a = l2_policy_evict_first();
            auto load_w = [&](const CUtensorMap* tm, int kcol, int row0) {
                if ((bit & 1u) == me) {
                    const uint32_t stage = bit % static_cast<uint32_t>(p.stages);
                    const uint32_t phase = (bit / static_cast<uint32_t>(p.stages)) & 1u;
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * kStageBytes);
                    for (uint32_t i = 0; i < gk; ++i) {
                        uint8_t* dst = ring + static_cast<size_t>(stage) * kStageBytes + i * kWBytes;
                        if (hint) tma_load_2d_pair_hint(dst, tm, &full_bar[stage], kcol + static_cast<int>(i) * 64,
                                                        row0 + static_cast<int>(crank) * 64, pol_w);
                        else tma_load_2d_pair(dst, tm, &full_bar[stage], kcol + static_cast<int>(i) * 64,
                                              row0 + static_cast<int>(crank) * 64);
                    }
                }
                ++bit;
            };
            for (int tile = first_tile; tile < p.num_tiles; tile += tile_step, ++tc) {
                const int m0 = tile * G2_BM;
                mbar_wait(aempty_bar, (tc & 1) ^ 1);                 // fc1 of the previous tile has read Abuf
                for (int kb = static_cast<int>(me); kb < KB; kb += 2) {
                    if (leader) mbar_arrive_expect_tx(&afull_bar[kb], 2 * G2_A_KB_BYTES);
                    if (hint) tma_load_2d_pair_hint(abuf + static_cast<size_t>(kb) * G2_A_KB_BYTES, &tmO, &afull_bar[kb], kb * 64, m0, pol_a);
                    else tma_load_2d_pair(abuf + static_cast<size_t>(kb) * G2_A_KB_BYTES, &tmO, &afull_bar[kb], kb * 64, m0);
                }
                for (int n = 0; n < NB; ++n)
                    for (int kb = 0; kb < KB; kb += p.gk) load_w(&tmWp, kb * 64, n * BT_BN);
                // MMA order of the MLP: fc1(0) | fc1(j+1), fc2 partial(j) ...  (fc1 of the next hidden block runs
                // while the epilogue applies GELU to the current one)
                for (int kb = 0; kb < KB; kb += p.gk) load_w(&tmW1, kb * 64, 0);
                // The tile's own HBM reads (O, residual) are over by now and its stores have not begun: pull the next
                // tile's O rows into L2 in this quiet window rather than on top of the current tile's loads.
                if (me == 0 && tile + tile_step < p.num_tiles && (p.flags & 1))
                    for (int kb = 0; kb < KB; ++kb) tma_prefetch_l2_2d(&tmO, kb * 64, (tile + tile_step) * G2_BM);
                for (int j = 0; j < NB; ++j) {
                    if (j + 1 < NB)
                        for (int kb = 0; kb < KB; kb += p.gk) load_w(&tmW1, kb * 64, (j + 1) * BT_BN);
                    for (int n = 0; n < NB; ++n)
                        for (int kk = 0; kk < 2; kk += p.gk) load_w(&tmW2, (2 * j + kk) * 64, n * BT_BN);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1 && leader) {
        // ================================================================= MMA issuer (pair leader)
        const uint32_t idesc = umma_idesc_16bit(256, BT_BN, F16);
        const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint64_t ad0 = umma_desc_sw128(smem_u32(abuf));
        const uint64_t hd0 = umma_desc_sw128(smem_u32(hbuf));
        const uint64_t bd0 = umma_desc_sw128(smem_u32(ring));
        const uint32_t kStageUnits = kStageBytes >> 4;
        constexpr uint32_t kWUnits = kWBytes >> 4;
        constexpr uint32_t kKbUnits = G2_A_KB_BYTES >> 4;
        uint32_t bit = 0, tc = 0, g = 0;          // g: running hidden-block index (H slot = g & 1)
        auto mma_step = [&](uint32_t d_tmem, uint64_t adesc, uint32_t accumulate) {
            const uint32_t stage = bit % static_cast<uint32_t>(p.stages);
            const uint32_t phase = (bit / static_cast<uint32_t>(p.stages)) & 1u;
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            if (elect_one()) {
                umma_kblock<2>(d_tmem, adesc, bd0 + static_cast<uint64_t>(stage * kStageUnits), idesc, accumulate);
                if (gk == 2)
                    umma_kblock<2>(d_tmem, adesc + kKbUnits, bd0 + static_cast<uint64_t>(stage * kStageUnits + kWUnits), idesc, 1u);
                umma_commit_pair(&empty_bar[stage]);
            }
            __syncwarp();
            ++bit;
        };
        for (int tile = first_tile; tile < p.num_tiles; tile += tile_step, ++tc) {
            if (lane == 0) bt_stamp(p, 0, tc * 16 + 0);
            // ---- proj: row block <- O * Wproj^T.  N tile n may start as soon as epi3 of the previous tile has
            // drained columns [128 n, 128 n + 128).
            for (int n = 0; n < NB; ++n) {
                mbar_wait(&tfree_bar[n], (tc & 1) ^ 1);
                tc_fence_after();
                for (int kb = 0; kb < KB; kb += p.gk) {
                    if (n == 0) {
                        mbar_wait(&afull_bar[kb], tc & 1);
                        if (gk == 2) mbar_wait(&afull_bar[kb + 1], tc & 1);
                        tc_fence_after();
                    }
                    mma_step(tb + static_cast<uint32_t>(n * BT_BN), ad0 + static_cast<uint64_t>(kb * kKbUnits), kb != 0);
                }
                if (elect_one()) umma_commit_pair(&proj_bar[n]);
                __syncwarp();
            }
            if (lane == 0) bt_stamp(p, 0, tc * 16 + 1);
            // ---- MLP: Abuf now receives LayerNorm2(X1) from the epilogue warps of both CTAs, K block by K block
            auto fc1 = [&](int j, uint32_t gj, bool first) {
                mbar_wait(t1free_bar, (gj & 1u) ^ 1u);                // fc1 accumulator read out by epi2(gj - 1)
                tc_fence_after();
                for (int kb = 0; kb < KB; kb += p.gk) {
                    if (first) {
                        mbar_wait(&ln2_bar[kb], tc & 1);
                        if (gk == 2) mbar_wait(&ln2_bar[kb + 1], tc & 1);
                        tc_fence_after();
                    }
                    mma_step(tb + static_cast<uint32_t>(C), ad0 + static_cast<uint64_t>(kb * kKbUnits), kb != 0);
                }
                if (elect_one()) {
                    umma_commit_pair(t1full_bar);
                    if (j == NB - 1) umma_commit_pair(aempty_bar);  // Abuf may take the next tile's O
                }
                __syncwarp();
                if (lane == 0) bt_stamp(p, 0, tc * 16 + 3 + 4 * j);
            };
            fc1(0, g, true);
            for (int j = 0; j < NB; ++j, ++g) {
                const uint32_t hs = g & 1u;
                if (j + 1 < NB) fc1(j + 1, g + 1, false);
                mbar_wait(&hready_bar[hs], (g >> 1) & 1u);            // epi2 wrote hidden block g into Hbuf[hs]
                tc_fence_after();
                if (lane == 0) bt_stamp(p, 0, tc * 16 + 4 + 4 * j);
                for (int n = 0; n < NB; ++n)
                    for (int kk = 0; kk < 2; kk += p.gk)
                        mma_step(tb + static_cast<uint32_t>(n * BT_BN),
                                 hd0 + static_cast<uint64_t>(hs * (BT_HSLOT_BYTES >> 4) + kk * kKbUnits), 1u);
                if (lane == 0) bt_stamp(p, 0, tc * 16 + 5 + 4 * j);
                if (elect_one()) {
                    umma_commit_pair(&hfree_bar[hs]);
                    if (j == NB - 1) umma_commit_pair(final_bar);
                }
                __syncwarp();
            }
        }
    } else if (warp >= 4) {
        // ================================================================= epilogue (NG groups of four warps)
        // Warp (q, hf): TMEM lane quarter q (rows 32q .. 32q+31 of the tile, thread = row), column group hf.  Groups take
        // 32-column chunks (fp32 passes) or 64-column chunks (16-bit passes) round-robin.  Hbuf doubles as staging:
        // SL = 4 / NG slots of 16 KB per group.
        constexpr int SL = 4 / NG;
        const int q = warp & 3;
        const int hf = (warp - 4) >> 2;
        const bool gleader = q == 0 && lane == 0;
        const int gbar = 1 + hf;                      // named barrier of the group (128 threads), ids 1 .. NG
        const int pbar = 8 + q;                       // named barrier of the NG warps that share lane quarter q
        const int trow = q * 32 + lane;
        const uint32_t lane_taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        uint8_t* gslots = hbuf + static_cast<size_t>(hf) * SL * G2_GSLOT_BYTES;
        uint64_t* my_res = res_bar + hf * SL;
        const int nchunk = C / 32;
        const int nper = (nchunk - hf + NG - 1) / NG;          // 32-column chunks of this group
        const float inv_c = 1.0f / static_cast<float>(C);
        float* red_mean = red + hf * 128 + trow;                // [2 stats][NG groups][128 rows]
        float* red_m2 = red + NG * 128 + hf * 128 + trow;
        uint32_t tc = 0, g = 0, ri = 0, res_phase = 0;
        const bool hint = (p.flags & 2) != 0;
        const uint64_t pol_x = l2_policy_evict_first();

        // LayerNorm statistics in one pass over the data: each group accumulates sum and sum of squares of its part of
        // the row around a shift K (the first value it sees, so nothing cancels), turns them into (mean, M2) of that
        // part, and the NG equal-sized parts are merged exactly (Chan et al.): mean = avg(m_i),
        // M2 = sum(M2_i) + n_part * sum((m_i - mean)^2).
        const float part_n = static_cast<float>(C) / static_cast<float>(NG);
        auto row_stats = [&](float shift, float s1, float s2, float& mean, float& rstd) {
            const float mg = shift + s1 / part_n;
            const float m2g = s2 - s1 * s1 / part_n;
            *red_mean = mg;
            *red_m2 = m2g;
            tc_fence_before();                  // the partner groups later read TMEM columns this warp stored
            named_bar_sync(pbar, 32 * NG);
            tc_fence_after();
            float ms[NG], msum = 0.f, m2 = 0.f;
#pragma unroll
            for (int i = 0; i < NG; ++i) {
                ms[i] = red[i * 128 + trow];
                msum += ms[i];
                m2 += red[NG * 128 + i * 128 + trow];
            }
            mean = msum * (1.0f / NG);
#pragma unroll
            for (int i = 0; i < NG; ++i) m2 = fmaf(part_n * (ms[i] - mean), ms[i] - mean, m2);
            rstd = rsqrtf(m2 * inv_c + p.ln_eps);
            named_bar_sync(pbar, 32 * NG);      // every partner has read the exchange slots before they are reused
        };

        for (int tile = first_tile; tile < p.num_tiles; tile += tile_step, ++tc) {
            const int m0 = tile * G2_BM;

            // ------------------------------------------------------------ epi1: X1 = acc + b_proj + X -> TMEM
            if (gleader) {
                bulk_wait_read<0>();               // the previous tile's X / XN stores have read the staging slots
                for (int k = 0; k < SL && k < nper; ++k) {               // all of the group's slots in flight
                    mbar_arrive_expect_tx(&my_res[k], G2_GSLOT_BYTES);
                    if (hint) tma_load_2d_hint(gslots + k * G2_GSLOT_BYTES, &tmXld, &my_res[k], (hf + NG * k) * 32, m0, pol_x);
                    else tma_load_2d(gslots + k * G2_GSLOT_BYTES, &tmXld, &my_res[k], (hf + NG * k) * 32, m0);
                }
            }
            if (warp == 4 && lane == 0) bt_stamp(p, 1, tc * 16 + 0);
            float shift = 0.f, s1 = 0.f, s2 = 0.f;
            int waited = -1;
#ifdef DDC_PROBE_TIMING
            long long res_wait = 0;                // probes only: cycles this warp waited for residual chunks
#endif
            for (int k = 0; k < nper; ++k) {
                const int c = hf + NG * k;
                const uint32_t s = static_cast<uint32_t>(k % SL);
                uint8_t* slot = gslots + s * G2_GSLOT_BYTES;
                const int n0 = c * 32;
                if (n0 / BT_BN != waited) {
                    waited = n0 / BT_BN;
                    mbar_wait(&proj_bar[waited], tc & 1);
                    tc_fence_after();
                }
                float v[32];
                tmem_ld32(lane_taddr + n0, v);
#ifdef DDC_PROBE_TIMING
                const long long w0 = (p.dbg != nullptr) ? clock64() : 0;
#endif
                mbar_wait(&my_res[s], (res_phase >> s) & 1);
#ifdef DDC_PROBE_TIMING
                if (p.dbg != nullptr) res_wait += clock64() - w0;
#endif
                res_phase ^= (1u << s);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float4 r = *reinterpret_cast<const float4*>(slot + slot_off(trow, j));
                    const float4 b = *reinterpret_cast<const float4*>(s_bproj + n0 + 4 * j);
                    v[4 * j] += r.x + b.x; v[4 * j + 1] += r.y + b.y;
                    v[4 * j + 2] += r.z + b.z; v[4 * j + 3] += r.w + b.w;
                }
                if (k == 0) shift = v[0];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float d = v[i] - shift;
                    s1 += d;
                    s2 = fmaf(d, d, s2);
                }
                tmem_st32(lane_taddr + n0, v);
                if (k + SL < nper) {
                    named_bar_sync(gbar, 128);     // every warp of the group has read the slot before it is refilled
                    if (gleader) {
                        mbar_arrive_expect_tx(&my_res[s], G2_GSLOT_BYTES);
                        if (hint) tma_load_2d_hint(slot, &tmXld, &my_res[s], (c + NG * SL) * 32, m0, pol_x);
                        else tma_load_2d(slot, &tmXld, &my_res[s], (c + NG * SL) * 32, m0);
                    }
                }
            }
            if (warp == 4 && lane == 0) {
                bt_stamp(p, 1, tc * 16 + 1);
#ifdef DDC_PROBE_TIMING
                if (p.dbg != nullptr && blockIdx.x == 0 && tc < 4) p.dbg[128 + tc] = res_wait;
#endif
            }
            float mean, rstd;
            row_stats(shift, s1, s2, mean, rstd);
            float nmr = -mean * rstd;              // (v - mean) * rstd  ==  fma(v, rstd, nmr)
            // LayerNorm2 -> Abuf (K-major, 128-byte swizzle: 64 columns = one K block).  All proj MMAs that read
            // the O tile in Abuf completed before proj_bar[NB-1] fired.
            for (int c = hf; c < C / 64; c += NG) {
                uint8_t* kblk = abuf + static_cast<size_t>(c) * G2_A_KB_BYTES;
                float v[64];                       // the K block's 64 columns: both loads in flight before the one wait
                tmem_ld32_nowait(lane_taddr + c * 64, v);
                tmem_ld32_nowait(lane_taddr + c * 64 + 32, v + 32);
                tmem_ld_wait64(v);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    uint4 o;
                    o.x = pack2<F16>(fmaf(v[8 * j], rstd, nmr), fmaf(v[8 * j + 1], rstd, nmr));
                    o.y = pack2<F16>(fmaf(v[8 * j + 2], rstd, nmr), fmaf(v[8 * j + 3], rstd, nmr));
                    o.z = pack2<F16>(fmaf(v[8 * j + 4], rstd, nmr), fmaf(v[8 * j + 5], rstd, nmr));
                    o.w = pack2<F16>(fmaf(v[8 * j + 6], rstd, nmr), fmaf(v[8 * j + 7], rstd, nmr));
                    *reinterpret_cast<uint4*>(kblk + slot_off(trow, j)) = o;
                }
                fence_proxy_async_smem();          // generic writes -> visible to the tensor core's operand reads
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_leader(&ln2_bar[c]);    // fc1 can consume this K block right away
            }
            if (warp == 4 && lane == 0) bt_stamp(p, 1, tc * 16 + 2);

            // ------------------------------------------------------------ epi2(j): H_j = gelu(acc + b_fc1) -> Hbuf
            constexpr int HW = BT_BN / NG;         // hidden columns of a block per group: 64 (two groups) or 32 (four)
            for (int j = 0; j < NB; ++j, ++g) {
                const uint32_t hs = g & 1u;
                mbar_wait(t1full_bar, g & 1u);
                tc_fence_after();
                if (warp == 4 && lane == 0) bt_stamp(p, 1, tc * 16 + 3 + 2 * j);
                if (g >= 2) mbar_wait(&hfree_bar[hs], ((g >> 1) & 1u) ^ 1u);   // fc2 partial(g-2) has read this slot
                // read this thread's accumulator columns, then hand the accumulator back at once: fc1 of the next
                // hidden block overlaps the GELU below
                float vv[HW / 32][32];
                if constexpr (HW == 64) {          // both loads in flight before the one wait
                    tmem_ld32_nowait(lane_taddr + static_cast<uint32_t>(C) + hf * HW, vv[0]);
                    tmem_ld32_nowait(lane_taddr + static_cast<uint32_t>(C) + hf * HW + 32, vv[1]);
                    tmem_ld_wait64(&vv[0][0]);
                } else {
#pragma unroll
                    for (int h = 0; h < HW / 32; ++h)
                        tmem_ld32(lane_taddr + static_cast<uint32_t>(C) + hf * HW + 32 * h, vv[h]);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_leader(t1free_bar);
#pragma unroll
                for (int h = 0; h < HW / 32; ++h) {
                    float (&v)[32] = vv[h];
                    const int c0 = hf * HW + 32 * h;                  // column inside the hidden block
                    uint8_t* kblk = hbuf + static_cast<size_t>(hs) * BT_HSLOT_BYTES + static_cast<size_t>(c0 >> 6) * G2_A_KB_BYTES;
                    const int jb = (c0 & 63) >> 3;                    // first 16-byte chunk inside the 128-byte row
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        const float4 b = *reinterpret_cast<const float4*>(s_bfc1 + j * BT_BN + c0 + i);
                        v[i] = gelu_erf(v[i] + b.x); v[i + 1] = gelu_erf(v[i + 1] + b.y);
                        v[i + 2] = gelu_erf(v[i + 2] + b.z); v[i + 3] = gelu_erf(v[i + 3] + b.w);
                    }
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        uint4 o;
                        o.x = pack2<F16>(v[8 * jj], v[8 * jj + 1]);
                        o.y = pack2<F16>(v[8 * jj + 2], v[8 * jj + 3]);
                        o.z = pack2<F16>(v[8 * jj + 4], v[8 * jj + 5]);
                        o.w = pack2<F16>(v[8 * jj + 6], v[8 * jj + 7]);
                        *reinterpret_cast<uint4*>(kblk + slot_off(trow, jb + jj)) = o;
                    }
                }
                fence_proxy_async_smem();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_leader(&hready_bar[hs]);
                if (warp == 4 && lane == 0) bt_stamp(p, 1, tc * 16 + 4 + 2 * j);
            }
            // the MLP phase moves no activation bytes: (measured) an L2 prefetch of the next tile's residual here is no gain
            if (gleader && tile + tile_step < p.num_tiles && (p.flags & 1))
                for (int c = hf; c < nchunk; c += NG) tma_prefetch_l2_2d(&tmXld, c * 32, (tile + tile_step) * G2_BM);

            // ------------------------------------------------------------ epi3: X' = TMEM + b_fc2 -> X, LN_next -> XN
            mbar_wait(final_bar, tc & 1);
            tc_fence_after();
            if (warp == 4 && lane == 0) bt_stamp(p, 1, tc * 16 + 9);
            s1 = 0.f; s2 = 0.f;
            for (int k = 0; k < nper; ++k, ++ri) {
                const int c = hf + NG * k;
                uint8_t* slot = gslots + (ri % SL) * G2_GSLOT_BYTES;
                const int n0 = c * 32;
                float v[32];
                tmem_ld32(lane_taddr + n0, v);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float4 b = *reinterpret_cast<const float4*>(s_bfc2 + n0 + 4 * j);
                    v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
                }
                if (k == 0) shift = v[0];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float d = v[i] - shift;
                    s1 += d;
                    s2 = fmaf(d, d, s2);
                }
                if (gleader) bulk_wait_read<SL - 1>();   // the store that last used this slot has read it
                named_bar_sync(gbar, 128);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<float4*>(slot + slot_off(trow, j)) =
                        make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                fence_proxy_async_smem();
                named_bar_sync(gbar, 128);
                if (gleader && m0 < p.M) {
                    if (hint) tma_store_2d_hint(&tmXst, slot, n0, m0, pol_x);
                    else tma_store_2d(&tmXst, slot, n0, m0);
                    bulk_commit();
                }
            }
            if (warp == 4 && lane == 0) bt_stamp(p, 1, tc * 16 + 10);
            row_stats(shift, s1, s2, mean, rstd);
            nmr = -mean * rstd;
            for (int c = hf; c < C / 64; c += NG, ++ri) {
                uint8_t* slot = gslots + (ri % SL) * G2_GSLOT_BYTES;
                if (gleader) bulk_wait_read<SL - 1>();   // the store that last used this slot has read it
                named_bar_sync(gbar, 128);
                {
                    float v[64];                   // both loads in flight before the one wait
                    const int n0 = c * 64;
                    tmem_ld32_nowait(lane_taddr + n0, v);
                    tmem_ld32_nowait(lane_taddr + n0 + 32, v + 32);
                    tmem_ld_wait64(v);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 a0 = *reinterpret_cast<const float4*>(s_bfc2 + n0 + 8 * j);
                        const float4 a1 = *reinterpret_cast<const float4*>(s_bfc2 + n0 + 8 * j + 4);
                        uint4 o;
                        o.x = pack2<F16>(fmaf(v[8 * j] + a0.x, rstd, nmr), fmaf(v[8 * j + 1] + a0.y, rstd, nmr));
                        o.y = pack2<F16>(fmaf(v[8 * j + 2] + a0.z, rstd, nmr), fmaf(v[8 * j + 3] + a0.w, rstd, nmr));
                        o.z = pack2<F16>(fmaf(v[8 * j + 4] + a1.x, rstd, nmr), fmaf(v[8 * j + 5] + a1.y, rstd, nmr));
                        o.w = pack2<F16>(fmaf(v[8 * j + 6] + a1.z, rstd, nmr), fmaf(v[8 * j + 7] + a1.w, rstd, nmr));
                        *reinterpret_cast<uint4*>(slot + slot_off(trow, j)) = o;
                    }
                }
                // the two 64-column halves of N tile c / 2 belong to two groups: once both are read out, the next tile's
                // proj may overwrite those TMEM columns while the remaining N tiles are still being normalised
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_leader(&tfree_bar[c >> 1]);
                fence_proxy_async_smem();
                named_bar_sync(gbar, 128);
                if (gleader && m0 < p.M) {
                    tma_store_2d(&tmXN, slot, c * 64, m0);
                    bulk_commit();
                }
            }
            if (warp == 4 && lane == 0) bt_stamp(p, 1, tc * 16 + 11);
        }
        if (gleader) bulk_wait_all();
        __syncwarp();
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 2) tmem_dealloc<2>(tmem_base, 512);
}

// Pack-time fold of a LayerNorm's affine part into the Linear that consumes it:  (n * gamma + beta) W^T + b  =
// n (W o gamma)^T + (b + W beta), n the plainly normalised activations.  w_out = 16-bit(gamma o W) [N][k_pad], zero padded;
// b_out = b + W beta (fp32, from the unrounded W).  One block per output row.
template <bool F16>
__global__ void __launch_bounds__(128) fold_ln_linear_kernel(const float* __restrict__ w, const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, const float* __restrict__ b,
                                                             bf16* __restrict__ w_out, float* __restrict__ b_out, int K, int k_pad) {
    const int o = blockIdx.x;
    float acc = 0.f;
    for (int k = threadIdx.x; k < k_pad; k += blockDim.x) {
        const float wv = k < K ? w[static_cast<size_t>(o) * K + k] : 0.f;
        const float gv = k < K ? gamma[k] : 0.f;
        reinterpret_cast<uint16_t*>(w_out)[static_cast<size_t>(o) * k_pad + k] = static_cast<uint16_t>(pack2<F16>(gv * wv, 0.f) & 0xFFFFu);
        if (k < K) acc = fmaf(beta[k], wv, acc);
    }
    __shared__ float red[4];
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) b_out[o] = b[o] + ((red[0] + red[1]) + (red[2] + red[3]));
}

__global__ void fill_f32_kernel(float* __restrict__ dst, float v, size_t n) {
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x) dst[i] = v;
}

}  // namespace ddc

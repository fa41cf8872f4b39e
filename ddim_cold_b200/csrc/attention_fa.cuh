// Attention for long token sequences on the tcgen05 tensor cores (ViT.py:107-114): S2 (257 tokens, head_dim 64) and
// S3 (626 tokens, head_dim 32), where the L^2 work is large enough for 128-row UMMA tiles to pay.
//
// Work item = (image, 64 qkv columns) = one head of 64 or a pair of heads of 32.  Everything arrives by TMA as
// [128 rows x 128 B] boxes in the 128-byte swizzle: Q tiles, and K / V in chunks of 128 keys through two rings.
// Per (query tile, key chunk):
//     S   = Q K_c^T           tcgen05.mma, both operands K-major, fp32 scores in TMEM (128 columns)
//     P   = exp2(S*c - m)     four softmax warps per tile, one query row per thread straight from TMEM: running max m
//                             and sum l in registers, P written as 16-bit rows of a K-major shared-memory tile
//     O  += P V_c             tcgen05.mma with V as it lies in memory ([key][dim], i.e. an MN-major B operand); O lives
//                             in TMEM across the chunks and is rescaled in place only when a row maximum jumps
// Two tiles w = 0, 1 run side by side on one tensor pipe: for head_dim 64 two consecutive query tiles of the same
// head, for head_dim 32 the two heads of the pair on the same query tile (the head is a 64-byte offset inside the
// 128-byte rows).  Each tile has two softmax warpgroups, one per 64-key half of the chunk.  The exponentials are the
// bound (16 MUFU results per clock per SM); the tensor pipe idles most of the time.
//
// A token count of the form 128 n + r with small r (S2: 257 = 2 * 128 + 1) would waste a whole 128-row tile on r
// rows: those rows are left to attention_rows_kernel, one warp per (image, head, row) on the CUDA cores.
//
// TMEM (512 columns allocated): warpgroup w uses S at [224 w, +128), O at [224 w + 128, +D) and the row sums (P times
// a block of ones, 16 identical columns) right after O.
#pragma once

#include "attention.cuh"
#include "gemm2.cuh"

namespace ddc {

constexpr int FA_THREADS = 640;            // warp 0 TMEM + TMA, warps 1-2 MMA (one per tile), warp 3 spare, warps 4-19 softmax
constexpr int FA_KCH = 128;                // keys per chunk
constexpr int FA_TILE = 128 * 128;         // bytes of a [128 rows x 128 B] tile
// Shared memory = 13 tiles of 16 KB (two Q buffers -- head_dim 64: two query tiles each -- a K ring, a V ring, P) plus a
// 2 KB block of ones.
// K and V have rings of their own because their lifetimes differ: a K chunk is done with as soon as Q K^T has run, a
// V chunk only after the softmax of that chunk.  One combined stage per chunk left a single load in flight ahead of the
// MMAs, and the whole kernel ran at the latency of a TMA load per chunk (measured: 4.3k cycles per chunk with the
// softmax arithmetic removed entirely).
template <int D> struct FaCfg {
    static constexpr int QT = D == 64 ? 2 : 1;         // query tiles per Q buffer
    static constexpr int NK = D == 64 ? 2 : 3;         // K ring stages (a K chunk lives from its load to the end of Q K^T)
    static constexpr int NV = D == 64 ? 3 : 4;         // V ring stages (a V chunk lives until the softmax of its chunk is done)
};
constexpr int FA_TMEM_W = 224;             // TMEM columns per warpgroup: S 128, O 64, row sums 32
constexpr float FA_MAX_SLACK = 8.f;        // see the softmax warps
constexpr int FA_SIDE_MAX = 8;             // remainder rows handed to the CUDA-core kernel
constexpr int FA_SMEM_BYTES = 1024 + 13 * FA_TILE + 2048 + 2048 + 512;

struct FaParams {
    int L, C, n_items, n_hs;       // tokens, embed_dim, work items, head slots (64 columns) per image
    int n_qi;                      // query iterations per (image, head slot): item = (image * n_hs + slot) * n_qi + iteration
    int n_qt;                      // 128-row query tiles handled by the tensor kernel
    int rows_main;                 // query rows [0, rows_main) of every image belong to those tiles
    int n_kc;                      // key chunks
    float scale_log2e;
    int dbg;                       // probes only (DDC_FA_DBG): 1 = no S read from TMEM, 2 = no exponentials, 4 = no P stores
};

inline void fa_split_rows(int L, int* n_qt, int* rows_main) {
    const int full = L / 128, r = L % 128;
    if (full > 0 && r > 0 && r <= FA_SIDE_MAX) { *n_qt = full; *rows_main = full * 128; }
    else { *n_qt = (L + 127) / 128; *rows_main = L; }
}

// Probe builds only (-DDDC_FA_TRACE): SM-clock stamps of CTA 0 -- [role][event index][8 stamps]
#ifdef DDC_FA_TRACE
__device__ long long g_fa_trace[3 * 64 * 8];
#define FA_STAMP(role, idx, k) do { if (blockIdx.x == 0 && lane == 0 && (idx) < 64) g_fa_trace[((role) * 64 + (idx)) * 8 + (k)] = clock64(); } while (0)
#else
#define FA_STAMP(role, idx, k)
#endif

template <int D, bool F16>
__global__ void __launch_bounds__(FA_THREADS, 1)
attention_fa_kernel(const __grid_constant__ CUtensorMap tm,       // qkv [rows, 3C] 16 bit, box 64 cols x 128 rows
                    bf16* __restrict__ out, const FaParams p) {
    extern __shared__ uint8_t fa_raw[];
    pdl_launch_dependents();
    const uint32_t ra = smem_u32(fa_raw);
    uint8_t* sm = fa_raw + (((ra + 1023u) & ~1023u) - ra);
    constexpr int QT = FaCfg<D>::QT, NK = FaCfg<D>::NK, NV = FaCfg<D>::NV;
    uint8_t* Qs = sm;                                   // 2 buffers x QT tiles
    uint8_t* Ks = Qs + 2 * QT * FA_TILE;                // NK tiles
    uint8_t* Vs = Ks + NK * FA_TILE;                    // NV tiles
    uint8_t* Ps = Vs + NV * FA_TILE;                    // per warpgroup: 2 K blocks of 64 keys
    uint8_t* Ones = Ps + 2 * 2 * FA_TILE;               // [16 rows x 128 B] of 1.0: the B operand of the row-sum MMAs
    float* xchg = reinterpret_cast<float*>(Ones + 2048);  // [tile][key half][128 rows] partial row maxima
    uint64_t* bars = reinterpret_cast<uint64_t*>(Ones + 4096);
    uint64_t* q_full = bars;             // [2]
    uint64_t* q_empty = bars + 2;        // [2]
    uint64_t* k_full = bars + 4;         // [4]
    uint64_t* k_empty = bars + 8;        // [4]
    uint64_t* v_full = bars + 12;        // [8]
    uint64_t* v_empty = bars + 20;       // [8]
    uint64_t* s_full = bars + 28;        // [2] S of warpgroup w complete (commit)
    uint64_t* s_free = bars + 30;        // [2] warpgroup w has read S out (4 warps)
    uint64_t* p_full = bars + 32;        // [2] P of warpgroup w written, O rescaled if need be (4 warps)
    uint64_t* o_full = bars + 34;        // [2] P V of warpgroup w complete (commit)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 36);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tm);
        for (int i = 0; i < 4; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 2); }        // released by both issuers
        for (int i = 0; i < 8; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 2); }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&q_full[i], 1);
            mbar_init(&q_empty[i], 2);
            mbar_init(&s_full[i], 1);
            mbar_init(&s_free[i], 8);
            mbar_init(&p_full[i], 8);
            mbar_init(&o_full[i], 1);
        }
        fence_barrier_init();
    }
    if (warp == 0) {
        __syncwarp();
        tmem_alloc<1>(tmem_slot, 512);
        tmem_relinquish<1>();
    }
    if (warp >= 4) {
        for (int i = threadIdx.x - 128; i < 2048 / 4; i += 512) reinterpret_cast<uint32_t*>(Ones)[i] = F16 ? 0x3C003C00u : 0x3F803F80u;
        fence_proxy_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                         // qkv comes from the previous kernel
    // 640 threads start with 96 registers each; the control warpgroup hands some over to the softmax threads, which hold
    // 64 scores each (128 x 56 + 512 x 104 = 60416 <= 640 x 96: registers only move inside what the CTA got at launch)
    // (issued at the top of each role's own branch: code reachable from both settings is compiled for the smaller one)

    const int L = p.L, C = p.C;
    // work item = (image, head slot, query iteration): a few divisions per item, nothing per chunk
    auto split = [&](int item, int& img, int& hs, int& qi) {
        qi = item % p.n_qi;
        const int t = item / p.n_qi;
        img = t / p.n_hs;
        hs = t - img * p.n_hs;
    };
    auto tile_of = [&](int qi, int w) { return D == 64 ? 2 * qi + w : qi; };

    if (warp < 4) asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (warp == 0) {
        // ================================================================= TMA producer
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            uint32_t q_it = 0, kv_it = 0;
            for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++q_it) {
                int img, hs, qi;
                split(item, img, hs, qi);
                const int row0 = img * L;
                const uint32_t qb = q_it & 1u;
                mbar_wait(&q_empty[qb], ((q_it >> 1) & 1u) ^ 1u);
                const int nq = (D == 64 && tile_of(qi, 1) < p.n_qt) ? 2 : 1;
                mbar_arrive_expect_tx(&q_full[qb], nq * FA_TILE);
                for (int j = 0; j < nq; ++j)
                    tma_load_2d_hint(Qs + (qb * QT + j) * FA_TILE, &tm, &q_full[qb], hs * 64, row0 + tile_of(qi, j) * 128, pol);
                for (int c = 0; c < p.n_kc; ++c, ++kv_it) {
                    const uint32_t ks = kv_it % NK, kph = (kv_it / NK) & 1u;
                    mbar_wait(&k_empty[ks], kph ^ 1);
                    mbar_arrive_expect_tx(&k_full[ks], FA_TILE);
                    tma_load_2d_hint(Ks + ks * FA_TILE, &tm, &k_full[ks], C + hs * 64, row0 + c * FA_KCH, pol);
                    const uint32_t vs = kv_it % NV, vph = (kv_it / NV) & 1u;
                    mbar_wait(&v_empty[vs], vph ^ 1);
                    mbar_arrive_expect_tx(&v_full[vs], FA_TILE);
                    tma_load_2d_hint(Vs + vs * FA_TILE, &tm, &v_full[vs], 2 * C + hs * 64, row0 + c * FA_KCH, pol);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1 || warp == 2) {
        // ================================================================= MMA issuers: warp 1 for tile 0, warp 2 for tile 1
        // One issuer per tile, each waiting only on its own tile's barriers: a single in-order issuer put eight blocking
        // steps of 250-400 cycles each between "P of this tile is written" and "the next S of this tile is ready"
        // (measured timeline), which was the chunk period.  The shared Q / K / V buffers are released by both (count 2).
        const int w = warp - 1;
        const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0) + static_cast<uint32_t>(w * FA_TMEM_W);
        const uint32_t idesc_o = umma_idesc_16bit(128, D, F16) | (1u << 16);      // B operand (V) is MN-major
        const uint32_t idesc_1 = umma_idesc_16bit(128, 16, F16);                  // row sums: P times 16 columns of ones
        const uint64_t od = umma_desc_sw128(smem_u32(Ones));
        const uint64_t qd0 = umma_desc_sw128(smem_u32(Qs));
        const uint64_t kd0 = umma_desc_sw128(smem_u32(Ks));
        const uint64_t vd0 = umma_desc_sw128(smem_u32(Vs));
        const uint64_t pd = umma_desc_sw128(smem_u32(Ps)) + static_cast<uint64_t>(w * 2 * (FA_TILE >> 4));
        constexpr uint32_t kTileU = FA_TILE >> 4;
        uint32_t q_it = 0, kv_it = 0, sfree_it = 0, pfull_it = 0;
        // P*V of a chunk is issued one step late, after the NEXT chunk's Q K^T -- also across item boundaries, so that
        // the first S of an item is under way while the softmax warps still work on the last chunk of the item before
        bool pend = false, pend_act = false;
        int pend_cc = 0;
        uint32_t pend_it = 0;
        // release a shared buffer: once this issuer's MMAs so far are complete, or at once if its tile has no work
        auto release = [&](uint64_t* bar, bool act) {
            if (elect_one()) {
                if (act) umma_commit(bar);
                else mbar_arrive(bar);
            }
            __syncwarp();
        };
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++q_it) {
            int img, hs, qi;
            split(item, img, hs, qi);
            const bool act = tile_of(qi, w) < p.n_qt;
            const uint32_t qb = q_it & 1u;
            const uint64_t qd = qd0 + static_cast<uint64_t>(qb * QT * kTileU);
            mbar_wait(&q_full[qb], (q_it >> 1) & 1u);
            tc_fence_after();
            auto issue_pv = [&](bool act, int cc, uint32_t it) {          // O_w += P_w(cc) V(cc), row sums += P_w(cc) 1
                mbar_wait(&v_full[it % NV], (it / NV) & 1u);
                if (act) {
                    const int kvalid = L - cc * FA_KCH < FA_KCH ? L - cc * FA_KCH : FA_KCH;
                    const int nks = (kvalid + 15) >> 4;
                    const uint64_t vd = vd0 + static_cast<uint64_t>((it % NV) * kTileU);
                    mbar_wait(&p_full[w], pfull_it & 1);
                    ++pfull_it;
                    tc_fence_after();
                    if (elect_one()) {          // whole K blocks of 64 keys: the softmax warps zero-fill a started block
                        const uint64_t vb = vd + static_cast<uint64_t>(D == 32 ? 4 * w : 0);
                        umma_pv_kblock(tb + 128, tb + 128 + D, pd, vb, od, idesc_o, idesc_1, cc > 0 ? 1u : 0u);
                        if (nks > 4) umma_pv_kblock(tb + 128, tb + 128 + D, pd + kTileU, vb + 4 * (2048 >> 4), od, idesc_o, idesc_1, 1u);
                        umma_commit(&o_full[w]);
                    }
                    __syncwarp();
                }
                release(&v_empty[it % NV], act);
            };
            for (int c = 0; c < p.n_kc; ++c, ++kv_it) {
                FA_STAMP(0, kv_it, 0);
                mbar_wait(&k_full[kv_it % NK], (kv_it / NK) & 1u);
                tc_fence_after();
                FA_STAMP(0, kv_it, 1);
                if (act) {                                      // S_w(c) = Q_w K(c)^T
                    const int kvalid = L - c * FA_KCH < FA_KCH ? L - c * FA_KCH : FA_KCH;
                    const uint32_t idesc_s = umma_idesc_16bit(128, static_cast<uint32_t>((kvalid + 15) & ~15), F16);
                    const uint64_t kd = kd0 + static_cast<uint64_t>((kv_it % NK) * kTileU);
                    mbar_wait(&s_free[w], (sfree_it & 1) ^ 1);  // S of the previous chunk has been read out
                    ++sfree_it;
                    tc_fence_after();
                    if (elect_one()) {
                        if (D == 64) umma_kblock<1>(tb, qd + static_cast<uint64_t>(w * kTileU), kd, idesc_s, 0u);
                        else umma_k32(tb, qd + 4 * w, kd + 4 * w, idesc_s);    // head w of the pair: bytes [64 w, +64) of every row
                        umma_commit(&s_full[w]);
                    }
                    __syncwarp();
                }
                FA_STAMP(0, kv_it, 2);
                release(&k_empty[kv_it % NK], act);
                if (c == p.n_kc - 1) release(&q_empty[qb], act);        // every S of this item is issued: Q may be replaced
                if (pend) issue_pv(pend_act, pend_cc, pend_it);
                pend = true; pend_act = act; pend_cc = c; pend_it = kv_it;
                FA_STAMP(0, kv_it, 3);
            }
            if (item + static_cast<int>(gridDim.x) >= p.n_items) issue_pv(pend_act, pend_cc, pend_it);      // last item: flush
        }
    } else if (warp >= 4) {
        // ================================================================= softmax warpgroups (warps 4-19)
        asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
        // Four warpgroups: tile w = g / 2 (see the head comment), key half hf = g % 2 of every 128-key chunk.  A thread
        // owns one query row and 64 of the chunk's keys: it takes those 64 scores from TMEM in one go, hands the S
        // columns back at once (the pair's last reader releases them, so the next chunk's Q K^T runs under the
        // exponentials) and never reads them again.  The two halves of a row exchange their maxima through shared
        // memory (one named barrier per chunk).  O stays in TMEM and accumulates there over the chunks (online softmax:
        // when a row's maximum grows, half 0 rescales the O rows in place, half 1 the row sums).  Row sums come from
        // the tensor core (P times a block of ones), i.e. from the rounded P that the numerator uses.  The MUFU pipe
        // (one unit per SM quarter, 128 exponentials per row and chunk) is the scarce resource; four softmax warps per
        // quarter keep it fed while their neighbours wait, load, reduce or store.
        const int g = (warp - 4) >> 2;
        const int w = g >> 1, hf = g & 1;
        const int q = warp & 3;                         // a warp reads the TMEM lane quarter (warp id % 4)
        const int trow = q * 32 + lane;
        const uint32_t lane_taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(w * FA_TMEM_W);
        uint8_t* Pw = Ps + static_cast<size_t>(w) * 2 * FA_TILE + static_cast<size_t>(hf) * FA_TILE;     // this half's K block of 64 keys
        float* my_max = xchg + (w * 2 + hf) * 128 + trow;
        const float* other_max = xchg + (w * 2 + (hf ^ 1)) * 128 + trow;
        const float sc = p.scale_log2e;
        uint32_t sfull_it = 0, ofull_it = 0;
        // The exponential phases of the two tiles take turns (named barriers 3 and 4 as a token): run at the same time
        // they would only halve each other's MUFU rate, run alternately each overlaps the other tile's loads, waits,
        // reductions and stores.  Tile 1 hands the token over once up front.
        if (w == 1) asm volatile("bar.arrive 4, 512;" ::: "memory");
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            int img, hs, qi;
            split(item, img, hs, qi);
            const int tile = tile_of(qi, w);
            if (tile >= p.n_qt) continue;
            const bool both = tile_of(qi, 1) < p.n_qt;          // the other tile works on this item too
            const int row = tile * 128 + trow;
            float m = -INFINITY;
            for (int c = 0; c < p.n_kc; ++c) {
                const int kvalid = L - c * FA_KCH < FA_KCH ? L - c * FA_KCH : FA_KCH;
                const int hvalid = kvalid - 64 * hf < 64 ? kvalid - 64 * hf : 64;           // valid keys of this half (may be <= 0)
                FA_STAMP(1 + w, sfull_it, 0);
                mbar_wait(&s_full[w], sfull_it & 1);
                FA_STAMP(1 + w, sfull_it, 1);
                ++sfull_it;
                tc_fence_after();
                float v[64];
                tmem_ld32_nowait(lane_taddr + 64 * hf, v);
                tmem_ld32_nowait(lane_taddr + 64 * hf + 32, v + 32);
                tmem_ld_wait64(v);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_free[w]);         // S is in registers (8 warps): the next chunk's S may be computed
                FA_STAMP(1 + w, sfull_it - 1, 2);
                if (hvalid < 64) {
#pragma unroll
                    for (int i = 0; i < 64; ++i) if (i >= hvalid) v[i] = -INFINITY;
                }
                float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
                for (int i = 0; i < 64; i += 4) {
                    m4[0] = fmaxf(m4[0], v[i]); m4[1] = fmaxf(m4[1], v[i + 1]);
                    m4[2] = fmaxf(m4[2], v[i + 2]); m4[3] = fmaxf(m4[3], v[i + 3]);
                }
                float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
                *my_max = mx;
                named_bar_sync(1 + w, 256);                     // both halves of tile w have published their maxima
                mx = fmaxf(mx, *other_max) * sc;                // scaled (log2) domain; scale > 0; identical in both halves
                float alpha = 1.f;
                // The running maximum only moves when a row's scores outgrow it by more than FA_MAX_SLACK (log2 units): P then
                // stays below 2^FA_MAX_SLACK, exact for the quotient P V / P 1 and well inside 16-bit range, and O is not
                // rescaled for the usual small drift of a row maximum from chunk to chunk (first chunk: m = -inf, always taken)
                if (mx > m + FA_MAX_SLACK) { alpha = ex2_approx(m - mx); m = mx; }
                const bool rescale = c > 0 && __any_sync(0xffffffffu, alpha != 1.f);
                FA_STAMP(1 + w, sfull_it - 1, 3);
                if (c > 0) {                        // P*V of the previous chunk is done: P may be overwritten, O is stable
                    mbar_wait(&o_full[w], ofull_it & 1);
                    ++ofull_it;
                    tc_fence_after();
                    if (rescale) {                  // half 0: the O columns; half 1: the 32 columns holding the row sums
                        const int h0 = hf == 0 ? 0 : D / 32, h1 = hf == 0 ? D / 32 : D / 32 + 1;
                        for (int h = h0; h < h1; ++h) {
                            float ov[32];
                            tmem_ld32(lane_taddr + 128 + 32 * h, ov);
#pragma unroll
                            for (int i = 0; i < 32; ++i) ov[i] *= alpha;
                            tmem_st32(lane_taddr + 128 + 32 * h, ov);
                        }
                    }
                }
                FA_STAMP(1 + w, sfull_it - 1, 4);
                const int nstore = hvalid <= 0 ? 0 : 8;       // a started half is written whole: the P*V MMAs run over whole K blocks
                const float nm = -m;
                if (both && !(p.dbg & 8)) named_bar_sync(w == 0 ? 4 : 3, 512);      // my turn on the MUFU pipe
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (j < nstore) {
                        float e[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) e[i] = ex2_approx(fmaf(v[8 * j + i], sc, nm));      // masked keys: 2^(-inf) = 0
                        uint4 u;
                        u.x = pack2<F16>(e[0], e[1]);
                        u.y = pack2<F16>(e[2], e[3]);
                        u.z = pack2<F16>(e[4], e[5]);
                        u.w = pack2<F16>(e[6], e[7]);
                        *reinterpret_cast<uint4*>(Pw + slot_off(trow, j)) = u;
                    }
                }
                if (both && !(p.dbg & 8)) {
                    if (w == 0) asm volatile("bar.arrive 3, 512;" ::: "memory");
                    else asm volatile("bar.arrive 4, 512;" ::: "memory");
                }
                fence_proxy_async_smem();           // P -> visible to the tensor core
                tc_fence_before();                  // and this warp's rescale precedes the next P*V
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_full[w]);
                FA_STAMP(1 + w, sfull_it - 1, 5);
            }
            mbar_wait(&o_full[w], ofull_it & 1);
            ++ofull_it;
            tc_fence_after();
            float lsum[32];
            tmem_ld32(lane_taddr + 128 + D, lsum);      // every one of the 16 sum columns holds the row sum
            const float inv = __frcp_rn(lsum[0]);
            // half hf writes output columns [hf * D/2, +D/2) of the row
            constexpr int HW = D / 2;
            uint4* orow = reinterpret_cast<uint4*>(out + (static_cast<size_t>(img) * L + row) * C + hs * 64 + (D == 32 ? 32 * w : 0) + hf * HW);
            float o[32];
            if (D == 64) tmem_ld32(lane_taddr + 128 + 32 * hf, o);
            else tmem_ld16(lane_taddr + 128 + 16 * hf, o);
            if (row < p.rows_main) {
#pragma unroll
                for (int j = 0; j < HW / 8; ++j) {
                    uint4 u;
                    u.x = pack2<F16>(o[8 * j] * inv, o[8 * j + 1] * inv);
                    u.y = pack2<F16>(o[8 * j + 2] * inv, o[8 * j + 3] * inv);
                    u.z = pack2<F16>(o[8 * j + 4] * inv, o[8 * j + 5] * inv);
                    u.w = pack2<F16>(o[8 * j + 6] * inv, o[8 * j + 7] * inv);
                    orow[j] = u;
                }
            }
            tc_fence_before();                      // the O reads precede the next item's first P*V (ordered through p_full)
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<1>(tmem_base, 512);
}

// Query rows [row0, L) of every (image, head) on the CUDA cores: one warp per row.  Lane k owns keys k, k + 32, ...
// (at most FA_ROWS_MAXJ * 32 keys) for both the scores and P * V; max, sum and the output vector are reduced across the warp.
constexpr int FA_ROWS_MAXJ = 20;

template <int D, bool F16>
__global__ void __launch_bounds__(256)
attention_rows_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ out, int B, int L, int C, int heads, int row0,
                      float scale_log2e) {
    pdl_launch_dependents();
    pdl_wait();
    const int lane = threadIdx.x & 31;
    const int nrow = L - row0;
    const int n_tasks = B * heads * nrow;
    const int wpb = blockDim.x >> 5;
    for (int task = blockIdx.x * wpb + (threadIdx.x >> 5); task < n_tasks; task += gridDim.x * wpb) {
        const int r = row0 + task % nrow;
        const int bh = task / nrow;
        const int h = bh % heads, img = bh / heads;
        const bf16* base = qkv + static_cast<size_t>(img) * L * 3 * C + h * D;
        float qf[D];
        {
            const uint4* qg = reinterpret_cast<const uint4*>(base + static_cast<size_t>(r) * 3 * C);
#pragma unroll
            for (int i = 0; i < D / 8; ++i) {
                const uint4 u = __ldg(qg + i);
                const uint32_t wd[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack2<F16>(wd[j]);
                    qf[8 * i + 2 * j] = f.x;
                    qf[8 * i + 2 * j + 1] = f.y;
                }
            }
        }
        float sc[FA_ROWS_MAXJ];
        float mx = -INFINITY;
#pragma unroll
        for (int j = 0; j < FA_ROWS_MAXJ; ++j) {
            const int key = j * 32 + lane;
            float acc = -INFINITY;
            if (key < L) {
                const uint4* kg = reinterpret_cast<const uint4*>(base + static_cast<size_t>(key) * 3 * C + C);
                acc = 0.f;
#pragma unroll
                for (int i = 0; i < D / 8; ++i) {
                    const uint4 u = __ldg(kg + i);
                    const uint32_t wd[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const float2 f = unpack2<F16>(wd[t]);
                        acc = fmaf(qf[8 * i + 2 * t], f.x, acc);
                        acc = fmaf(qf[8 * i + 2 * t + 1], f.y, acc);
                    }
                }
            }
            sc[j] = acc;
            mx = fmaxf(mx, acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
#pragma unroll
        for (int j = 0; j < FA_ROWS_MAXJ; ++j) {
            sc[j] = ex2_approx((sc[j] - mx) * scale_log2e);      // masked keys: 2^(-inf) = 0
            sum += sc[j];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        // P * V: every lane accumulates its own keys over all D dimensions (independent 16-byte loads that pipeline),
        // then the 32 partial vectors are summed across the warp
        float acc[D];
#pragma unroll
        for (int i = 0; i < D; ++i) acc[i] = 0.f;
#pragma unroll
        for (int j = 0; j < FA_ROWS_MAXJ; ++j) {
            const int key = j * 32 + lane;
            if (key < L) {
                const uint4* vg = reinterpret_cast<const uint4*>(base + static_cast<size_t>(key) * 3 * C + 2 * C);
#pragma unroll
                for (int i = 0; i < D / 8; ++i) {
                    const uint4 u = __ldg(vg + i);
                    const uint32_t wd[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const float2 f = unpack2<F16>(wd[t]);
                        acc[8 * i + 2 * t] = fmaf(sc[j], f.x, acc[8 * i + 2 * t]);
                        acc[8 * i + 2 * t + 1] = fmaf(sc[j], f.y, acc[8 * i + 2 * t + 1]);
                    }
                }
            }
        }
        // transpose-reduce: after step o, lane keeps the half of the vector selected by its bit o; 5 steps leave lane e
        // with the total of dimensions {e, e + 32, ...} -- here plain butterfly sums, D * 5 shuffles
#pragma unroll
        for (int i = 0; i < D; ++i) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
        }
        const float inv = __frcp_rn(sum);
        if (lane == 0) {
            uint4* orow = reinterpret_cast<uint4*>(out + (static_cast<size_t>(img) * L + r) * C + h * D);
#pragma unroll
            for (int j = 0; j < D / 8; ++j) {
                uint4 u;
                u.x = pack2<F16>(acc[8 * j] * inv, acc[8 * j + 1] * inv);
                u.y = pack2<F16>(acc[8 * j + 2] * inv, acc[8 * j + 3] * inv);
                u.z = pack2<F16>(acc[8 * j + 4] * inv, acc[8 * j + 5] * inv);
                u.w = pack2<F16>(acc[8 * j + 6] * inv, acc[8 * j + 7] * inv);
                orow[j] = u;
            }
        }
    }
}

}  // namespace ddc

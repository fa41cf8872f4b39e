// Attention for long token sequences on the tcgen05 tensor cores (ViT.py:107-114): S2 (257 tokens, head_dim 64) and
// S3 (626 tokens, head_dim 32), where the L^2 work is large enough for 128-row UMMA tiles to pay.
//
// Work item = (image, 64 qkv columns) = one head of 64 or a pair of heads of 32.  Everything arrives by TMA as
// [128 rows x 128 B] boxes in the 128-byte swizzle: Q tiles, and K / V in chunks of 128 keys through a 3-stage ring.
// Per (query tile, key chunk):
//     S   = Q K_c^T           tcgen05.mma, both operands K-major, fp32 scores in TMEM (128 columns)
//     P   = exp2(S*c - m)     four softmax warps per tile, one query row per thread straight from TMEM: running max m
//                             and sum l in registers, P written as 16-bit rows of a K-major shared-memory tile
//     O_c = P V_c             tcgen05.mma with V as it lies in memory ([key][dim], i.e. an MN-major B operand), fresh
//                             accumulator per chunk; the softmax warps fold O_c into fp32 registers one chunk late
//                             (O = (O + O_{c-1}) * 2^(m_old - m_new)), so the P*V latency hides behind the max pass
// Two softmax warpgroups run side by side on one tensor pipe: for head_dim 64 they take two consecutive query tiles
// of the same head, for head_dim 32 the two heads of the pair on the same query tile (the head is a 64-byte offset
// inside the 128-byte rows).  The exponentials are the bound (16 MUFU results per clock per SM); the tensor pipe
// idles about three quarters of the time, which is why nothing here is double-buffered beyond the K/V ring.
//
// A token count of the form 128 n + r with small r (S2: 257 = 2 * 128 + 1) would waste a whole 128-row tile on r
// rows: those rows are left to attention_rows_kernel, one warp per (image, head, row) on the CUDA cores.
//
// TMEM (512 columns allocated): warpgroup w uses S at [192 w, +128) and O_c at [192 w + 128, +64).
#pragma once

#include "attention.cuh"
#include "gemm2.cuh"

namespace ddc {

constexpr int FA_THREADS = 384;            // warp 0 TMA, warp 1 MMA, warp 2 TMEM, warp 3 spare, warps 4-7 / 8-11 softmax
constexpr int FA_KCH = 128;                // keys per chunk
constexpr int FA_TILE = 128 * 128;         // bytes of a [128 rows x 128 B] tile
constexpr int FA_STAGES = 3;               // K/V ring stages (K tile + V tile each)
constexpr int FA_TMEM_W = 192;             // TMEM columns per warpgroup
constexpr int FA_SIDE_MAX = 8;             // remainder rows handed to the CUDA-core kernel
constexpr int FA_SMEM_BYTES = 1024 + 2 * FA_TILE + FA_STAGES * 2 * FA_TILE + 2 * 2 * FA_TILE + 512;

struct FaParams {
    int L, C, n_items, n_hs;       // tokens, embed_dim, images * head slots, head slots (64 columns) per image
    uint32_t n_hs_magic;
    int n_qt;                      // 128-row query tiles handled by the tensor kernel
    int rows_main;                 // query rows [0, rows_main) of every image belong to those tiles
    int n_kc;                      // key chunks
    float scale_log2e;
};

inline void fa_split_rows(int L, int* n_qt, int* rows_main) {
    const int full = L / 128, r = L % 128;
    if (full > 0 && r > 0 && r <= FA_SIDE_MAX) { *n_qt = full; *rows_main = full * 128; }
    else { *n_qt = (L + 127) / 128; *rows_main = L; }
}

template <int D, bool F16>
__global__ void __launch_bounds__(FA_THREADS, 1)
attention_fa_kernel(const __grid_constant__ CUtensorMap tm,       // qkv [rows, 3C] 16 bit, box 64 cols x 128 rows
                    bf16* __restrict__ out, const FaParams p) {
    extern __shared__ uint8_t fa_raw[];
    pdl_launch_dependents();
    const uint32_t ra = smem_u32(fa_raw);
    uint8_t* sm = fa_raw + (((ra + 1023u) & ~1023u) - ra);
    uint8_t* Qs = sm;                                   // 2 tiles
    uint8_t* KV = Qs + 2 * FA_TILE;                     // FA_STAGES x (K tile, V tile)
    uint8_t* Ps = KV + FA_STAGES * 2 * FA_TILE;         // per warpgroup: 2 K blocks of 64 keys
    uint64_t* bars = reinterpret_cast<uint64_t*>(Ps + 2 * 2 * FA_TILE);
    uint64_t* q_full = bars;             // [1]
    uint64_t* q_empty = bars + 1;        // [1]
    uint64_t* kv_full = bars + 2;        // [3]
    uint64_t* kv_empty = bars + 5;       // [3]
    uint64_t* s_full = bars + 8;         // [2] S of warpgroup w complete (commit)
    uint64_t* s_free = bars + 10;        // [2] warpgroup w has read S out (4 warps)
    uint64_t* p_full = bars + 12;        // [2] P of warpgroup w written, O_c of the previous chunk read (4 warps)
    uint64_t* o_full = bars + 14;        // [2] O_c of warpgroup w complete (commit)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tm);
        mbar_init(q_full, 1);
        mbar_init(q_empty, 1);
        for (int i = 0; i < FA_STAGES; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1); }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s_full[i], 1);
            mbar_init(&s_free[i], 4);
            mbar_init(&p_full[i], 4);
            mbar_init(&o_full[i], 1);
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc<1>(tmem_slot, 512);
        tmem_relinquish<1>();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                         // qkv comes from the previous kernel

    const int L = p.L, C = p.C;
    const int n_qi = D == 64 ? (p.n_qt + 1) / 2 : p.n_qt;        // query iterations per item
    auto split = [&](int item, int& img, int& hs) {
        img = p.n_hs == 1 ? item : static_cast<int>(__umulhi(static_cast<uint32_t>(item), p.n_hs_magic));
        hs = item - img * p.n_hs;
    };
    auto tile_of = [&](int qi, int w) { return D == 64 ? 2 * qi + w : qi; };

    if (warp == 0) {
        // ================================================================= TMA producer
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            uint32_t q_it = 0, kv_it = 0;
            for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
                int img, hs;
                split(item, img, hs);
                const int row0 = img * L;
                for (int qi = 0; qi < n_qi; ++qi, ++q_it) {
                    mbar_wait(q_empty, (q_it & 1) ^ 1);
                    const int nq = (D == 64 && tile_of(qi, 1) < p.n_qt) ? 2 : 1;
                    mbar_arrive_expect_tx(q_full, nq * FA_TILE);
                    for (int j = 0; j < nq; ++j)
                        tma_load_2d_hint(Qs + j * FA_TILE, &tm, q_full, hs * 64, row0 + tile_of(qi, j) * 128, pol);
                    for (int c = 0; c < p.n_kc; ++c, ++kv_it) {
                        const uint32_t st = kv_it % FA_STAGES, ph = (kv_it / FA_STAGES) & 1u;
                        mbar_wait(&kv_empty[st], ph ^ 1);
                        mbar_arrive_expect_tx(&kv_full[st], 2 * FA_TILE);
                        uint8_t* dst = KV + static_cast<size_t>(st) * 2 * FA_TILE;
                        tma_load_2d_hint(dst, &tm, &kv_full[st], C + hs * 64, row0 + c * FA_KCH, pol);
                        tma_load_2d_hint(dst + FA_TILE, &tm, &kv_full[st], 2 * C + hs * 64, row0 + c * FA_KCH, pol);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ================================================================= MMA issuer
        const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t idesc_o = umma_idesc_16bit(128, D, F16) | (1u << 16);      // B operand (V) is MN-major
        const uint64_t qd = umma_desc_sw128(smem_u32(Qs));
        const uint64_t kvd = umma_desc_sw128(smem_u32(KV));
        const uint64_t pd = umma_desc_sw128(smem_u32(Ps));
        constexpr uint32_t kTileU = FA_TILE >> 4;
        uint32_t q_it = 0, kv_it = 0, sfree_it[2] = {0, 0}, pfull_it[2] = {0, 0};
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            for (int qi = 0; qi < n_qi; ++qi, ++q_it) {
                const bool act[2] = {true, tile_of(qi, 1) < p.n_qt};
                mbar_wait(q_full, q_it & 1);
                tc_fence_after();
                // O_c(cc) = P(cc) V(cc) for both warpgroups, then the K/V stage goes back to the producer
                auto issue_pv = [&](int cc, uint32_t it) {
                    const uint32_t st = it % FA_STAGES;
                    const int kvalid = L - cc * FA_KCH < FA_KCH ? L - cc * FA_KCH : FA_KCH;
                    const int nks = (kvalid + 15) >> 4;
                    const uint64_t vd = kvd + static_cast<uint64_t>((st * 2 + 1) * kTileU);
#pragma unroll
                    for (int w = 0; w < 2; ++w) {
                        if (!act[w]) continue;
                        mbar_wait(&p_full[w], pfull_it[w] & 1);
                        ++pfull_it[w];
                        tc_fence_after();
                        if (elect_one()) {
                            for (int ks = 0; ks < nks; ++ks) {
                                const uint64_t a = pd + static_cast<uint64_t>(w * 2 * kTileU + (ks >> 2) * kTileU + (ks & 3) * 2);
                                const uint64_t b = vd + static_cast<uint64_t>(ks * (2048 >> 4) + (D == 32 ? 4 * w : 0));
                                umma_bf16_ss(tb + w * FA_TMEM_W + 128, a, b, idesc_o, ks > 0 ? 1u : 0u);
                            }
                            umma_commit(&o_full[w]);
                        }
                        __syncwarp();
                    }
                    if (elect_one()) umma_commit(&kv_empty[st]);
                    __syncwarp();
                };
                for (int c = 0; c < p.n_kc; ++c, ++kv_it) {
                    const uint32_t st = kv_it % FA_STAGES, ph = (kv_it / FA_STAGES) & 1u;
                    const int kvalid = L - c * FA_KCH < FA_KCH ? L - c * FA_KCH : FA_KCH;
                    const uint32_t idesc_s = umma_idesc_16bit(128, static_cast<uint32_t>((kvalid + 15) & ~15), F16);
                    mbar_wait(&kv_full[st], ph);
                    tc_fence_after();
                    const uint64_t kd = kvd + static_cast<uint64_t>(st * 2 * kTileU);
#pragma unroll
                    for (int w = 0; w < 2; ++w) {
                        if (!act[w]) continue;
                        mbar_wait(&s_free[w], (sfree_it[w] & 1) ^ 1);       // S of the previous chunk has been read out
                        ++sfree_it[w];
                        tc_fence_after();
                        if (elect_one()) {
                            if (D == 64) {
                                const uint64_t a = qd + static_cast<uint64_t>(w * kTileU);
#pragma unroll
                                for (int ks = 0; ks < 4; ++ks) umma_bf16_ss(tb + w * FA_TMEM_W, a + 2 * ks, kd + 2 * ks, idesc_s, ks);
                            } else {            // head w of the pair: bytes [64 w, 64 w + 64) of every row
#pragma unroll
                                for (int ks = 0; ks < 2; ++ks)
                                    umma_bf16_ss(tb + w * FA_TMEM_W, qd + 4 * w + 2 * ks, kd + 4 * w + 2 * ks, idesc_s, ks);
                            }
                            umma_commit(&s_full[w]);
                        }
                        __syncwarp();
                    }
                    if (c == p.n_kc - 1) {                  // every S of this query iteration is issued: Q may be replaced
                        if (elect_one()) umma_commit(q_empty);
                        __syncwarp();
                    }
                    if (c > 0) issue_pv(c - 1, kv_it - 1);
                }
                issue_pv(p.n_kc - 1, kv_it - 1);
            }
        }
    } else if (warp >= 4) {
        // ================================================================= softmax warpgroups
        const int w = (warp - 4) >> 2;
        const int q = warp & 3;
        const int trow = q * 32 + lane;
        const uint32_t lane_taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(w * FA_TMEM_W);
        uint8_t* Pw = Ps + static_cast<size_t>(w) * 2 * FA_TILE;
        const float sc = p.scale_log2e;
        uint32_t sfull_it = 0, ofull_it = 0;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            int img, hs;
            split(item, img, hs);
            for (int qi = 0; qi < n_qi; ++qi) {
                const int tile = tile_of(qi, w);
                if (tile >= p.n_qt) continue;
                const int row = tile * 128 + trow;
                float m = -INFINITY, l = 0.f;
                float o[D];
#pragma unroll
                for (int i = 0; i < D; ++i) o[i] = 0.f;
                auto fold_o = [&]() {                   // o += O_c of the chunk before
                    mbar_wait(&o_full[w], ofull_it & 1);
                    ++ofull_it;
                    tc_fence_after();
#pragma unroll
                    for (int h = 0; h < D / 32; ++h) {
                        float v[32];
                        tmem_ld32(lane_taddr + 128 + 32 * h, v);
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[32 * h + i] += v[i];
                    }
                };
                for (int c = 0; c < p.n_kc; ++c) {
                    const int kvalid = L - c * FA_KCH < FA_KCH ? L - c * FA_KCH : FA_KCH;
                    const int npiece = (kvalid + 31) >> 5;
                    mbar_wait(&s_full[w], sfull_it & 1);
                    ++sfull_it;
                    tc_fence_after();
                    // ---- pass 1: row maximum of the raw scores
                    float mx = -INFINITY;
                    for (int pc = 0; pc < npiece; ++pc) {
                        float v[32];
                        tmem_ld32(lane_taddr + 32 * pc, v);
                        if (32 * pc + 32 > kvalid) {
#pragma unroll
                            for (int i = 0; i < 32; ++i) if (32 * pc + i >= kvalid) v[i] = -INFINITY;
                        }
                        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
                            m4[0] = fmaxf(m4[0], v[i]); m4[1] = fmaxf(m4[1], v[i + 1]);
                            m4[2] = fmaxf(m4[2], v[i + 2]); m4[3] = fmaxf(m4[3], v[i + 3]);
                        }
                        mx = fmaxf(mx, fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])));
                    }
                    const float m_new = fmaxf(m, mx * sc);          // scaled (log2) domain; scale > 0
                    if (c > 0) fold_o();
                    const float alpha = ex2_approx(m - m_new);      // first chunk: 2^(-inf) = 0 on an all-zero state
                    m = m_new;
                    l *= alpha;
#pragma unroll
                    for (int i = 0; i < D; ++i) o[i] *= alpha;
                    // ---- pass 2: P = 2^(S*c - m) -> 16-bit rows of the K-major P tile
                    for (int pc = 0; pc < npiece; ++pc) {
                        float v[32];
                        tmem_ld32(lane_taddr + 32 * pc, v);
                        if (pc == npiece - 1) {                     // S is in registers: the next chunk's S may overwrite it
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&s_free[w]);
                        }
                        float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                float e = ex2_approx(fmaf(v[i + j], sc, -m));
                                if (32 * pc + 32 > kvalid && 32 * pc + i + j >= kvalid) e = 0.f;
                                v[i + j] = e;
                                s4[j] += e;
                            }
                        }
                        l += (s4[0] + s4[1]) + (s4[2] + s4[3]);
                        uint8_t* kblk = Pw + (pc >> 1) * FA_TILE;
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            uint4 u;
                            u.x = pack2<F16>(v[8 * j], v[8 * j + 1]);
                            u.y = pack2<F16>(v[8 * j + 2], v[8 * j + 3]);
                            u.z = pack2<F16>(v[8 * j + 4], v[8 * j + 5]);
                            u.w = pack2<F16>(v[8 * j + 6], v[8 * j + 7]);
                            *reinterpret_cast<uint4*>(kblk + slot_off(trow, 4 * (pc & 1) + j)) = u;
                        }
                    }
                    fence_proxy_async_smem();           // P -> visible to the tensor core
                    tc_fence_before();                  // and this warp's reads of O_c precede the next P*V
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&p_full[w]);
                }
                fold_o();
                if (row < p.rows_main) {
                    const float inv = __frcp_rn(l);
                    uint4* orow = reinterpret_cast<uint4*>(out + (static_cast<size_t>(img) * L + row) * C + hs * 64 + (D == 32 ? 32 * w : 0));
#pragma unroll
                    for (int j = 0; j < D / 8; ++j) {
                        uint4 u;
                        u.x = pack2<F16>(o[8 * j] * inv, o[8 * j + 1] * inv);
                        u.y = pack2<F16>(o[8 * j + 2] * inv, o[8 * j + 3] * inv);
                        u.z = pack2<F16>(o[8 * j + 4] * inv, o[8 * j + 5] * inv);
                        u.w = pack2<F16>(o[8 * j + 6] * inv, o[8 * j + 7] * inv);
                        orow[j] = u;
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<1>(tmem_base, 512);
}

// Query rows [row0, L) of every (image, head) on the CUDA cores: one warp per row.  Lane k scores keys k, k + 32, ...
// (at most FA_ROWS_MAXJ * 32 keys), the warp reduces max and sum, then lane e accumulates output dimension e (and e + 32).
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
        const uint16_t* vg = reinterpret_cast<const uint16_t*>(base + 2 * C) + lane;
        float a0 = 0.f, a1 = 0.f;
#pragma unroll
        for (int j = 0; j < FA_ROWS_MAXJ; ++j) {
            if (j * 32 >= L) break;
            for (int t = 0; t < 32; ++t) {
                const int key = j * 32 + t;
                if (key >= L) break;
                const float pk = __shfl_sync(0xffffffffu, sc[j], t);
                const uint16_t* vrow = vg + static_cast<size_t>(key) * 3 * C;
                a0 = fmaf(pk, unpack1<F16>(__ldg(vrow)), a0);
                if (D == 64) a1 = fmaf(pk, unpack1<F16>(__ldg(vrow + 32)), a1);
            }
        }
        const float inv = __frcp_rn(sum);
        uint16_t* orow = reinterpret_cast<uint16_t*>(out + (static_cast<size_t>(img) * L + r) * C + h * D) + lane;
        orow[0] = static_cast<uint16_t>(pack2<F16>(a0 * inv, 0.f) & 0xFFFFu);
        if (D == 64) orow[32] = static_cast<uint16_t>(pack2<F16>(a1 * inv, 0.f) & 0xFFFFu);
    }
}

}  // namespace ddc

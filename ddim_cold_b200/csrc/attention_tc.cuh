// Short-sequence attention on the tcgen05 tensor cores (ViT.py:107-114) for head_dim 32 and L = 65 tokens (S1: 64
// patches + the class token).  The warp-level mma.sync kernel in attention.cuh tops out near 40 % of the legacy tensor pipe with
// the MUFU pipe just as busy; here the two contractions run as UMMA and four warps do nothing but softmax.
//
// Work item = (image, 64 qkv columns = two heads of 32).  One 128-row UMMA tile serves BOTH heads: the Q rows of
// tokens 0..63 are loaded twice (tile rows 0..63 and 64..127); S_A = Qtile * K_A^T and S_B = Qtile * K_B^T are two
// MMAs over the same A tile (head selected by a 64-byte start offset inside the 128-byte rows), and softmax warps
// 0,1 read rows 0..63 of S_A while warps 2,3 read rows 64..127 of S_B -- all four TMEM lane quarters do useful work.
// P (bf16, unnormalised) goes to a K-major shared-memory tile (rows 0..63 head A, 64..127 head B); O_h = P * V_h with
// V_h^T produced by an explicit shared-memory transpose (both UMMA operands stay K-major, 128-byte swizzle).
// Token 64 (L = 65 is 64 + the class token) is one row per head: two side warps compute it on the CUDA cores straight
// from global memory, beside the tensor path.  Q/K/V shared memory is released as soon as the S MMAs and the transposes
// are done, so the next item's TMA loads overlap softmax, P*V and the output stage; two CTAs share an SM.
//
// TMEM (256 columns per CTA): S_A [0,80), S_B [96,176), O_A [192,224), O_B [224,256).
#pragma once

#include "attention.cuh"
#include "gemm2.cuh"

namespace ddc {

constexpr int ATC_THREADS = 256;      // warp 0 TMA, warp 1 MMA, warps 2-3 side rows (warp 2 also owns TMEM), warps 4-7 softmax
constexpr int ATC_KEYS = 80;          // key rows per box (multiple of 16 >= L)
constexpr int ATC_Q_BYTES = 2 * 64 * 128;                 // the 64 query rows, twice
constexpr int ATC_KV_BYTES = ATC_KEYS * 128;              // 10 KB, a multiple of 1024
constexpr int ATC_VT_BYTES = 2 * 2 * 32 * 128;            // [head][K block][32 rows x 128 B]
constexpr int ATC_P_BYTES = 2 * 128 * 128;                // [K block][128 rows x 128 B]
constexpr int ATC_SMEM_BYTES = 1024 + ATC_Q_BYTES + 2 * ATC_KV_BYTES + ATC_VT_BYTES + ATC_P_BYTES + 256;

inline bool attention_tc_usable(int L, int C, int heads) {
    return heads > 0 && C % heads == 0 && C / heads == 32 && C % 64 == 0 && L == 65;     // 64 tile rows + one side row
}

// probe builds (DDC_PROBE_TIMING) stamp the SM clock of CTA 0's softmax warp 4 into this buffer: [item][8]
#ifdef DDC_PROBE_TIMING
__device__ long long* g_atc_dbg = nullptr;
#define ATC_STAMP(k) do { if (g_atc_dbg != nullptr && blockIdx.x == 0 && warp == 4 && lane == 0 && it < 8) g_atc_dbg[it * 8 + (k)] = clock64(); } while (0)
#else
#define ATC_STAMP(k)
#endif

template <bool F16>
__global__ void __launch_bounds__(ATC_THREADS, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQ,      // qkv [rows, 3C], box 64 cols x 64 rows
                    const __grid_constant__ CUtensorMap tmKV,     // qkv [rows, 3C], box 64 cols x 80 rows
                    const bf16* __restrict__ qkv, bf16* __restrict__ out, int L, int C, int n_items,
                    uint32_t n_hg_magic, float scale_log2e) {
    extern __shared__ uint8_t atc_raw[];
    pdl_launch_dependents();
    const uint32_t ra = smem_u32(atc_raw);
    uint8_t* sm = atc_raw + (((ra + 1023u) & ~1023u) - ra);
    uint8_t* Qs = sm;
    uint8_t* Ks = Qs + ATC_Q_BYTES;
    uint8_t* Vs = Ks + ATC_KV_BYTES;
    uint8_t* VT = Vs + ATC_KV_BYTES;
    uint8_t* Ps = VT + ATC_VT_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(Ps + ATC_P_BYTES);
    uint64_t* full_bar = bars;          // Q, K, V landed (TMA bytes)
    uint64_t* s_bar = bars + 1;         // both S MMAs complete (commit): Q and K no longer read by the tensor core
    uint64_t* rel_bar = bars + 2;       // transposes (4 warps) done with V
    uint64_t* p_bar = bars + 3;         // P and V^T written (4 warps)
    uint64_t* o_bar = bars + 4;         // both O MMAs complete (commit)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_hg = C / 64;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmKV);
        mbar_init(full_bar, 1);
        mbar_init(s_bar, 1);
        mbar_init(rel_bar, 4);
        mbar_init(p_bar, 4);
        mbar_init(o_bar, 1);
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc<1>(tmem_slot, 256);
        tmem_relinquish<1>();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                         // qkv comes from the previous kernel

    auto split = [&](int item, int& img, int& hg) {
        img = n_hg == 1 ? item : static_cast<int>(__umulhi(static_cast<uint32_t>(item), n_hg_magic));
        hg = item - img * n_hg;
    };

    if (warp == 0) {
        // ================================================================= TMA producer
        if (lane == 0) {
            int it = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
                if (it > 0) {           // the previous item's readers of Q / K / V are done
                    mbar_wait(s_bar, (it - 1) & 1);
                    mbar_wait(rel_bar, (it - 1) & 1);
                }
                int img, hg;
                split(item, img, hg);
#ifdef DDC_PROBE_TIMING
                if (g_atc_dbg != nullptr && blockIdx.x == 0 && it < 8) g_atc_dbg[it * 8 + 7] = clock64();
#endif
                mbar_arrive_expect_tx(full_bar, ATC_Q_BYTES + 2 * ATC_KV_BYTES);
                tma_load_2d(Qs, &tmQ, full_bar, hg * 64, img * L);
                tma_load_2d(Qs + 64 * 128, &tmQ, full_bar, hg * 64, img * L);
                tma_load_2d(Ks, &tmKV, full_bar, C + hg * 64, img * L);
                tma_load_2d(Vs, &tmKV, full_bar, 2 * C + hg * 64, img * L);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ================================================================= MMA issuer
        const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t idesc_s = umma_idesc_16bit(128, ATC_KEYS, F16);
        const uint32_t idesc_o = umma_idesc_16bit(128, 32, F16);
        const uint64_t qd = umma_desc_sw128(smem_u32(Qs));
        const uint64_t kd = umma_desc_sw128(smem_u32(Ks));
        const uint64_t pd = umma_desc_sw128(smem_u32(Ps));
        const uint64_t vd = umma_desc_sw128(smem_u32(VT));
        int it = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
            mbar_wait(full_bar, it & 1);
            tc_fence_after();
            if (elect_one()) {
                // S_h = Q[:, head h] * K[:, head h]^T : K = 32 = two K16 steps; head h starts 64 B into the 128-byte rows
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks)
                        umma_bf16_ss(tb + h * 96, qd + 4 * h + 2 * ks, kd + 4 * h + 2 * ks, idesc_s, ks);
                umma_commit(s_bar);
            }
            __syncwarp();
            mbar_wait(p_bar, it & 1);
            tc_fence_after();
            if (elect_one()) {
                // O_h = P * V_h : K = 80 keys = four K16 steps in K block 0 and one in K block 1
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const uint64_t vh = vd + static_cast<uint64_t>(h * ((2 * 32 * 128) >> 4));
#pragma unroll
                    for (int ks = 0; ks < 5; ++ks) {
                        const uint64_t a = pd + (ks < 4 ? 2 * ks : ((128 * 128) >> 4));
                        const uint64_t b = vh + (ks < 4 ? 2 * ks : ((32 * 128) >> 4));
                        umma_bf16_ss(tb + 192 + h * 32, a, b, idesc_o, ks);
                    }
                }
                umma_commit(o_bar);
            }
            __syncwarp();
        }
    } else if (warp < 4) {
        // ================================================================= side rows: token 64 of head (warp - 2)
        // Entirely from global memory (L2: the same lines the TMA loads fetch) and without any barrier: these warps
        // run at their own pace beside the tensor pipeline and never hold up the release of the Q/K/V stage.
        const int h = warp - 2;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int img, hg;
            split(item, img, hg);
            const bf16* base = qkv + static_cast<size_t>(img) * L * 3 * C + hg * 64 + h * 32;
            const uint4* qg = reinterpret_cast<const uint4*>(base + static_cast<size_t>(64) * 3 * C);
            float qf[32];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint4 qv = __ldg(qg + i);
                const uint32_t w[4] = {qv.x, qv.y, qv.z, qv.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack2<F16>(w[j]);
                    qf[8 * i + 2 * j] = f.x;
                    qf[8 * i + 2 * j + 1] = f.y;
                }
            }
            // scores: lane handles keys lane, lane + 32, lane + 64
            float sc[3];
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const int key = lane + 32 * r;
                float acc = -INFINITY;
                if (key < L) {
                    const uint4* kg = reinterpret_cast<const uint4*>(base + static_cast<size_t>(key) * 3 * C + C);
                    acc = 0.f;
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const uint4 kv = __ldg(kg + c);
                        const uint32_t w[4] = {kv.x, kv.y, kv.z, kv.w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float2 f = unpack2<F16>(w[j]);
                            acc = fmaf(qf[8 * c + 2 * j], f.x, acc);
                            acc = fmaf(qf[8 * c + 2 * j + 1], f.y, acc);
                        }
                    }
                }
                sc[r] = acc;
            }
            float mx = fmaxf(sc[0], fmaxf(sc[1], sc[2]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            float pr[3], sum = 0.f;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                pr[r] = ex2_approx((sc[r] - mx) * scale_log2e);      // exp2(-inf) = 0 for the masked keys
                sum += pr[r];
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            // out[e = lane] = sum_k p_k * V[k][h*32 + e]: one coalesced 64-byte read per key, fully unrolled
            const uint16_t* vg = reinterpret_cast<const uint16_t*>(base + 2 * C) + lane;
            float acc = 0.f, acc2 = 0.f;
#pragma unroll
            for (int r = 0; r < 2; ++r) {
#pragma unroll
                for (int k = 0; k < 32; k += 2) {
                    const int key = k + 32 * r;
                    const float p0 = __shfl_sync(0xffffffffu, pr[r], k);
                    const float p1 = __shfl_sync(0xffffffffu, pr[r], k + 1);
                    acc = fmaf(p0, unpack1<F16>(__ldg(vg + static_cast<size_t>(key) * 3 * C)), acc);
                    acc2 = fmaf(p1, unpack1<F16>(__ldg(vg + static_cast<size_t>(key + 1) * 3 * C)), acc2);
                }
            }
            for (int key = 64; key < L; ++key) {
                const float p0 = __shfl_sync(0xffffffffu, pr[2], key - 64);
                acc = fmaf(p0, unpack1<F16>(__ldg(vg + static_cast<size_t>(key) * 3 * C)), acc);
            }
            const float o = (acc + acc2) * __frcp_rn(sum);
            const float on = __shfl_down_sync(0xffffffffu, o, 1);
            if ((lane & 1) == 0)
                *reinterpret_cast<uint32_t*>(out + (static_cast<size_t>(img) * L + 64) * C + hg * 64 + h * 32 + lane) = pack2<F16>(o, on);
        }
    } else {
        // ================================================================= softmax warps (TMEM lane quarter q)
        const int q = warp & 3;
        const int h = q >> 1;                         // rows 0..63 belong to head A, 64..127 to head B
        const int trow = q * 32 + lane;               // tile row
        const int token = trow & 63;
        const uint32_t lane_taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        int it = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
            int img, hg;
            split(item, img, hg);
            ATC_STAMP(0);
            mbar_wait(full_bar, it & 1);
            ATC_STAMP(1);
            // ---- V^T: thread `tid` < 80 owns key row tid and scatters its 64 values into the two heads' K-major tiles
            // (V^T and P of the previous item are free: this warp waited for o_bar(it - 1) in its epilogue)
            // one ldmatrix.x4.trans + stmatrix.x4 moves an [8 keys x 32 dims] block of one head: lane l reads row
            // (8g + l%8) chunk (4*hh + l/8) of V and writes the 16-byte chunk "keys 8g..8g+7" of V^T row l
#pragma unroll
            for (int k5 = 0; k5 < (2 * (ATC_KEYS / 8)) / 4; ++k5) {
                const int pi = q + 4 * k5;
                const int g8 = pi >> 1, hh = pi & 1;
                const int r = lane & 7, m = lane >> 3;
                uint32_t t0, t1, t2, t3;
                ldmatrix_x4_trans(t0, t1, t2, t3, smem_u32(Vs + (8 * g8 + r) * 128 + (((4 * hh + m) ^ r) << 4)));
                stmatrix_x4(smem_u32(VT + hh * (2 * 32 * 128) + (g8 >> 3) * (32 * 128) + lane * 128 + (((g8 & 7) ^ (lane & 7)) << 4)),
                            t0, t1, t2, t3);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(rel_bar);               // done with V
            // ---- softmax of this thread's row
            ATC_STAMP(2);
            mbar_wait(s_bar, it & 1);
            tc_fence_after();
            ATC_STAMP(3);
            float s[ATC_KEYS];
            tmem_ld32(lane_taddr + h * 96, s);
            tmem_ld32(lane_taddr + h * 96 + 32, s + 32);
            tmem_ld16(lane_taddr + h * 96 + 64, s + 64);
#pragma unroll
            for (int i = 64; i < ATC_KEYS; ++i)                 // L > 64: only the last chunk holds padded keys
                if (i >= L) s[i] = -INFINITY;
            float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};     // four chains instead of one of 80
#pragma unroll
            for (int i = 0; i < ATC_KEYS; i += 4) {
                m4[0] = fmaxf(m4[0], s[i]); m4[1] = fmaxf(m4[1], s[i + 1]);
                m4[2] = fmaxf(m4[2], s[i + 2]); m4[3] = fmaxf(m4[3], s[i + 3]);
            }
            const float mb = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * scale_log2e;
            float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < ATC_KEYS; i += 4) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    s[i + j] = ex2_approx(fmaf(s[i + j], scale_log2e, -mb));
                    s4[j] += s[i + j];
                }
            }
            const float sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
#pragma unroll
            for (int c = 0; c < ATC_KEYS / 8; ++c) {           // ten 16-byte chunks: eight in K block 0, two in K block 1
                uint4 o;
                o.x = pack2<F16>(s[8 * c], s[8 * c + 1]);
                o.y = pack2<F16>(s[8 * c + 2], s[8 * c + 3]);
                o.z = pack2<F16>(s[8 * c + 4], s[8 * c + 5]);
                o.w = pack2<F16>(s[8 * c + 6], s[8 * c + 7]);
                *reinterpret_cast<uint4*>(Ps + (c >> 3) * (128 * 128) + slot_off(trow, c & 7)) = o;
            }
            fence_proxy_async_smem();                           // P and V^T -> visible to the tensor core
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_bar);
            ATC_STAMP(4);
            // ---- output: O / sum -> staging -> 128-byte row segments
            const float inv = __frcp_rn(sum);
            mbar_wait(o_bar, it & 1);
            tc_fence_after();
            ATC_STAMP(5);
            float o[32];
            tmem_ld32(lane_taddr + 192 + h * 32, o);
            // each thread owns 64 contiguous bytes of the output row (32 columns of its head): straight to global
            uint4* orow = reinterpret_cast<uint4*>(out + (static_cast<size_t>(img) * L + token) * C + hg * 64 + h * 32);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint4 w;
                w.x = pack2<F16>(o[8 * c] * inv, o[8 * c + 1] * inv);
                w.y = pack2<F16>(o[8 * c + 2] * inv, o[8 * c + 3] * inv);
                w.z = pack2<F16>(o[8 * c + 4] * inv, o[8 * c + 5] * inv);
                w.w = pack2<F16>(o[8 * c + 6] * inv, o[8 * c + 7] * inv);
                orow[c] = w;
            }
            tc_fence_before();
            ATC_STAMP(6);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<1>(tmem_base, 256);
}

}  // namespace ddc

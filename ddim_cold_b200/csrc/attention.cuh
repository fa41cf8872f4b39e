// Short-sequence fused attention (ViT.py:107-114): softmax(q k^T * d^-0.5) v per (image, head).
//
// Token counts here are 65 / 257 / 626 with head_dim 32 or 64, so K and V of one head fit in
// shared memory and the [L, L] score matrix never exists in HBM.  Each warp owns 16-query
// tiles, keeps scores / probabilities in registers (flash-style running max and sum in fp32,
// exp2 with the log2e-folded scale) and feeds the probability fragments straight back in as
// the A operand of P*V.  Tiles this small (K = head_dim = 32, M = 65) cannot fill a 128-row
// tcgen05 tile, so the contraction runs on warp-level mma.sync m16n8k16 (bf16 or fp16
// operands, fp32 accumulation).
//
// Two launch shapes:
//   * attention_stream_kernel: persistent CTAs, work item = (image, 64 qkv columns), Q / K / V tiles by TMA into a
//     double-buffered stage, every (head, q-tile) pair a warp task.  Used for short sequences (L <= 128: S1, 65 tokens).
//   * attention_head_kernel: one CTA per (image, head), K/V in shared memory, Q fragments read
//     directly from global.  Used for long sequences (S2 257 tokens, S3 626 tokens).
//
// Layouts: qkv [rows, 3C] with column s*C + head*D + e (s = q,k,v) exactly as the fused qkv
// Linear emits it (ViT.py:107); out [rows, C] with column head*D + e (ViT.py:114).
#pragma once

#include "common.cuh"

namespace ddc {

// 2^x, one MUFU instruction (exp2f adds range fix-ups that softmax does not need; ex2(-inf) = +0)
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// two exponentials per MUFU operation on a packed pair of IEEE halves; the result is the packed pair
__device__ __forceinline__ uint32_t ex2_h2(uint32_t x) {
    uint32_t y;
    asm("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x));
    return y;
}

constexpr int ATT_MAX_WARPS = 8;
constexpr int ATT_GROUP_WARPS = 8;     // group kernel: one warp per 16-query tile, each walking the G heads

// One 16-query tile against all keys of one head.  Ks/Vs point at the head's first column in
// shared memory, RS is the row pitch in bytes, keys >= L are masked, rows [L, lk_pad) must be
// zero.  On return o holds the normalised output fragments.
// SWZ = false: rows of PITCH bytes, Ks/Vs already offset to the head's first column (cbase = 0).
// SWZ = true : Ks/Vs are [rows x 128 B] tiles in the TMA 128-byte swizzle (PITCH = 128); `cbase` is the head's
//              first 16-byte chunk inside the 128-byte row.
// All ldmatrix row offsets are multiples of 8 plus (lane & 7), so the swizzle term (chunk ^ (row & 7)) is a
// per-lane constant: addresses are one add away from a base computed once per tile.
template <bool SWZ>
__device__ __forceinline__ uint32_t att_addr(const uint8_t* base, int row, int chunk, int RS) {
    return SWZ ? smem_u32(base + row * 128 + ((chunk ^ (row & 7)) << 4)) : smem_u32(base + row * RS + (chunk << 4));
}

// SINGLE: the whole key range is one chunk and KC - 16 < L <= KC, so only the last two n-tiles can reach past L
// and the mask is resolved at compile time for all the others (the caller checks the condition).
template <int D, int KC, bool F16, bool SWZ, int PITCH, bool SINGLE>
__device__ __forceinline__ void attend_qtile(const uint32_t (&qa)[D / 16][4], const uint8_t* Ks, const uint8_t* Vs,
                                             int L, int lk_pad, float scale_log2e, int lane,
                                             float (&o)[D / 8][4], int cbase = 0) {
    constexpr int NT = KC / 8;    // score n-tiles per key chunk
    constexpr int DT = D / 8;     // output n-tiles
    constexpr int DK = D / 16;    // k-steps of q k^T
    const int t = lane & 3;
    const int l7 = lane & 7;
    const uint32_t kbase = smem_u32(Ks) + static_cast<uint32_t>((l7 + ((lane >> 4) << 3)) * PITCH);
    const uint32_t vbase = smem_u32(Vs) + static_cast<uint32_t>((l7 + (((lane >> 3) & 1) << 3)) * PITCH);
    uint32_t kx[DK], vx[DT / 2];
#pragma unroll
    for (int ks = 0; ks < DK; ++ks) {
        const int chunk = cbase + ks * 2 + ((lane >> 3) & 1);
        kx[ks] = static_cast<uint32_t>((SWZ ? (chunk ^ l7) : chunk) << 4);
    }
#pragma unroll
    for (int d2 = 0; d2 < DT / 2; ++d2) {
        const int chunk = cbase + d2 * 2 + (lane >> 4);
        vx[d2] = static_cast<uint32_t>((SWZ ? (chunk ^ l7) : chunk) << 4);
    }
#pragma unroll
    for (int i = 0; i < DT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    // row sums ride on the tensor cores: P times a column of ones (every lane then holds its rows' sums)
    constexpr uint32_t kOnes = F16 ? 0x3C003C00u : 0x3F803F80u;
    float lsum[4] = {0.f, 0.f, 0.f, 0.f};
    float m0 = -INFINITY, m1 = -INFINITY;
    for (int kc0 = 0; kc0 < lk_pad; kc0 += KC) {
        float s[NT][4];
#pragma unroll
        for (int i = 0; i < NT; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
        for (int nt = 0; nt < NT; nt += 2) {
#pragma unroll
            for (int ks = 0; ks < DK; ++ks) {
                uint32_t b0, b1, b2, b3;
                ldmatrix_x4(b0, b1, b2, b3, kbase + static_cast<uint32_t>((kc0 + nt * 8) * PITCH) + kx[ks]);
                mma_16816<F16>(s[nt], qa[ks], b0, b1);
                mma_16816<F16>(s[nt + 1], qa[ks], b2, b3);
            }
        }
        // mask only the n-tiles that reach past the last key (raw scores; the scale is folded into the exponent)
        if (SINGLE) {
#pragma unroll
            for (int nt = (NT >= 2 ? NT - 2 : 0); nt < NT; ++nt) {
                const int key = nt * 8 + 2 * t;
                if (key >= L) { s[nt][0] = -INFINITY; s[nt][2] = -INFINITY; }
                if (key + 1 >= L) { s[nt][1] = -INFINITY; s[nt][3] = -INFINITY; }
            }
        } else if (kc0 + KC > L) {          // only the last chunk reaches past L
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const int key = kc0 + nt * 8 + 2 * t;
                if (key >= L) { s[nt][0] = -INFINITY; s[nt][2] = -INFINITY; }
                if (key + 1 >= L) { s[nt][1] = -INFINITY; s[nt][3] = -INFINITY; }
            }
        }
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
            mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        mx0 = fmaxf(m0, mx0 * scale_log2e);       // running max in the scaled (log2) domain; scale > 0
        mx1 = fmaxf(m1, mx1 * scale_log2e);
        const float c0 = ex2_approx(m0 - mx0), c1 = ex2_approx(m1 - mx1);
        m0 = mx0; m1 = mx1;
        if (!SINGLE) {
            lsum[0] *= c0; lsum[2] *= c1;
#pragma unroll
            for (int i = 0; i < DT; ++i) { o[i][0] *= c0; o[i][1] *= c0; o[i][2] *= c1; o[i][3] *= c1; }
        }
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            s[nt][0] = ex2_approx(fmaf(s[nt][0], scale_log2e, -m0));
            s[nt][1] = ex2_approx(fmaf(s[nt][1], scale_log2e, -m0));
            s[nt][2] = ex2_approx(fmaf(s[nt][2], scale_log2e, -m1));
            s[nt][3] = ex2_approx(fmaf(s[nt][3], scale_log2e, -m1));
        }
#pragma unroll
        for (int kk = 0; kk < KC / 16; ++kk) {
            uint32_t pa[4];
            pa[0] = pack2<F16>(s[2 * kk][0], s[2 * kk][1]);
            pa[1] = pack2<F16>(s[2 * kk][2], s[2 * kk][3]);
            pa[2] = pack2<F16>(s[2 * kk + 1][0], s[2 * kk + 1][1]);
            pa[3] = pack2<F16>(s[2 * kk + 1][2], s[2 * kk + 1][3]);
            mma_16816<F16>(lsum, pa, kOnes, kOnes);
#pragma unroll
            for (int dt = 0; dt < DT; dt += 2) {
                uint32_t b0, b1, b2, b3;
                ldmatrix_x4_trans(b0, b1, b2, b3, vbase + static_cast<uint32_t>((kc0 + kk * 16) * PITCH) + vx[dt / 2]);
                mma_16816<F16>(o[dt], pa, b0, b1);
                mma_16816<F16>(o[dt + 1], pa, b2, b3);
            }
        }
    }
    const float i0 = __frcp_rn(lsum[0]), i1 = __frcp_rn(lsum[2]);     // sums are >= 1 (the row maximum contributes 1)
#pragma unroll
    for (int i = 0; i < DT; ++i) { o[i][0] *= i0; o[i][1] *= i0; o[i][2] *= i1; o[i][3] *= i1; }
}

// ---------------------------------------------------------------------------- persistent, double-buffered
// Work item = (image, 64 qkv columns = 64/D heads).  A CTA walks items grid-strided and keeps TWO items' Q/K/V tiles
// in shared memory: the TMA loads of item i+1 are in flight while the warps compute item i, so the DRAM latency that
// a one-shot CTA pays at its start (nothing to do until its only tiles land) is hidden behind tensor-core work.  Three
// such CTAs share an SM.  Q, K and V of an item arrive as three [lk_pad x 64] boxes in the 128-byte swizzle behind
// one mbarrier per stage; every (head, q-tile) pair is a warp task; the output tile is staged over the Q tile and
// leaves as 128-byte row segments.
inline size_t att_stream_smem_bytes(int L, int KC) {
    const int lk = (L + KC - 1) / KC * KC;
    return 1024 + static_cast<size_t>(2 * 3) * lk * 128 + 64;
}

template <int D, int KC, bool F16>
__global__ void __launch_bounds__(ATT_GROUP_WARPS * 32, 2)
attention_stream_kernel(const __grid_constant__ CUtensorMap tmQKV, bf16* __restrict__ out, int L, int C, int n_items,
                        uint32_t n_hg_magic, float scale_log2e) {
    extern __shared__ uint8_t att_raw[];
    pdl_launch_dependents();
    pdl_wait();        // qkv comes from the previous kernel; nothing worth doing before it
    constexpr int GH = 64 / D;            // heads per item
    constexpr int DK = D / 16, DT = D / 8;
    const uint32_t ra = smem_u32(att_raw);
    uint8_t* sm = att_raw + (((ra + 1023u) & ~1023u) - ra);
    const int n_hg = C / 64;
    const int qtiles = (L + 15) / 16;
    const int lk_pad = (L + KC - 1) / KC * KC;      // == box rows (single key chunk: L <= KC)
    const uint32_t tile_bytes = static_cast<uint32_t>(lk_pad) * 128u;
    uint64_t* full = reinterpret_cast<uint64_t*>(sm + 6 * tile_bytes);     // [2] stage loaded (TMA bytes)
    uint64_t* empty = full + 2;                                            // [2] stage released (one arrival per warp)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nwarps = blockDim.x >> 5;
    // item -> (image, 64-column group); n_hg_magic = 2^32 / n_hg + 1 (host), exact for item * n_hg < 2^32
    auto split = [&](int item, int& img, int& hg) {
        img = n_hg == 1 ? item : static_cast<int>(__umulhi(static_cast<uint32_t>(item), n_hg_magic));
        hg = item - img * n_hg;
    };
    auto issue = [&](int item, int s) {           // one thread: the three boxes of `item` into stage s
        int img, hg;
        split(item, img, hg);
        mbar_arrive_expect_tx(&full[s], 3 * tile_bytes);
        // QKV is read exactly once: stream it through L2 (evict_first) so that the attention output written below and
        // the weights stay resident for the block tail that follows
        const uint64_t pol = l2_policy_evict_first();
#pragma unroll
        for (int m = 0; m < 3; ++m)
            tma_load_2d_hint(sm + (s * 3 + m) * tile_bytes, &tmQKV, &full[s], m * C + hg * 64, img * L, pol);
    };
    if (threadIdx.x == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        mbar_init(&empty[0], nwarps);
        mbar_init(&empty[1], nwarps);
        fence_barrier_init();
        if (static_cast<int>(blockIdx.x) < n_items) issue(blockIdx.x, 0);
    }
    __syncthreads();

    // No CTA-wide barrier in the loop: a warp owns whole q-tiles (all GH heads of 16 query rows), stages its output
    // over the Q rows only it reads, writes those rows out itself as 128-byte segments, then releases the stage.
    const int g = lane >> 2, t = lane & 3;
    int it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        const int s = it & 1;
        if (threadIdx.x == 0 && item + static_cast<int>(gridDim.x) < n_items) {
            if (it >= 1) mbar_wait(&empty[s ^ 1], ((it - 1) >> 1) & 1);    // every warp is done with item it-1
            issue(item + gridDim.x, s ^ 1);
        }
        __syncwarp();
        mbar_wait(&full[s], (it >> 1) & 1);
        uint8_t* Qt = sm + (s * 3 + 0) * tile_bytes;
        const uint8_t* Kt = sm + (s * 3 + 1) * tile_bytes;
        const uint8_t* Vt = sm + (s * 3 + 2) * tile_bytes;
        int img, hg;
        split(item, img, hg);
        bf16* ob = out + static_cast<size_t>(img) * L * C + hg * 64;
        for (int qt = warp; qt < qtiles; qt += nwarps) {
            const int q0 = qt * 16;
#pragma unroll
            for (int head = 0; head < GH; ++head) {
                const int cbase = (head * D) / 8;         // first 16-byte chunk of this head in the 128-byte row
                uint32_t qa[DK][4];
#pragma unroll
                for (int ks = 0; ks < DK; ++ks) {
                    const int qrow = q0 + (lane & 7) + (((lane >> 3) & 1) << 3);
                    ldmatrix_x4(qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], att_addr<true>(Qt, qrow, cbase + ks * 2 + (lane >> 4), 0));
                }
                float o[DT][4];
                if (L > KC - 16) attend_qtile<D, KC, F16, true, 128, true>(qa, Kt, Vt, L, lk_pad, scale_log2e, lane, o, cbase);
                else attend_qtile<D, KC, F16, true, 128, false>(qa, Kt, Vt, L, lk_pad, scale_log2e, lane, o, cbase);
                __syncwarp();                             // all lanes have their Q fragments: reuse the patch as stage
#pragma unroll
                for (int dt = 0; dt < DT; ++dt) {
                    const int r0 = q0 + g, r1 = q0 + g + 8;
                    *reinterpret_cast<uint32_t*>(Qt + r0 * 128 + (((cbase + dt) ^ (r0 & 7)) << 4) + 4 * t) = pack2<F16>(o[dt][0], o[dt][1]);
                    *reinterpret_cast<uint32_t*>(Qt + r1 * 128 + (((cbase + dt) ^ (r1 & 7)) << 4) + 4 * t) = pack2<F16>(o[dt][2], o[dt][3]);
                }
            }
            __syncwarp();
#pragma unroll
            for (int i = lane; i < 16 * 8; i += 32) {     // this warp's 16 rows x 128 B
                const int r = q0 + (i >> 3), j = i & 7;
                if (r < L)
                    *reinterpret_cast<uint4*>(ob + static_cast<size_t>(r) * C + j * 8) =
                        *reinterpret_cast<const uint4*>(Qt + r * 128 + ((j ^ (r & 7)) << 4));
            }
        }
        fence_proxy_async_smem();     // generic accesses to this stage are ordered before the TMA refill
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
}

// ---------------------------------------------------------------------------- one CTA per (image, head)
inline size_t att_head_smem_bytes(int L, int D, int KC) {
    const int lk = (L + KC - 1) / KC * KC;
    // K, V, and (head_dim 64) a warp-private output staging tile; with head_dim 32 the staging measured slower (S3: 566 ->
    // 704 us: K and V of 626 tokens leave no room for it beside a second CTA on the SM)
    return static_cast<size_t>(2) * lk * (D + 8) * 2 + (D == 64 ? static_cast<size_t>(ATT_MAX_WARPS) * 16 * (D + 8) * 2 : 0);
}

template <int D, int KC, bool F16>
__global__ void __launch_bounds__(ATT_MAX_WARPS * 32)
attention_head_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ out, int L, int C, float scale_log2e) {
    extern __shared__ __align__(16) uint8_t att_smem[];
    pdl_launch_dependents();
    pdl_wait();        // qkv comes from the previous kernel; nothing worth doing before it
    constexpr int RS = (D + 8) * 2;
    constexpr int DT = D / 8, DK = D / 16;
    const int head = blockIdx.x;
    const int img = blockIdx.y;
    const int lk_pad = (L + KC - 1) / KC * KC;
    uint8_t* Ks = att_smem;
    uint8_t* Vs = att_smem + static_cast<size_t>(lk_pad) * RS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nwarps = blockDim.x >> 5;
    const size_t ld = static_cast<size_t>(3) * C;
    const bf16* base = qkv + static_cast<size_t>(img) * L * ld + head * D;

    constexpr int CPR = D / 8;    // 16-byte chunks per row
    for (int i = threadIdx.x; i < lk_pad * CPR; i += blockDim.x) {
        const int r = i / CPR, c = i - r * CPR;
        if (r < L) {
            cp_async16(smem_u32(Ks + r * RS + c * 16), base + r * ld + C + c * 8);
            cp_async16(smem_u32(Vs + r * RS + c * 16), base + r * ld + 2 * C + c * 8);
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
        const int r0 = q0 + g, r1 = q0 + g + 8;
        uint32_t qa[DK][4];
        {
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
        attend_qtile<D, KC, F16, false, RS, false>(qa, Ks, Vs, L, lk_pad, scale_log2e, lane, o);
        // The fragments hold 4-byte pieces of 16 rows: stored from here an instruction would write 8 rows x 16 B.  Through a
        // warp-private [16 rows x D] tile (row pitch padded by 16 B) every store instruction writes whole 16-byte chunks of
        // full rows instead (uncoalesced stores clog the SM's load/store path, see qkv_attn_tc.cuh).
        bf16* ob = out + static_cast<size_t>(img) * L * C + head * D;
        if (D != 64) {
#pragma unroll
            for (int dt = 0; dt < DT; ++dt) {
                if (r0 < L)
                    *reinterpret_cast<uint32_t*>(ob + static_cast<size_t>(r0) * C + dt * 8 + 2 * t) = pack2<F16>(o[dt][0], o[dt][1]);
                if (r1 < L)
                    *reinterpret_cast<uint32_t*>(ob + static_cast<size_t>(r1) * C + dt * 8 + 2 * t) = pack2<F16>(o[dt][2], o[dt][3]);
            }
            continue;
        }
        uint8_t* stg = Vs + static_cast<size_t>(lk_pad) * RS + static_cast<size_t>(warp) * 16 * RS;
#pragma unroll
        for (int dt = 0; dt < DT; ++dt) {
            *reinterpret_cast<uint32_t*>(stg + g * RS + dt * 16 + 4 * t) = pack2<F16>(o[dt][0], o[dt][1]);
            *reinterpret_cast<uint32_t*>(stg + (g + 8) * RS + dt * 16 + 4 * t) = pack2<F16>(o[dt][2], o[dt][3]);
        }
        __syncwarp();
        for (int i = lane; i < 16 * CPR; i += 32) {
            const int rr = i / CPR, ch = i - rr * CPR;
            if (q0 + rr < L)
                *reinterpret_cast<uint4*>(ob + static_cast<size_t>(q0 + rr) * C + ch * 8) = *reinterpret_cast<const uint4*>(stg + rr * RS + ch * 16);
        }
        __syncwarp();
    }
}

}  // namespace ddc

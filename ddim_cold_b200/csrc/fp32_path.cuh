// "fp32 mode" of the predictor (operand_format = "fp32"): every contraction in fp32 on the CUDA cores, for callers that
// need the reference's own arithmetic precision (the reference computes in fp32 throughout, ViT.py:208-218; north_star
// asks for per-step eps within 1e-4 of it in this mode).  tcgen05 has no fp32 MMA: the 16-bit tensor-core formats reach
// 2e-3 (fp16) / 2e-2 (bf16), TF32 about 1e-3.  So this path trades the tensor cores for exactness: a register-tiled SGEMM
// with the bias / GELU / residual / token-assembly epilogues, two-pass LayerNorm, exact erff GELU, expf softmax -- fp32
// accumulation everywhere, no 16-bit value anywhere.  It is a precision mode, not a fast path (about 3 % of the 16-bit
// path's throughput); the step chain, the step table and the unpatchify + clamp + DDIM kernel are shared.
//
// The weights are read where the caller keeps them (the fp32 state_dict tensors handed to ddc_pack_weights): nothing is
// copied, so those tensors must stay alive and unchanged while the plan is used in this mode.
#pragma once

#include "common.cuh"

namespace ddc {

// ---------------------------------------------------------------------------- patchify, fp32 (ViT.py:150,154)
// ap[b*L + 1 + ph*Wp + pw, c*p*p + p1*p + p2] = x[b, c, ph*p + p1, pw*p + p2]; row b*L (the cls slot) = 0.
// Block (0,0) thread 0 advances the device-side step counter like the 16-bit patchify kernels do.
__global__ void __launch_bounds__(256)
patchify_f32_kernel(const float* __restrict__ x, float* __restrict__ ap, int H, int W, int p, int L, int P, int* step_ctr) {
    const int ph = blockIdx.x, b = blockIdx.y;
    const int Wp = W / p, pp = p * p;
    if (step_ctr != nullptr && ph == 0 && b == 0 && threadIdx.x == 0) {
        const int nxt = step_ctr[1];
        step_ctr[0] = nxt;
        step_ctr[1] = nxt + 1;
    }
    const size_t row0 = static_cast<size_t>(b) * L + 1 + static_cast<size_t>(ph) * Wp;
    for (int i = threadIdx.x; i < Wp * P; i += blockDim.x) {
        const int pw = i / P, kc = i - pw * P;
        const int c = kc / pp, rem = kc - c * pp;
        const int p1 = rem / p, p2 = rem - p1 * p;
        ap[(row0 + pw) * P + kc] = x[((static_cast<size_t>(b) * 3 + c) * H + ph * p + p1) * W + pw * p + p2];
    }
    if (ph == 0)
        for (int i = threadIdx.x; i < P; i += blockDim.x) ap[static_cast<size_t>(b) * L * P + i] = 0.f;
}

// ---------------------------------------------------------------------------- SGEMM with fused epilogues
//   acc[m, n] = sum_k A[m, k] * Wt[n, k]        A: [M, K] row-major, Wt: [N, K] row-major (nn.Linear / flattened conv weight)
//   F_BIAS   out = acc + bias                                   qkv, head   (ViT.py:107, 213)
//   F_GELU   out = gelu_erf(acc + bias)                         fc1         (ViT.py:85-86)
//   F_RESID  out += acc + bias   (in place on the residual)     proj, fc2   (ViT.py:136-137)
//   F_EMBED  out = acc + posb[m % L] + time[t(m / L)]           patch embed (ViT.py:154,202-205); posb holds cls / conv bias
// 128 x 128 tile per block of 256 threads, 8 x 8 outputs per thread, K in steps of 8 through double-buffered shared memory.
enum { F_BIAS = 0, F_GELU = 1, F_RESID = 2, F_EMBED = 3 };

struct SgemmEmbed {
    const float* posb;       // [>= L, N]
    const float* time_tab;   // [T, N]
    const int* t_img;        // [B] per-image t, or null -> tbl[*cur].t
    const StepEntry* tbl;
    const int* cur;
    int L, T;
};

template <int EPI>
__global__ void __launch_bounds__(256)
sgemm_f32_kernel(const float* __restrict__ A, const float* __restrict__ Wt, const float* __restrict__ bias, float* __restrict__ out,
                 int M, int N, int K, const SgemmEmbed e) {
    __shared__ float As[2][8][128 + 4];
    __shared__ float Bs[2][8][128 + 4];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * 128, n0 = blockIdx.x * 128;
    const int lrow = tid >> 1, lk = (tid & 1) * 4;          // this thread stages A[m0 + lrow][k0 + lk .. +4] and the same of Wt
    const int ty = tid >> 4, tx = tid & 15;                 // outputs rows ty*8.., cols tx*8..
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    const bool a_ok = m0 + lrow < M, b_ok = n0 + lrow < N;
    const float* ap = A + static_cast<size_t>(a_ok ? m0 + lrow : 0) * K + lk;
    const float* bp = Wt + static_cast<size_t>(b_ok ? n0 + lrow : 0) * K + lk;
    auto stage = [&](int buf, int k0) {
        const float4 av = a_ok ? *reinterpret_cast<const float4*>(ap + k0) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 bv = b_ok ? *reinterpret_cast<const float4*>(bp + k0) : make_float4(0.f, 0.f, 0.f, 0.f);
        As[buf][lk][lrow] = av.x; As[buf][lk + 1][lrow] = av.y; As[buf][lk + 2][lrow] = av.z; As[buf][lk + 3][lrow] = av.w;
        Bs[buf][lk][lrow] = bv.x; Bs[buf][lk + 1][lrow] = bv.y; Bs[buf][lk + 2][lrow] = bv.z; Bs[buf][lk + 3][lrow] = bv.w;
    };
    stage(0, 0);
    __syncthreads();
    const int nk = K / 8;
    for (int kt = 0; kt < nk; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < nk) stage(buf ^ 1, (kt + 1) * 8);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
            float a[8], b[8];
            const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 8]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 8 + 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 8]);
            const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 8 + 4]);
            a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
            b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + ty * 8 + i;
        if (m >= M) continue;
        const float* trow = nullptr;
        const float* prow = nullptr;
        if (EPI == F_EMBED) {
            const int b = m / e.L;
            int t = e.t_img != nullptr ? e.t_img[b] : e.tbl[*e.cur].t;
            t = t < 0 ? 0 : (t >= e.T ? e.T - 1 : t);
            trow = e.time_tab + static_cast<size_t>(t) * N;
            prow = e.posb + static_cast<size_t>(m - b * e.L) * N;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int n = n0 + tx * 8 + j;
            if (n >= N) continue;
            float v = acc[i][j];
            float* o = out + static_cast<size_t>(m) * N + n;
            if (EPI == F_EMBED) v += prow[n] + trow[n];
            else v += bias[n];
            if (EPI == F_GELU) v = 0.5f * v * (1.0f + erff(v * 0.70710678118654752440f));
            if (EPI == F_RESID) v += *o;
            *o = v;
        }
    }
}

// ---------------------------------------------------------------------------- LayerNorm, fp32, two passes (ViT.py:124,128,182)
// one warp per row; biased variance around the mean, eps inside the square root
__global__ void __launch_bounds__(256)
layernorm_f32_kernel(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ b, float* __restrict__ y,
                     int M, int C, float eps) {
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= M) return;
    const float* xr = x + static_cast<size_t>(row) * C;
    float s = 0.f;
    for (int i = lane; i < C; i += 32) s += xr[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / static_cast<float>(C);
    float v = 0.f;
    for (int i = lane; i < C; i += 32) {
        const float d = xr[i] - mean;
        v = fmaf(d, d, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const float rstd = 1.0f / sqrtf(v / static_cast<float>(C) + eps);
    float* yr = y + static_cast<size_t>(row) * C;
    for (int i = lane; i < C; i += 32) yr[i] = (xr[i] - mean) * rstd * g[i] + b[i];
}

// ---------------------------------------------------------------------------- attention, fp32 (ViT.py:107-114)
// One block per (image, head): K and V of the head in shared memory (row pitch D + 1: conflict-free for both access
// patterns), one warp per query row at a time.  scores: lane k owns keys k, k + 32, ...; softmax with expf; P*V: lane e owns
// output dimension e (and e + 32), the probabilities pass through a per-warp shared-memory row.
template <int D>
__global__ void __launch_bounds__(256)
attention_f32_kernel(const float* __restrict__ qkv, float* __restrict__ out, int L, int C, float scale) {
    extern __shared__ float att_sm[];
    const int h = blockIdx.x, img = blockIdx.y;
    const int PITCH = D + 1;
    float* Ks = att_sm;
    float* Vs = Ks + static_cast<size_t>(L) * PITCH;
    float* Pr = Vs + static_cast<size_t>(L) * PITCH;           // [8 warps][L]
    const float* base = qkv + static_cast<size_t>(img) * L * 3 * C + h * D;
    for (int i = threadIdx.x; i < L * D; i += blockDim.x) {
        const int r = i / D, c = i - r * D;
        Ks[r * PITCH + c] = base[static_cast<size_t>(r) * 3 * C + C + c];
        Vs[r * PITCH + c] = base[static_cast<size_t>(r) * 3 * C + 2 * C + c];
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* pr = Pr + static_cast<size_t>(warp) * L;
    for (int r = warp; r < L; r += 8) {
        float q[D];
        const float* qg = base + static_cast<size_t>(r) * 3 * C;
#pragma unroll
        for (int i = 0; i < D; ++i) q[i] = qg[i];
        float mx = -INFINITY;
        for (int k = lane; k < L; k += 32) {
            float acc = 0.f;
#pragma unroll
            for (int i = 0; i < D; ++i) acc = fmaf(q[i], Ks[k * PITCH + i], acc);
            acc *= scale;
            pr[k] = acc;
            mx = fmaxf(mx, acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
        for (int k = lane; k < L; k += 32) {
            const float pv = expf(pr[k] - mx);
            pr[k] = pv;
            sum += pv;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        __syncwarp();
        float a0 = 0.f, a1 = 0.f;
        for (int k = 0; k < L; ++k) {
            const float pv = pr[k];
            a0 = fmaf(pv, Vs[k * PITCH + lane], a0);
            if (D == 64) a1 = fmaf(pv, Vs[k * PITCH + lane + 32], a1);
        }
        const float inv = 1.0f / sum;
        float* orow = out + (static_cast<size_t>(img) * L + r) * C + h * D;
        orow[lane] = a0 * inv;
        if (D == 64) orow[lane + 32] = a1 * inv;
        __syncwarp();
    }
}

inline size_t attention_f32_smem(int L, int D) { return (static_cast<size_t>(2) * L * (D + 1) + 8 * static_cast<size_t>(L)) * sizeof(float); }

}  // namespace ddc

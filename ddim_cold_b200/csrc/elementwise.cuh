// HBM-bound passes either side of the GEMM chain, plus the one-off weight packing kernels.
#pragma once

#include "common.cuh"

namespace ddc {

// ---------------------------------------------------------------------------- patchify (im2col)
// Conv2d(kernel = stride = p) lowered to a GEMM operand (ViT.py:150,154):
//   Ap[b*L + 1 + ph*Wp + pw, c*p*p + p1*p + p2] = bf16(x[b, c, ph*p + p1, pw*p + p2])
// Row b*L (the cls slot, ViT.py:202-203) is all zeros so that the patch GEMM's row index equals
// the token row and its epilogue can add cls/pos/time without any remapping.  Columns
// [3p^2, Kp) are zero padding up to the 64-element K block.
// grid (Hp, B).  Block (0,0) thread 0 also advances the device-side step counter that every
// later kernel of the step reads: cur <- next, next <- next + 1.
template <bool F16>
__global__ void __launch_bounds__(256)
patchify_kernel(const float* __restrict__ x, bf16* __restrict__ ap, int H, int W, int p, int L, int Kp,
                int* step_ctr) {
    const int ph = blockIdx.x, b = blockIdx.y;
    const int Wp = W / p;
    const int pp = p * p;
    if (step_ctr != nullptr && ph == 0 && b == 0 && threadIdx.x == 0) {
        const int nxt = step_ctr[1];
        step_ctr[0] = nxt;
        step_ctr[1] = nxt + 1;
    }
    const size_t row0 = static_cast<size_t>(b) * L + 1 + static_cast<size_t>(ph) * Wp;
    const int half = Kp >> 1;   // bf16 pairs per row
    for (int i = threadIdx.x; i < Wp * half; i += blockDim.x) {
        const int pw = i / half;
        const int kc = (i - pw * half) * 2;
        float v0 = 0.f, v1 = 0.f;
        if (kc < 3 * pp) {      // p is even for every shape here, so a pair never straddles a row
            const int c = kc / pp;
            const int rem = kc - c * pp;
            const int p1 = rem / p, p2 = rem - p1 * p;
            const float2 v = *reinterpret_cast<const float2*>(
                x + ((static_cast<size_t>(b) * 3 + c) * H + ph * p + p1) * W + pw * p + p2);
            v0 = v.x;
            v1 = v.y;
        }
        *reinterpret_cast<uint32_t*>(ap + (row0 + pw) * Kp + kc) = pack2<F16>(v0, v1);
    }
    if (ph == 0) {
        for (int i = threadIdx.x; i < half; i += blockDim.x)
            *reinterpret_cast<uint32_t*>(ap + static_cast<size_t>(b) * L * Kp + 2 * i) = 0u;
    }
}

// ---------------------------------------------------------------------------- unpatchify + clamp + DDIM
// One pass over the image state (ViT.py:214-217, 229, 233-234):
//   x0[b,c,ph*p+p1,pw*p+p2] = y[b*L + 1 + ph*Wp + pw, (p1*p+p2)*3 + c]      (channel fastest)
//   MODE_FORWARD : out = x0                        (forward(): no clamp)
//   MODE_STEP    : x0 = clamp(x0,-1,1);  x <- a*x + b*x0 in place  (a, b from the step table);
//                  optional taps: x0 of this step, eps = (x - sqrt_at*x0) * inv_sqrt_1mat,
//                  and the result buffer on the flagged last step.
// grid (Hp, B); the block stages its Wp head rows in shared memory (row stride P+1) so that
// both the y reads and the image reads / writes are contiguous.
enum { MODE_FORWARD = 0, MODE_STEP = 1 };

struct HeadDdimParams {
    const float* y;     // [rows, ldy] fp32 head output
    int ldy;
    float* x;           // [B,3,H,W] state (MODE_STEP) or output (MODE_FORWARD)
    int H, W, p, L;
    const StepEntry* tbl;   // device step table indexed by *cur, or null -> use `one`
    const int* cur;
    StepEntry one;
    float* x0_last;     // [B,3,H,W] or null
    float* x0_seq;      // [steps,B,3,H,W] or null
    float* eps_seq;     // [steps,B,3,H,W] or null
    int B;
};

template <int MODE>
__global__ void __launch_bounds__(256) head_ddim_kernel(const HeadDdimParams q) {
    extern __shared__ float ys[];
    const int ph = blockIdx.x, b = blockIdx.y;
    const int p = q.p, W = q.W, H = q.H;
    const int Wp = W / p;
    const int P = 3 * p * p;
    const int PS = P + 1;
    const float* yrow = q.y + (static_cast<size_t>(b) * q.L + 1 + static_cast<size_t>(ph) * Wp) * q.ldy;
    for (int i = threadIdx.x; i < Wp * P; i += blockDim.x) {
        const int pw = i / P, f = i - pw * P;
        ys[pw * PS + f] = yrow[static_cast<size_t>(pw) * q.ldy + f];
    }
    __syncthreads();
    StepEntry e;
    int step = 0;
    if (MODE == MODE_STEP) {
        if (q.tbl != nullptr) {
            step = *q.cur;
            e = q.tbl[step];
        } else {
            e = q.one;
        }
    }
    const size_t img_elems = static_cast<size_t>(3) * H * W;
    const size_t seq_off = static_cast<size_t>(step) * q.B * img_elems;
    for (int i = threadIdx.x; i < 3 * p * W; i += blockDim.x) {
        const int c = i / (p * W);
        const int rem = i - c * p * W;
        const int p1 = rem / W, xx = rem - p1 * W;
        const int pw = xx / p, p2 = xx - pw * p;
        float v = ys[pw * PS + (p1 * p + p2) * 3 + c];
        const size_t gi = (static_cast<size_t>(b) * 3 + c) * H * W + static_cast<size_t>(ph * p + p1) * W + xx;
        if (MODE == MODE_FORWARD) {
            q.x[gi] = v;
        } else {
            v = fminf(fmaxf(v, -1.f), 1.f);
            const float xt = q.x[gi];
            q.x[gi] = e.a * xt + e.b * v;
            if (q.x0_seq) q.x0_seq[seq_off + gi] = v;
            if (q.eps_seq) q.eps_seq[seq_off + gi] = (xt - e.sqrt_at * v) * e.inv_sqrt_1mat;
            if ((e.flags & 1) && q.x0_last) q.x0_last[gi] = v;
        }
    }
}

// ---------------------------------------------------------------------------- weight packing (one-off)
// fp32 [rows, cols] -> 16-bit operand format [rows, cols_pad], zero padded
template <bool F16>
__global__ void pack_16bit_kernel(const float* __restrict__ src, bf16* __restrict__ dst, int rows, int cols,
                                  int cols_pad) {
    const size_t n = static_cast<size_t>(rows) * cols_pad;
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const int r = static_cast<int>(i / cols_pad), c = static_cast<int>(i - static_cast<size_t>(r) * cols_pad);
        const float v = c < cols ? src[static_cast<size_t>(r) * cols + c] : 0.f;
        reinterpret_cast<uint16_t*>(dst)[i] = static_cast<uint16_t>(pack2<F16>(v, 0.f) & 0xFFFFu);
    }
}

// posb[0] = cls + pos[0];  posb[l>0] = pos[l] + conv_bias   (ViT.py:150,202-205)
__global__ void pack_posb_kernel(const float* __restrict__ pos, const float* __restrict__ cls,
                                 const float* __restrict__ conv_bias, float* __restrict__ posb, int L, int C) {
    const int n = L * C;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int l = i / C, c = i - l * C;
        posb[i] = pos[i] + (l == 0 ? cls[c] : conv_bias[c]);
    }
}

__global__ void copy_f32_kernel(const float* __restrict__ src, float* __restrict__ dst, size_t n) {
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<size_t>(gridDim.x) * blockDim.x)
        dst[i] = src[i];
}

}  // namespace ddc

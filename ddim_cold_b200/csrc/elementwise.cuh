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
    pdl_launch_dependents();
    pdl_wait();                 // reads x, writes Ap and the step counter: all after the previous kernel
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
    // MODE_STEP, vector kernel only: also emit the NEXT step's patch matrix (what patchify would make of the updated x:
    // 16-bit [B*L, Kp], column c*p*p + p1*p + p2), so that the loop runs patchify only once, and advance the step counter
    // the way the next step's patchify would have (ctr[0] = ctr[1]; ctr[1] += 1; ctr[2] is the block ticket).
    bf16* ap_next;      // or null
    int Kp, f16;
    int* ctr_adv;       // or null
};

template <int MODE>
__global__ void __launch_bounds__(256) head_ddim_kernel(const HeadDdimParams q) {
    extern __shared__ float ys[];
    pdl_launch_dependents();
    pdl_wait();
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

// ---------------------------------------------------------------------------- row-group variants (vector access)
// The two kernels above touch 4 bytes per thread per access and spend most of their instructions on index
// arithmetic.  When W, p and ldy are multiples of 4 the variants below are used instead: a block owns R whole patch
// rows of one image, moves the image in float4 lines (fully coalesced), and exchanges the patch-major / image-major
// orders through shared memory, so the GEMM-side rows (Ap, Y) are read and written as contiguous 16-byte vectors
// too.  Divisions by runtime shape values are replaced by multiply-high with host-made reciprocals.
struct FastDiv {
    uint32_t d, m;      // m = floor(2^32 / d) + 1; exact for n * d < 2^32 (n < 2^21, d < 2^11 here)
};
inline FastDiv make_fastdiv(int d) {
    FastDiv f;
    f.d = static_cast<uint32_t>(d);
    f.m = d > 1 ? static_cast<uint32_t>((1ull << 32) / static_cast<uint32_t>(d)) + 1u : 0u;
    return f;
}
__device__ __forceinline__ uint32_t fdiv(uint32_t n, const FastDiv f) { return f.d == 1 ? n : __umulhi(n, f.m); }

struct RowGroup {
    int R;                      // patch rows per block
    FastDiv w4, rp, p, p4;      // W/4, R*p, p, P/4
};

// largest divisor R of Hp whose staging tile fits `budget` bytes; 0 if even one patch row does not
inline int rows_per_block(int Hp, size_t bytes_per_patch_row, size_t budget) {
    int best = 0;
    for (int r = 1; r <= Hp; ++r)
        if (Hp % r == 0 && static_cast<size_t>(r) * bytes_per_patch_row <= budget) best = r;
    return best;
}

template <bool F16>
__global__ void __launch_bounds__(256)
patchify_rows_kernel(const float* __restrict__ x, bf16* __restrict__ ap, int H, int W, int p, int L, int Kp,
                     const RowGroup rg, int* step_ctr) {
    extern __shared__ __align__(16) uint8_t pr_tile[];      // [R*Wp][Kp] 16-bit, exactly the block's Ap rows
    pdl_launch_dependents();
    pdl_wait();                 // reads x, writes Ap and the step counter: all after the previous kernel
    const int b = blockIdx.y, ph0 = blockIdx.x * rg.R;
    const int Wp = W / p, pp = p * p, R = rg.R;
    if (step_ctr != nullptr && blockIdx.x == 0 && b == 0 && threadIdx.x == 0) {
        const int nxt = step_ctr[1];
        step_ctr[0] = nxt;
        step_ctr[1] = nxt + 1;
    }
    const int rows = R * Wp;
    const int padw = (Kp - 3 * pp) >> 1;                     // zero K padding, in 32-bit words per row
    for (int i = threadIdx.x; i < rows * padw; i += blockDim.x) {
        const int r = i / padw, j = i - r * padw;
        *reinterpret_cast<uint32_t*>(pr_tile + (static_cast<size_t>(r) * Kp + 3 * pp) * 2 + 4 * j) = 0u;
    }
    const uint32_t w4 = rg.w4.d;
    const uint32_t nitems = 3u * R * p * w4;
    const float* xb = x + static_cast<size_t>(b) * 3 * H * W;
    for (uint32_t i = threadIdx.x; i < nitems; i += blockDim.x) {
        const uint32_t line = fdiv(i, rg.w4), x4 = i - line * w4;
        const uint32_t c = fdiv(line, rg.rp), yy = line - c * rg.rp.d;
        const uint32_t r = fdiv(yy, rg.p), p1 = yy - r * p;
        const uint32_t xx = x4 * 4, pw = fdiv(xx, rg.p), p2 = xx - pw * p;
        const float4 v = *reinterpret_cast<const float4*>(xb + (static_cast<size_t>(c) * H + ph0 * p + yy) * W + xx);
        uint2 o;
        o.x = pack2<F16>(v.x, v.y);
        o.y = pack2<F16>(v.z, v.w);
        *reinterpret_cast<uint2*>(pr_tile + (static_cast<size_t>(r * Wp + pw) * Kp + c * pp + p1 * p + p2) * 2) = o;
    }
    __syncthreads();
    uint4* dst = reinterpret_cast<uint4*>(ap + (static_cast<size_t>(b) * L + 1 + static_cast<size_t>(ph0) * Wp) * Kp);
    const int nvec = rows * Kp / 8;
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = reinterpret_cast<const uint4*>(pr_tile)[i];
    if (blockIdx.x == 0) {                                   // the cls slot's all-zero row
        uint4* z = reinterpret_cast<uint4*>(ap + static_cast<size_t>(b) * L * Kp);
        for (int i = threadIdx.x; i < Kp / 8; i += blockDim.x) z[i] = make_uint4(0, 0, 0, 0);
    }
}

template <int MODE>
__global__ void __launch_bounds__(256) head_ddim_rows_kernel(const HeadDdimParams q, const RowGroup rg) {
    extern __shared__ float ys[];
    pdl_launch_dependents();
    pdl_wait();
    const int b = blockIdx.y, ph0 = blockIdx.x * rg.R;
    const int p = q.p, W = q.W, H = q.H, R = rg.R;
    const int Wp = W / p;
    const int P = 3 * p * p, PS = P + 1;
    const float* yrow = q.y + (static_cast<size_t>(b) * q.L + 1 + static_cast<size_t>(ph0) * Wp) * q.ldy;
    // the next step's patch-matrix rows of this block, behind the y tile (16-byte aligned)
    uint8_t* apt = reinterpret_cast<uint8_t*>(ys) + ((static_cast<size_t>(R) * Wp * PS * sizeof(float) + 15) & ~static_cast<size_t>(15));
    if (MODE == MODE_STEP && q.ap_next != nullptr) {
        const int padw = (q.Kp - P) >> 1;                    // zero K padding, in 32-bit words per row
        for (int i = threadIdx.x; i < R * Wp * padw; i += blockDim.x) {
            const int rr = i / padw, j = i - rr * padw;
            *reinterpret_cast<uint32_t*>(apt + (static_cast<size_t>(rr) * q.Kp + P) * 2 + 4 * j) = 0u;
        }
    }
    const uint32_t p4 = rg.p4.d;
    const uint32_t nld = static_cast<uint32_t>(R * Wp) * p4;
    for (uint32_t i = threadIdx.x; i < nld; i += blockDim.x) {
        const uint32_t row = fdiv(i, rg.p4), f4 = i - row * p4;
        const float4 v = *reinterpret_cast<const float4*>(yrow + static_cast<size_t>(row) * q.ldy + 4 * f4);
        float* d = ys + row * PS + 4 * f4;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
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
    const uint32_t w4 = rg.w4.d;
    const uint32_t nitems = 3u * R * p * w4;
    for (uint32_t i = threadIdx.x; i < nitems; i += blockDim.x) {
        const uint32_t line = fdiv(i, rg.w4), x4 = i - line * w4;
        const uint32_t c = fdiv(line, rg.rp), yy = line - c * rg.rp.d;
        const uint32_t r = fdiv(yy, rg.p), p1 = yy - r * p;
        const uint32_t xx = x4 * 4, pw = fdiv(xx, rg.p), p2 = xx - pw * p;
        const float* src = ys + (r * Wp + pw) * PS + (p1 * p + p2) * 3 + c;
        float4 v = make_float4(src[0], src[3], src[6], src[9]);
        const size_t gi = (static_cast<size_t>(b) * 3 + c) * H * W + static_cast<size_t>(ph0 * p + yy) * W + xx;
        if (MODE == MODE_FORWARD) {
            *reinterpret_cast<float4*>(q.x + gi) = v;
        } else {
            v.x = fminf(fmaxf(v.x, -1.f), 1.f); v.y = fminf(fmaxf(v.y, -1.f), 1.f);
            v.z = fminf(fmaxf(v.z, -1.f), 1.f); v.w = fminf(fmaxf(v.w, -1.f), 1.f);
            const float4 xt = *reinterpret_cast<const float4*>(q.x + gi);
            const float4 xn = make_float4(e.a * xt.x + e.b * v.x, e.a * xt.y + e.b * v.y, e.a * xt.z + e.b * v.z, e.a * xt.w + e.b * v.w);
            *reinterpret_cast<float4*>(q.x + gi) = xn;
            if (q.x0_seq) *reinterpret_cast<float4*>(q.x0_seq + seq_off + gi) = v;
            if (q.eps_seq)
                *reinterpret_cast<float4*>(q.eps_seq + seq_off + gi) =
                    make_float4((xt.x - e.sqrt_at * v.x) * e.inv_sqrt_1mat, (xt.y - e.sqrt_at * v.y) * e.inv_sqrt_1mat,
                                (xt.z - e.sqrt_at * v.z) * e.inv_sqrt_1mat, (xt.w - e.sqrt_at * v.w) * e.inv_sqrt_1mat);
            if ((e.flags & 1) && q.x0_last) *reinterpret_cast<float4*>(q.x0_last + gi) = v;
            if (q.ap_next != nullptr) {          // the updated pixels, in the patch matrix's layout and operand format
                uint2 o;
                if (q.f16) { o.x = pack2<true>(xn.x, xn.y); o.y = pack2<true>(xn.z, xn.w); }
                else { o.x = pack2<false>(xn.x, xn.y); o.y = pack2<false>(xn.z, xn.w); }
                *reinterpret_cast<uint2*>(apt + (static_cast<size_t>(r * Wp + pw) * q.Kp + c * p * p + p1 * p + p2) * 2) = o;
            }
        }
    }
    if (MODE == MODE_STEP && q.ap_next != nullptr) {
        __syncthreads();
        uint4* dst = reinterpret_cast<uint4*>(q.ap_next + (static_cast<size_t>(b) * q.L + 1 + static_cast<size_t>(ph0) * Wp) * q.Kp);
        const int nvec = R * Wp * q.Kp / 8;
        for (int i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = reinterpret_cast<const uint4*>(apt)[i];
    }
    if (MODE == MODE_STEP && q.ctr_adv != nullptr) {
        // every block read the step index at its start; the block that finishes LAST moves the counter on
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned total = gridDim.x * gridDim.y;
            if (atomicAdd(reinterpret_cast<unsigned*>(q.ctr_adv + 2), 1u) == total - 1) {
                volatile int* cv = q.ctr_adv;
                const int nxt = cv[1];
                cv[0] = nxt;
                cv[1] = nxt + 1;
                cv[2] = 0;
                __threadfence();
            }
        }
    }
}

constexpr size_t EW_TILE_BUDGET = 40 * 1024;     // staging tile per block (several blocks per SM stay resident)

inline RowGroup make_rowgroup(int R, int W, int p) {
    RowGroup rg;
    rg.R = R;
    rg.w4 = make_fastdiv(W / 4);
    rg.rp = make_fastdiv(R * p);
    rg.p = make_fastdiv(p);
    rg.p4 = make_fastdiv(3 * p * p / 4);
    return rg;
}

// Launchers: vector variant when the shape allows it, else the scalar kernels above.
template <bool F16>
inline cudaError_t launch_patchify(const float* x, bf16* ap, int B, int H, int W, int p, int L, int Kp, int* step_ctr,
                                   cudaStream_t st) {
    const int Hp = H / p, Wp = W / p;
    const bool aligned = (reinterpret_cast<uintptr_t>(x) % 16 == 0) && (reinterpret_cast<uintptr_t>(ap) % 16 == 0);
    const int R = (W % 4 == 0 && p % 4 == 0 && Kp % 8 == 0 && aligned)
                      ? rows_per_block(Hp, static_cast<size_t>(Wp) * Kp * 2, EW_TILE_BUDGET) : 0;
    if (R > 0)
        return launch_chain(patchify_rows_kernel<F16>, dim3(Hp / R, B), dim3(256), static_cast<size_t>(R) * Wp * Kp * 2, st, 1,
                            x, ap, H, W, p, L, Kp, make_rowgroup(R, W, p), step_ctr);
    return launch_chain(patchify_kernel<F16>, dim3(Hp, B), dim3(256), 0, st, 1, x, ap, H, W, p, L, Kp, step_ctr);
}

// Patch rows per block of the vector head_ddim kernel (0 = the shape needs the scalar kernel).  With `with_ap` the block
// also holds its rows of the next step's patch matrix.
inline int head_ddim_rows(const HeadDdimParams& q, bool with_ap) {
    const int Hp = q.H / q.p, Wp = q.W / q.p;
    const size_t row_bytes = static_cast<size_t>(Wp) * (3 * q.p * q.p + 1) * sizeof(float) +
                             (with_ap ? static_cast<size_t>(Wp) * q.Kp * 2 + 16 : 0);
    auto al16 = [](const void* ptr) { return reinterpret_cast<uintptr_t>(ptr) % 16 == 0; };
    const bool aligned = al16(q.y) && al16(q.x) && al16(q.x0_last) && al16(q.x0_seq) && al16(q.eps_seq) && al16(q.ap_next);
    if (with_ap && (q.Kp % 8 != 0 || (3 * q.p * q.p) % 2 != 0)) return 0;
    return (q.W % 4 == 0 && q.p % 4 == 0 && q.ldy % 4 == 0 && aligned) ? rows_per_block(Hp, row_bytes, EW_TILE_BUDGET) : 0;
}

template <int MODE>
inline cudaError_t launch_head_ddim(const HeadDdimParams& q, cudaStream_t st) {
    const int Hp = q.H / q.p, Wp = q.W / q.p;
    const size_t y_bytes = static_cast<size_t>(Wp) * (3 * q.p * q.p + 1) * sizeof(float);
    const bool with_ap = MODE == MODE_STEP && q.ap_next != nullptr;
    const int R = head_ddim_rows(q, with_ap);
    if (R > 0) {
        const size_t smem = ((R * y_bytes + 15) & ~static_cast<size_t>(15)) + (with_ap ? static_cast<size_t>(R) * Wp * q.Kp * 2 : 0);
        return launch_chain(head_ddim_rows_kernel<MODE>, dim3(Hp / R, q.B), dim3(256), smem, st, 1, q, make_rowgroup(R, q.W, q.p));
    }
    if (with_ap) return cudaErrorInvalidValue;          // the caller checks head_ddim_rows() before asking for the patch matrix
    return launch_chain(head_ddim_kernel<MODE>, dim3(Hp, q.B), dim3(256), y_bytes, st, 1, q);
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

// posb[0] = cls + pos[0];  posb[l>0] = pos[l] + conv_bias   (ViT.py:150,202-205), written for `rows` >= L table rows
// with row i holding position i % L: token rows m0 .. m0+127 of any 128-row GEMM tile are then the contiguous table
// rows (m0 % L) .. (m0 % L) + 127, one TMA box.
__global__ void pack_posb_kernel(const float* __restrict__ pos, const float* __restrict__ cls,
                                 const float* __restrict__ conv_bias, float* __restrict__ posb, int L, int C, int rows) {
    const int n = rows * C;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int r = i / C, c = i - r * C;
        const int l = r % L;
        posb[i] = pos[l * C + c] + (l == 0 ? cls[c] : conv_bias[c]);
    }
}

// out = (in + add) * mul, float4 lanes.  The samplers return (x0 + 1) / 2 (ViT.py:237): one pass on the device
// before the copy back instead of a 50 MB pass on the host.  Not on the step chain (plain launch, no PDL).
__global__ void __launch_bounds__(256) affine_kernel(const float* __restrict__ in, float* __restrict__ out, size_t n,
                                                     float add, float mul) {
    const size_t n4 = n >> 2;
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
        const float4 v = reinterpret_cast<const float4*>(in)[i];
        reinterpret_cast<float4*>(out)[i] = make_float4((v.x + add) * mul, (v.y + add) * mul, (v.z + add) * mul, (v.w + add) * mul);
    }
    for (size_t i = (n4 << 2) + blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += stride)
        out[i] = (in[i] + add) * mul;
}

// out[i] = a * z[i] + b * draft[i % period]: the draft->drawing start state x = N(0,1) * sqrt(1 - alpha) + draft * sqrt(alpha)
// (ViT_draft2drawing.py:396) for a batch of noise draws and one draft (period = 3*H*W) or one draft per image (period = n).
// Separate roundings for the two products and the sum, like the reference's tensor expression (no FMA contraction).
__global__ void __launch_bounds__(256) mix_kernel(const float* __restrict__ z, const float* __restrict__ draft, float* __restrict__ out,
                                                  size_t n, size_t period, float a, float b) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += stride)
        out[i] = __fadd_rn(__fmul_rn(z[i], a), __fmul_rn(draft[i % period], b));
}

// mean over n elements of smooth_l1(pred - target) with beta = 1 (F.smooth_l1_loss defaults; the trainer's evaluate(),
// multi_gpu_trainer.py:42): 0.5 d^2 for |d| < 1, |d| - 0.5 otherwise.  One pass over both tensors (float4 lanes), a
// block sum per CTA into scratch[0 .. gridDim.x), and the CTA that finishes last adds the block sums in index order in
// double precision -- the result does not depend on the order in which CTAs finish.  scratch[1023] is the ticket
// counter: zero before the first call, left at zero by every call.
__global__ void __launch_bounds__(256) smooth_l1_kernel(const float* __restrict__ pred, const float* __restrict__ target, size_t n,
                                                        float* __restrict__ scratch, float* __restrict__ out) {
    const size_t n4 = n >> 2;
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    auto term = [](float a, float b) {
        const float d = fabsf(a - b);
        return d < 1.f ? 0.5f * d * d : d - 0.5f;
    };
    float acc = 0.f;
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
        const float4 a = reinterpret_cast<const float4*>(pred)[i], b = reinterpret_cast<const float4*>(target)[i];
        acc += (term(a.x, b.x) + term(a.y, b.y)) + (term(a.z, b.z) + term(a.w, b.w));
    }
    for (size_t i = (n4 << 2) + blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += stride)
        acc += term(pred[i], target[i]);
    __shared__ float warp_sum[8];
    __shared__ bool last;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float b = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) b += warp_sum[w];
        scratch[blockIdx.x] = b;
        __threadfence();
        unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + 1023);
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        double t = 0.0;
        for (unsigned int i = 0; i < gridDim.x; ++i) t += static_cast<double>(reinterpret_cast<volatile float*>(scratch)[i]);
        out[0] = static_cast<float>(t / static_cast<double>(n));
        *reinterpret_cast<unsigned int*>(scratch + 1023) = 0u;
    }
}

__global__ void copy_f32_kernel(const float* __restrict__ src, float* __restrict__ dst, size_t n) {
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<size_t>(gridDim.x) * blockDim.x)
        dst[i] = src[i];
}

}  // namespace ddc

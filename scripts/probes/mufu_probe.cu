// Probe: MUFU throughput per SM (ex2.approx.f32, ex2.approx.f16x2, rcp.approx.f32) with 4 / 8 / 16 warps per SM.
// Each thread runs 8 independent dependency chains; results per clock per SM = warps * 32 * ops / cycles.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, long long* cyc, int iters) {
    float a[8];
    unsigned h[8];
    for (int i = 0; i < 8; ++i) { a[i] = -0.001f * (threadIdx.x + i + 1); h[i] = 0x3800b800u + threadIdx.x + i; }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[i]));
            if (MODE == 2) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            if (MODE == 3) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i]));
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += a[i] + __uint_as_float(h[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    const int iters = 4096;
    const char* names[4] = {"ex2.f32", "ex2.f16x2 (instructions; 2 results each)", "rcp.f32", "fma.f32"};
    for (int mode = 0; mode < 4; ++mode)
        for (int warps = 4; warps <= 16; warps *= 2) {
            for (int rep = 0; rep < 2; ++rep) {
                if (mode == 0) k<0><<<148, warps * 32>>>(out, cyc, iters);
                if (mode == 1) k<1><<<148, warps * 32>>>(out, cyc, iters);
                if (mode == 2) k<2><<<148, warps * 32>>>(out, cyc, iters);
                if (mode == 3) k<3><<<148, warps * 32>>>(out, cyc, iters);
                cudaDeviceSynchronize();
            }
            long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
            printf("%-45s warps/SM %2d: %.2f instr-lanes / clk / SM\n", names[mode], warps, (double)warps * 32 * 8 * iters / c);
        }
    return 0;
}

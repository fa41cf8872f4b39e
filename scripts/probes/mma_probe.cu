// Micro-benchmark: back-to-back tcgen05.mma (M=128, N, K=16, SS) cycles per instruction.
#include "../../ddim_cold_b200/csrc/common.cuh"
#include <cstdio>
using namespace ddc;
__global__ void __launch_bounds__(128, 1) probe(int n, int iters, int kblocks, long long* out) {
    extern __shared__ uint8_t raw[];
    __shared__ uint32_t slot;
    __shared__ uint64_t bar;
    uint32_t ra = smem_u32(raw);
    uint8_t* sm = raw + (((ra + 1023u) & ~1023u) - ra);
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (96 + 32) * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = 0x3c003c00u;
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    fence_proxy_async_smem();
    tc_fence_before(); __syncthreads(); tc_fence_after();
    uint32_t tb = __shfl_sync(0xffffffffu, slot, 0);
    if (warp == 1) {
        const uint32_t idesc = umma_idesc_16bit(128, n, false);
        const uint64_t ad0 = umma_desc_sw128(smem_u32(sm)), bd0 = umma_desc_sw128(smem_u32(sm) + 96 * 1024);
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            for (int kb = 0; kb < kblocks; ++kb) {
                if (elect_one()) umma_kblock(tb + (it & 1) * 256, ad0 + kb * 1024, bd0, idesc, kb != 0);
                __syncwarp();
            }
        }
        if (elect_one()) umma_commit(&bar);
        __syncwarp();
        mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (lane == 0 && blockIdx.x == 0) out[0] = t1 - t0;
    }
    tc_fence_before(); __syncthreads();
    if (warp == 0) tmem_dealloc(tb, 512);
}
int main() {
    long long* d; cudaMalloc(&d, 8);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int grid : {1, 148}) for (int n : {64, 128, 192, 256}) {
        int iters = 200, kb = 6;
        probe<<<grid, 128, 130 * 1024, 0>>>(n, iters, kb, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long c = 0; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
        printf("grid=%d N=%d: %s  %.1f cycles per MMA (K=16)  -> %.0f flop/cycle/SM\n", grid, n, cudaGetErrorString(e),
               double(c) / (iters * kb * 4), 2.0 * 128 * n * 16 / (double(c) / (iters * kb * 4)));
    }
    return 0;
}

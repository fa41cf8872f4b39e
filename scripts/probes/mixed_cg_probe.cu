// Probe: may one kernel issue tcgen05.mma.cta_group::2 (pair GEMM) and tcgen05.mma.cta_group::1 (per-CTA MMAs) into
// tensor memory that was allocated with cta_group::2?  Every operand tile holds one constant, so the expected
// accumulators are exact small integers whatever the swizzle.  mode 0: both kinds in flight at once; mode 1: the
// per-CTA MMAs start after the pair MMAs have completed.
#include "../../ddim_cold_b200/csrc/common.cuh"
#include <cstdio>
#include <cuda_fp16.h>
using namespace ddc;

__global__ void __launch_bounds__(256, 1) probe(int iters, int mode, int* out) {
    extern __shared__ uint8_t raw[];
    __shared__ uint32_t slot;
    __shared__ uint64_t bar_pair, bar_local;
    const uint32_t ra = smem_u32(raw);
    uint8_t* sm = raw + (((ra + 1023u) & ~1023u) - ra);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t r = cluster_ctarank();
    const float a = r == 0 ? 1.f : 2.f, bp = r == 0 ? 1.f : 3.f, bl = r == 0 ? 5.f : 7.f;
    __half* A = reinterpret_cast<__half*>(sm);                 // [128][64]
    __half* Bp = reinterpret_cast<__half*>(sm + 16384);        // [32][64]   this CTA's half of the pair B operand
    __half* Bl = reinterpret_cast<__half*>(sm + 16384 + 4096); // [64][64]   per-CTA B operand
    for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) A[i] = __float2half(a);
    for (int i = threadIdx.x; i < 32 * 64; i += blockDim.x) Bp[i] = __float2half(bp);
    for (int i = threadIdx.x; i < 64 * 64; i += blockDim.x) Bl[i] = __float2half(bl);
    if (warp == 0) { tmem_alloc<2>(&slot, 128); tmem_relinquish<2>(); }
    if (threadIdx.x == 0) { mbar_init(&bar_pair, 1); mbar_init(&bar_local, 1); fence_barrier_init(); }
    fence_proxy_async_smem();
    tc_fence_before(); __syncthreads(); cluster_sync_all(); tc_fence_after();
    const uint32_t tb = slot;
    if (warp == 1 && r == 0) {
        const uint32_t idesc = umma_idesc_16bit(256, 64, true);
        const uint64_t ad = umma_desc_sw128(smem_u32(A)), bd = umma_desc_sw128(smem_u32(Bp));
        for (int it = 0; it < iters; ++it) {
            if (elect_one()) umma_kblock<2>(tb, ad, bd, idesc, it != 0);
            __syncwarp();
            if ((it & 7) == 3) __nanosleep(200);
        }
        if (elect_one()) umma_commit_pair(&bar_pair);
        __syncwarp();
    }
    if (warp == 2) {
        if (mode == 1) mbar_wait(&bar_pair, 0);
        tc_fence_after();
        const uint32_t idesc = umma_idesc_16bit(128, 64, true);
        const uint64_t ad = umma_desc_sw128(smem_u32(A)), bd = umma_desc_sw128(smem_u32(Bl));
        for (int it = 0; it < iters; ++it) {
            if (elect_one()) umma_kblock<1>(tb + 64, ad, bd, idesc, it != 0);
            __syncwarp();
            if ((it & 7) == 5) __nanosleep(150);
        }
        if (elect_one()) umma_commit(&bar_local);
        __syncwarp();
    }
    if (warp >= 4) {
        mbar_wait(&bar_pair, 0);
        mbar_wait(&bar_local, 0);
        tc_fence_after();
        const int q = warp & 3;
        int bad = 0;
        for (int c = 0; c < 4; ++c) {
            float v[32];
            tmem_ld32(tb + (static_cast<uint32_t>(q * 32) << 16) + 32 * c, v);
            for (int j = 0; j < 32; ++j) {
                const int col = 32 * c + j;
                const float want = col < 32 ? iters * 64.f * a * 1.f : (col < 64 ? iters * 64.f * a * 3.f : iters * 64.f * a * bl);
                if (v[j] != want) {
                    ++bad;
                    if (bad == 1 && lane == 0) printf("cta %u warp %d col %d: got %f want %f\n", r, warp, col, v[j], want);
                }
            }
        }
        atomicAdd(&out[r], bad);
    }
    tc_fence_before(); __syncthreads(); cluster_sync_all();
    if (warp == 0) tmem_dealloc<2>(tb, 128);
}

int main() {
    int* d; cudaMalloc(&d, 8);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int mode : {1, 0}) for (int grid : {2, 148}) {
        cudaMemset(d, 0, 8);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 40 * 1024;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, probe, 50, mode, d);
        cudaError_t e2 = cudaDeviceSynchronize();
        int h[2] = {-1, -1}; cudaMemcpy(h, d, 8, cudaMemcpyDeviceToHost);
        printf("mode=%d grid=%d: launch %s, sync %s, mismatches cta0=%d cta1=%d\n", mode, grid, cudaGetErrorString(e), cudaGetErrorString(e2), h[0], h[1]);
        if (e2 != cudaSuccess) return 1;
    }
    return 0;
}

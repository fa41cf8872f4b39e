// Probe: throughput of warp reads of tensor memory (tcgen05.ld 32x32b.x32 = 4 KB per warp instruction) with 1, 2 and 4
// warps per TMEM lane quarter, alone and under a stream of tcgen05.mma (M = 128, N = 256) writing other columns.
#include "../../ddim_cold_b200/csrc/gemm2.cuh"
#include <cstdio>
using namespace ddc;

#define LD32REGS(NAME, SHAPE) \
__device__ __forceinline__ void NAME(uint32_t taddr, float* v) { \
    asm volatile("tcgen05.ld.sync.aligned." SHAPE ".b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\ttcgen05.wait::ld.sync.aligned;" \
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]), "=f"(v[8]), "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15]), "=f"(v[16]), "=f"(v[17]), "=f"(v[18]), "=f"(v[19]), "=f"(v[20]), "=f"(v[21]), "=f"(v[22]), "=f"(v[23]), "=f"(v[24]), "=f"(v[25]), "=f"(v[26]), "=f"(v[27]), "=f"(v[28]), "=f"(v[29]), "=f"(v[30]), "=f"(v[31]) : "r"(taddr) : "memory"); }
LD32REGS(ld_16x256b_x8, "16x256b.x8")
LD32REGS(ld_16x128b_x16, "16x128b.x16")
LD32REGS(ld_16x64b_x32, "16x64b.x32")

__global__ void __launch_bounds__(640, 1) probe(int warps_per_quarter, int with_mma, int iters, long long* out, int shape) {
    extern __shared__ uint8_t raw[];
    __shared__ uint32_t slot;
    __shared__ uint64_t bar;
    __shared__ int stop;
    const uint32_t ra = smem_u32(raw);
    uint8_t* sm = raw + (((ra + 1023u) & ~1023u) - ra);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = 0x3c003c00u;
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); stop = 0; }
    fence_proxy_async_smem();
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tb = slot;
    if (warp == 1 && with_mma) {
        // MMAs into columns [256, 512) until the readers are done
        const uint32_t idesc = umma_idesc_16bit(128, 256, true);
        const uint64_t ad = umma_desc_sw128(smem_u32(sm)), bd = umma_desc_sw128(smem_u32(sm) + 16 * 1024);
        long long n = 0;
        while (!*reinterpret_cast<volatile int*>(&stop)) {
            for (int k = 0; k < 8; ++k) {
                if (elect_one()) umma_kblock<1>(tb + 256, ad, bd, idesc, 1u);
                __syncwarp();
                ++n;
            }
            if (elect_one()) umma_commit(&bar);
            __syncwarp();
            mbar_wait(&bar, static_cast<uint32_t>((n / 8 - 1) & 1));
        }
        if (lane == 0) out[2] = n * 4;                       // MMAs (K = 16 each) executed meanwhile
    }
    if (warp >= 4 && (warp - 4) / 4 < warps_per_quarter) {
        const int q = warp & 3;
        const uint32_t t = tb + (static_cast<uint32_t>(q * 32) << 16);
        float v[32], acc = 0.f;
        named_bar_sync(1, 128 * warps_per_quarter);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const uint32_t a = t + 32 * (it & 7);             // columns [0, 256)
            if (shape == 0) tmem_ld32(a, v);
            else if (shape == 1) ld_16x256b_x8(a + ((it & 8) ? (16u << 16) : 0u), v);      // 16 lanes x 64 columns
            else if (shape == 2) ld_16x128b_x16(a + ((it & 8) ? (16u << 16) : 0u), v);
            else ld_16x64b_x32(a + ((it & 8) ? (16u << 16) : 0u), v);
            acc += v[0] + v[17] + v[31];
        }
        const long long t1 = clock64();
        named_bar_sync(1, 128 * warps_per_quarter);
        if (warp == 4 && lane == 0) { out[0] = t1 - t0; stop = 1; }
        if (acc == 12345.f) out[1] = 1;
    }
    tc_fence_before(); __syncthreads();
    if (warp == 0) tmem_dealloc(tb, 512);
}

int main() {
    long long* d; cudaMalloc(&d, 32);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
    const int iters = 2000;
    for (int shape : {0, 1, 2, 3}) for (int mma : {0, 1}) for (int w : {1, 2, 4}) {
        if (shape > 0 && (mma || w != 1)) continue;
        cudaMemset(d, 0, 32);
        probe<<<1, 640, 66 * 1024>>>(w, mma, iters, d, shape);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[4] = {0, 0, 0, 0}; cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
        const double bytes = 4096.0 * 4 * w * iters;          // 4 quarters x w warps x 4 KB per load
        printf("shape=%d (0: 32x32b.x32, 1: 16x256b.x8, 2: 16x128b.x16, 3: 16x64b.x32) warps/quarter=%d mma=%d: %s  %.0f cycles per load per warp, %.1f B/clk per SM", shape, w, mma, cudaGetErrorString(e),
               double(h[0]) / iters, bytes / double(h[0]));
        if (mma) printf(", %.1f cycles per MMA (N = 256; 128 alone)", double(h[0]) / double(h[2] > 0 ? h[2] : 1));
        printf("\n");
        if (e != cudaSuccess) return 1;
    }
    return 0;
}

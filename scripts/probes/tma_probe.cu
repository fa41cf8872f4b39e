// Micro-benchmark: sustained TMA load bandwidth L2 -> shared memory, all SMs, L2-resident source.
#include "../../ddim_cold_b200/csrc/ddimcold.cu"
#include <cstdio>
__global__ void __launch_bounds__(128, 1) tma_stream(const __grid_constant__ CUtensorMap tm, int box_rows, int stages,
                                                     int iters, int src_rows, long long* out) {
    extern __shared__ uint8_t raw[];
    __shared__ uint64_t bars[32];
    uint32_t ra = smem_u32(raw);
    uint8_t* sm = raw + (((ra + 1023u) & ~1023u) - ra);
    const uint32_t bytes = box_rows * 128;
    if (threadIdx.x == 0) { for (int i = 0; i < stages; ++i) mbar_init(&bars[i], 1); fence_barrier_init(); }
    __syncthreads();
    if (threadIdx.x < 32) {
        long long t0 = clock64();
        uint32_t phase = 0; int stage = 0;
        const int nbox = src_rows / box_rows;
        for (int it = 0; it < iters + stages; ++it) {
            if (it >= stages) mbar_wait(&bars[stage], phase);           // oldest load has landed
            if (it < iters && elect_one()) {
                mbar_arrive_expect_tx(&bars[stage], bytes);
                tma_load_2d(sm + stage * bytes, &tm, &bars[stage], ((it + blockIdx.x) % 6) * 64, ((it * 7 + blockIdx.x) % nbox) * box_rows);
            }
            __syncwarp();
            if (++stage == stages) { stage = 0; if (it >= stages) phase ^= 1; }
        }
        long long t1 = clock64();
        if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    }
}
int main() {
    const int rows = 1152, K = 384;
    void* W; cudaMalloc(&W, (size_t)rows * K * 2); cudaMemset(W, 0, (size_t)rows * K * 2);
    long long* d; cudaMalloc(&d, 148 * 8);
    cudaFuncSetAttribute(tma_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    for (int box_rows : {64, 96, 128, 192}) for (int stages : {2, 4, 8, 16}) {
        if ((size_t)box_rows * 128 * stages > 200 * 1024) continue;
        CUtensorMap tm;
        if (make_tmap(&tm, false, W, rows, K, box_rows)) { printf("%s\n", ddc_last_error()); return 1; }
        const int iters = 2000;
        tma_stream<<<148, 128, 210 * 1024>>>(tm, box_rows, stages, iters, rows, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
        printf("box %3d rows (%5d B) stages %2d: %s  %.1f cycles/load  %.1f B/clk/SM  (in flight %d KB)\n", box_rows, box_rows * 128,
               stages, cudaGetErrorString(e), avg / iters, (double)box_rows * 128 * iters / avg, box_rows * 128 * stages / 1024);
    }
    return 0;
}

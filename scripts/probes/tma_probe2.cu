// Micro-benchmark 2: what serialises TMA loads? (per instruction vs per iteration, descriptor location, #issuers)
#include "../../ddim_cold_b200/csrc/ddimcold.cu"
#include <cstdio>
// mode 0: 1 load / iteration, param tensormap.  mode 1: 2 loads / iteration (half box each).  mode 2: tensormap in global
// memory.  mode 3: no wait on completion at all (pure issue rate, all into the same stage/barrier with huge tx count).
__global__ void __launch_bounds__(128, 1) tma_stream(const __grid_constant__ CUtensorMap tm, const CUtensorMap* gtm, int box_rows,
                                                     int stages, int iters, int src_rows, int mode, long long* out) {
    extern __shared__ uint8_t raw[];
    __shared__ uint64_t bars[32];
    uint32_t ra = smem_u32(raw);
    uint8_t* sm = raw + (((ra + 1023u) & ~1023u) - ra);
    const uint32_t bytes = box_rows * 128;
    if (threadIdx.x == 0) { for (int i = 0; i < 32; ++i) mbar_init(&bars[i], 1); fence_barrier_init(); }
    __syncthreads();
    const int warp = threadIdx.x >> 5;
    const int nissue = mode == 4 ? 2 : 1;           // mode 4: two warps issue independently
    if (warp < nissue) {
        const CUtensorMap* map = mode == 2 ? gtm : &tm;
        uint64_t* mybars = bars + warp * 16;
        uint8_t* mysm = sm + warp * 100 * 1024;
        long long t0 = clock64();
        uint32_t phase = 0; int stage = 0;
        const int nbox = src_rows / box_rows;
        int rb = blockIdx.x % nbox, kc = 0;
        for (int it = 0; it < iters + stages; ++it) {
            if (mode != 3 && it >= stages) mbar_wait(&mybars[stage], phase);
            if (++rb >= nbox) rb = 0;
            if (++kc >= 5) kc = 0;
            if (it < iters && elect_one()) {
                const int r = rb * box_rows;
                if (mode == 3) {
                    tma_load_2d(mysm + stage * bytes, map, &mybars[0], kc * 64, r);
                } else if (mode == 1) {
                    mbar_arrive_expect_tx(&mybars[stage], 2 * bytes);
                    tma_load_2d(mysm + stage * 2 * bytes, map, &mybars[stage], kc * 64, r);
                    tma_load_2d(mysm + stage * 2 * bytes + bytes, map, &mybars[stage], (kc + 1) * 64, r);
                } else {
                    mbar_arrive_expect_tx(&mybars[stage], bytes);
                    tma_load_2d(mysm + stage * bytes, map, &mybars[stage], kc * 64, r);
                }
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
    CUtensorMap* gtm; cudaMalloc(&gtm, sizeof(CUtensorMap));
    cudaFuncSetAttribute(tma_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const char* names[5] = {"1 load/iter", "2 loads/iter", "tensormap in gmem", "issue only (no waits)", "2 issuing warps"};
    for (int grid : {148, 1}) for (int mode = 0; mode < 5; ++mode) for (int box_rows : {8, 64, 192}) {
        const int stages = 4;
        CUtensorMap tm;
        if (make_tmap(&tm, false, W, rows, K, box_rows)) { printf("%s\n", ddc_last_error()); return 1; }
        cudaMemcpy(gtm, &tm, sizeof(tm), cudaMemcpyHostToDevice);
        const int iters = 2000;
        tma_stream<<<grid, 128, 210 * 1024>>>(tm, gtm, box_rows, stages, iters, rows, mode, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < grid; ++i) avg += h[i]; avg /= grid;
        printf("grid %3d  %-22s box %3d rows: %s  %.1f cycles/iter\n", grid, names[mode], box_rows, cudaGetErrorString(e), avg / iters);
    }
    return 0;
}

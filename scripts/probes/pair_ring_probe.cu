// Probe: throughput of the W-tile ring exactly as the pair GEMM drives it, minus the MMAs:
// both CTAs of a cluster issue cta_group::2 TMA loads crediting the leader's full barrier; the leader's
// "consumer" thread waits for the stage and releases it to both producers (local + remote mbarrier arrive).
#include "../../ddim_cold_b200/csrc/ddimcold.cu"
#include <cstdio>
__device__ __forceinline__ void mbar_arrive_peer(uint64_t* bar, uint32_t peer_rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(bar)), "r"(peer_rank));
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
ring(const __grid_constant__ CUtensorMap tm, int box_rows, int stages, int iters, int producers, int mode, long long* out) {
    extern __shared__ uint8_t raw[];
    __shared__ uint64_t full[16], empty[16];
    uint32_t ra = smem_u32(raw);
    uint8_t* sm = raw + (((ra + 1023u) & ~1023u) - ra);
    const uint32_t bytes = box_rows * 128;
    const uint32_t crank = cluster_ctarank();
    if (threadIdx.x == 0) {
        for (int i = 0; i < 16; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        fence_barrier_init();
    }
    __syncthreads();
    cluster_sync_all();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    long long t0 = clock64();
    if (warp < producers && lane == 0) {                       // producers: alternate loads
        for (int it = warp; it < iters; it += producers) {
            const int st = it % stages; const uint32_t ph = (it / stages) & 1;
            mbar_wait(&empty[st], ph ^ 1);
            if (crank == 0) mbar_arrive_expect_tx(&full[st], 2 * bytes);
            tma_load_2d_pair(sm + st * bytes, &tm, &full[st], (it % 6) * 64, ((it * 3) % 8) * box_rows + crank * box_rows);
        }
    } else if (warp == 2 && lane == 0 && crank == 0) {         // consumer on the leader
        for (int it = 0; it < iters; ++it) {
            const int st = it % stages; const uint32_t ph = (it / stages) & 1;
            mbar_wait(&full[st], ph);
            if (mode == 0) { mbar_arrive(&empty[st]); mbar_arrive_peer(&empty[st], 1); }
            else umma_commit_pair(&empty[st]);              // the release path the GEMM uses (no MMAs pending here)
        }
        out[blockIdx.x / 2] = clock64() - t0;
    }
    __syncthreads();
    cluster_sync_all();
}
int main() {
    const int rows = 4096, K = 384;
    void* W; cudaMalloc(&W, (size_t)rows * K * 2); cudaMemset(W, 0, (size_t)rows * K * 2);
    long long* d; cudaMalloc(&d, 74 * 8);
    cudaFuncSetAttribute(ring, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    for (int mode : {1}) for (int producers : {1, 2}) for (int box_rows : {64, 128}) for (int stages : {2, 4, 8}) {
        if ((size_t)box_rows * 128 * stages > 200 * 1024) continue;
        CUtensorMap tm;
        if (make_tmap(&tm, false, W, rows, K, box_rows)) { printf("%s\n", ddc_last_error()); return 1; }
        const int iters = 3000;
        ring<<<148, 128, 210 * 1024>>>(tm, box_rows, stages, iters, producers, mode, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[74]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < 74; ++i) avg += h[i]; avg /= 74;
        printf("mode %d producers %d box %3d rows (%5d B/CTA) stages %d: %s %.0f cycles/stage  %.1f B/clk/SM\n", mode, producers, box_rows,
               box_rows * 128, stages, cudaGetErrorString(e), avg / iters, (double)box_rows * 128 * iters / avg);
    }
    return 0;
}

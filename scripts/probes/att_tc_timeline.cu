// Probe: SM-clock timeline of the tcgen05 attention kernel (S1 shape), softmax warp 4 of CTA 0, first items.
#define DDC_PROBE_TIMING 1
#include "../../ddim_cold_b200/csrc/ddimcold.cu"
#include <cstdio>
int main() {
    const int B = 1024, L = 65, C = 384, heads = 12;
    const size_t rows = (size_t)B * L + 128;
    bf16 *qkv, *out; long long* dbg;
    cudaMalloc(&qkv, rows * 3 * C * 2); cudaMalloc(&out, rows * C * 2); cudaMalloc(&dbg, 64 * 8);
    cudaMemset(qkv, 0, rows * 3 * C * 2); cudaMemset(dbg, 0, 64 * 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 3; ++rep) {
        long long* d = rep == 2 ? dbg : nullptr;
        cudaMemcpyToSymbol(g_atc_dbg, &d, sizeof(d));
        cudaEventRecord(e0);
        int rc = ddc_k_attention(0, qkv, out, B, L, C, heads, nullptr);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("rep %d rc=%d %s %.1f us\n", rep, rc, cudaGetErrorString(e), ms * 1e3);
    }
    long long h[64];
    cudaMemcpy(h, dbg, sizeof(h), cudaMemcpyDeviceToHost);
    const char* nm[8] = {"top", "full", "vt done", "s ready", "p written", "o ready", "stored", "LOAD ISSUED"};
    for (int it = 0; it < 6; ++it) {
        printf("item %d:", it);
        for (int k = 0; k < 8; ++k) printf("  %s=%lld", nm[k], h[it * 8 + k] - h[0]);
        printf("\n");
    }
    return 0;
}

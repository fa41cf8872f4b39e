// Probe: timeline of the tcgen05 attention kernel (CTA 0).  Build: nvcc -DDDC_FA_TRACE ... (see scripts/probes/README).
#include "../../ddim_cold_b200/csrc/ddimcold.cu"
#include <vector>
int main(int argc, char** argv) {
    const int B = argc > 1 ? atoi(argv[1]) : 256, L = argc > 2 ? atoi(argv[2]) : 626, C = argc > 3 ? atoi(argv[3]) : 384, H = argc > 4 ? atoi(argv[4]) : 12;
    size_t nq = (size_t)B * L * 3 * C;
    std::vector<__half> h(nq);
    for (size_t i = 0; i < nq; ++i) h[i] = __float2half(((i * 2654435761u) % 2001) / 1000.0f - 1.0f);
    void *qkv, *out;
    cudaMalloc(&qkv, nq * 2); cudaMalloc(&out, (size_t)B * L * C * 2);
    cudaMemcpy(qkv, h.data(), nq * 2, cudaMemcpyHostToDevice);
    for (int r = 0; r < 3; ++r) {
        int rc = ddc_k_attention(1, qkv, out, B, L, C, H, 0);
        if (rc) { printf("err %s\n", ddc_last_error()); return 1; }
        cudaDeviceSynchronize();
    }
    static long long t[3 * 64 * 8];
    cudaMemcpyFromSymbol(t, ddc::g_fa_trace, sizeof(t));
    long long t0 = t[0];
    const char* names[3] = {"mma", "wg0", "wg1"};
    for (int role = 0; role < 3; ++role) {
        printf("== %s (cycles since first stamp; columns = stamps 0..6)\n", names[role]);
        for (int i = 0; i < 24; ++i) {
            printf("%2d:", i);
            for (int k = 0; k < 7; ++k) printf(" %8lld", t[(role * 64 + i) * 8 + k] ? t[(role * 64 + i) * 8 + k] - t0 : 0);
            printf("\n");
        }
    }
    return 0;
}

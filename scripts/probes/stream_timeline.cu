// Probe: SM-clock timeline of the streaming GEMM (qkv shape: M = 66560, N = 1152, K = 384) for CTA 0 of the first pair.
#define DDC_PROBE_TIMING 1
#include "../../ddim_cold_b200/csrc/ddimcold.cu"
#include <cstdio>
int main(int argc, char** argv) {
    const int M = 66560, C = 384, N = 3 * C, K = C, cs = 2;
    const size_t Mp = (M + 127) / 128 * 128;
    void *A, *W, *OUT; float* bias; long long* dbg;
    cudaMalloc(&A, Mp * K * 2); cudaMalloc(&W, (size_t)N * K * 2); cudaMalloc(&OUT, Mp * N * 2); cudaMalloc(&bias, N * 4);
    cudaMalloc(&dbg, 2 * 4 * 64 * 8);
    cudaMemset(A, 0, Mp * K * 2); cudaMemset(W, 0, (size_t)N * K * 2); cudaMemset(bias, 0, N * 4);
    const int bn = pick_bn_stream(N);
    CUtensorMap tmA, tmB, tmO;
    if (make_tmap(&tmA, false, A, Mp, K, G2_BM) || make_tmap(&tmB, false, W, N, K, bn / cs) || make_tmap(&tmO, false, OUT, M, N, G2_BM)) {
        printf("tmap: %s\n", ddc_last_error()); return 1;
    }
    Gemm2Params p{};
    p.M = M; p.N = N; p.K = K; p.bn = bn; p.stages = pick_stages2(K, N, bn / cs, false); p.bias = bias;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 3; ++rep) {
        cudaMemset(dbg, 0, 2 * 4 * 64 * 8);
        p.dbg = rep == 2 ? dbg : nullptr;
        cudaEventRecord(e0);
        int rc = launch_gemm2<S_BF16>(tmA, tmB, tmO, tmA, tmO, p, false, cs, 0);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("rep %d rc=%d %s  %.1f us  (bn=%d stages=%d)\n", rep, rc, cudaGetErrorString(e), ms * 1e3, bn, p.stages);
    }
    static long long h[2 * 4 * 64];
    cudaMemcpy(h, dbg, sizeof(h), cudaMemcpyDeviceToHost);
    const long long t0 = h[(0 * 4 + 1) * 64 + 0];
    for (int t = 0; t < 4; ++t) {
        printf("tile %d\n  MMA  [begin,end] per N tile:", t);
        for (int nt = 0; nt < 8; ++nt) { const long long* v = &h[(0 * 4 + 1) * 64 + (t * 8 + nt) * 2]; if (v[0]) printf("  %lld-%lld", v[0] - t0, v[1] - t0); }
        printf("\n  EPI  [begin,end] per N tile:");
        for (int nt = 0; nt < 8; ++nt) { const long long* v = &h[(0 * 4 + 2) * 64 + (t * 8 + nt) * 2]; if (v[0]) printf("  %lld-%lld", v[0] - t0, v[1] - t0); }
        printf("\n");
    }
    const char* nm[8] = {"loop/other", "wait tfull", "wait store-read", "bar (chunk begin)", "tmem_ld", "math+STS", "fence+bar", "store issue"};
    long long tot = 0;
    for (int i = 0; i < 8; ++i) tot += h[3 * 64 + i];
    for (int i = 0; i < 8; ++i) printf("  %-18s %8lld cycles  %5.1f %%\n", nm[i], h[3 * 64 + i], 100.0 * h[3 * 64 + i] / tot);
    return 0;
}

// Probe: per-role SM-clock timeline of the streaming GEMM (qkv shape) for the first CTA pair.
#include "../../ddim_cold_b200/csrc/ddimcold.cu"
#include <cstdio>
int main(int argc, char** argv) {
    const int M = 66560, N = 1152, K = 384;
    const int cs = argc > 1 ? atoi(argv[1]) : 2;
    void *A, *W, *O; float* bias; long long* dbg;
    const size_t Mp = (M + 127) / 128 * 128;
    cudaMalloc(&A, Mp * K * 2); cudaMalloc(&W, (size_t)N * K * 2); cudaMalloc(&O, (size_t)M * N * 2);
    cudaMalloc(&bias, N * 4); cudaMalloc(&dbg, 2 * 4 * 64 * 8);
    cudaMemset(A, 0, Mp * K * 2); cudaMemset(W, 0, (size_t)N * K * 2); cudaMemset(bias, 0, N * 4);
    const int bn = pick_bn_stream(N);
    CUtensorMap tmA, tmB, tmO;
    if (make_tmap_operand(&tmA, A, Mp, K, G2_BM, pick_gk(K)) || make_tmap_operand(&tmB, W, N, K, bn / cs, pick_gk(K)) || make_tmap(&tmO, false, O, M, N, 32)) {
        printf("tmap: %s\n", ddc_last_error()); return 1;
    }
    Gemm2Params p{};
    p.M = M; p.N = N; p.K = K; p.bn = bn; p.stages = pick_stages2(K, N, bn / cs, false); p.bias = bias;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 3; ++rep) {
        cudaMemset(dbg, 0, 2 * 4 * 64 * 8);
        p.dbg = rep == 2 ? dbg : nullptr;
        cudaEventRecord(e0);
        int rc = launch_gemm2<S_BF16>(tmA, tmB, tmO, tmA, tmA, p, false, cs, 0);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("rep %d rc=%d %s  %.1f us  (bn=%d stages=%d cs=%d)\n", rep, rc, cudaGetErrorString(e), ms * 1e3, bn, p.stages, cs);
    }
    static long long h[2 * 4 * 64];
    cudaMemcpy(h, dbg, sizeof(h), cudaMemcpyDeviceToHost);
    for (int cta = 0; cta < 2; ++cta) {
        long long t0 = h[(cta * 4 + 0) * 64 + 0];
        printf("== CTA %d (cycles since its first producer stamp)\n", cta);
        const char* names[3] = {"producer: tile start / per N-tile", "mma: [begin,end] per N-tile", "epilogue w4: [begin,end] per N-tile"};
        for (int role = 0; role < 3; ++role) {
            printf("  %s\n   ", names[role]);
            for (int i = 0; i < 64; ++i) {
                long long v = h[(cta * 4 + role) * 64 + i];
                if (v) printf(" %d:%lld", i, v - t0);
            }
            printf("\n");
        }
    }
    return 0;
}

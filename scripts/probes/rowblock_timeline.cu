// Probe: per-role SM-clock timeline of the row-block GEMM (proj / fc2 shape) for the first CTA pair.
#include "../../ddim_cold_b200/csrc/ddimcold.cu"
#include <cstdio>
int main(int argc, char** argv) {
    const int flags = argc > 1 ? atoi(argv[1]) : 0;
    const int M = 66560, C = 384, K = 384, cs = 2;
    const size_t Mp = (M + 127) / 128 * 128;
    void *A, *W, *XN; float *bias, *X, *g, *b; long long* dbg;
    cudaMalloc(&A, Mp * K * 2); cudaMalloc(&W, (size_t)C * K * 2); cudaMalloc(&XN, Mp * C * 2);
    cudaMalloc(&X, Mp * C * 4); cudaMalloc(&bias, C * 4); cudaMalloc(&g, C * 4); cudaMalloc(&b, C * 4);
    cudaMalloc(&dbg, 2 * 4 * 64 * 8);
    cudaMemset(A, 0, Mp * K * 2); cudaMemset(W, 0, (size_t)C * K * 2); cudaMemset(X, 0, Mp * C * 4);
    cudaMemset(bias, 0, C * 4); cudaMemset(g, 0, C * 4); cudaMemset(b, 0, C * 4);
    const int bn = pick_bn_row(C);
    CUtensorMap tmA, tmB, tmO, tmR, tmN;
    if (make_tmap(&tmA, false, A, Mp, K, G2_BM) || make_tmap(&tmB, false, W, C, K, bn / cs) || make_tmap(&tmO, true, X, M, C, G2_BM) ||
        make_tmap(&tmR, true, X, Mp, C, G2_BM) || make_tmap(&tmN, false, XN, M, C, G2_BM)) { printf("tmap: %s\n", ddc_last_error()); return 1; }
    Gemm2Params p{};
    p.M = M; p.N = C; p.K = K; p.bn = bn; p.stages = pick_stages2(K, C, bn / cs, true); p.bias = bias;
    p.ln_g = g; p.ln_b = b; p.ln_eps = 1e-5f; p.dbg_flags = flags;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 3; ++rep) {
        cudaMemset(dbg, 0, 2 * 4 * 64 * 8);
        p.dbg = rep == 2 ? dbg : nullptr;
        cudaEventRecord(e0);
        int rc = launch_gemm2<R_RESID>(tmA, tmB, tmO, tmR, tmN, p, false, cs, 0);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep) printf("flags %d rep %d rc=%d %s  %.1f us  (bn=%d stages=%d)\n", flags, rep, rc, cudaGetErrorString(e), ms * 1e3, bn, p.stages);
    }
    static long long h[2 * 4 * 64];
    cudaMemcpy(h, dbg, sizeof(h), cudaMemcpyDeviceToHost);
    long long t0 = h[(0 * 4 + 1) * 64 + 0];
    printf("CTA 0, cycles since first MMA stamp\n  mma [begin,end] per (tile, N-tile):");
    for (int i = 0; i < 64; ++i) { long long v = h[(0 * 4 + 1) * 64 + i]; if (v) printf(" %d:%lld", i, v - t0); }
    printf("\n  epilogue warp 4 per tile [pass1 begin, pass1 end, stats end, pass3 end]:");
    for (int i = 0; i < 64; ++i) { long long v = h[(0 * 4 + 2) * 64 + i]; if (v) printf(" %d:%lld", i, v - t0); }
    printf("\n");
    return 0;
}

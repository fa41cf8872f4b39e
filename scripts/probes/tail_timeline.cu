// Probe: SM-clock timeline of the fused block-tail kernel, CTA 0 (S1 shape).
#define DDC_PROBE_TIMING 1
#include "../../ddim_cold_b200/csrc/ddimcold.cu"
#include <cstdio>
int main() {
    const int M = 66560, C = 384;
    const size_t Mp = (M + 127) / 128 * 128;
    void *O, *Wp, *W1, *W2, *XN; float *X, *v; long long* dbg;
    cudaMalloc(&O, Mp * C * 2); cudaMalloc(&Wp, C * C * 2); cudaMalloc(&W1, C * C * 2); cudaMalloc(&W2, C * C * 2);
    cudaMalloc(&XN, Mp * C * 2); cudaMalloc(&X, Mp * C * 4); cudaMalloc(&v, C * 4); cudaMalloc(&dbg, 3 * 64 * 8);
    cudaMemset(O, 0, Mp * C * 2); cudaMemset(Wp, 0, C * C * 2); cudaMemset(W1, 0, C * C * 2); cudaMemset(W2, 0, C * C * 2);
    cudaMemset(X, 0, Mp * C * 4); cudaMemset(v, 0, C * 4);
    CUtensorMap tmO, tmWp, tmW1, tmW2, tmXld, tmXst, tmXN;
    if (make_tmap(&tmO, false, O, Mp, C, G2_BM) || make_tmap(&tmWp, false, Wp, C, C, 64) || make_tmap(&tmW1, false, W1, C, C, 64) ||
        make_tmap(&tmW2, false, W2, C, C, 64) || make_tmap(&tmXld, true, X, Mp, C, G2_BM) || make_tmap(&tmXst, true, X, M, C, G2_BM) ||
        make_tmap(&tmXN, false, XN, M, C, G2_BM)) { printf("%s\n", ddc_last_error()); return 1; }
    BlockTailParams bt{};
    bt.M = M; bt.C = C; bt.ln_eps = 1e-5f; bt.b_proj = bt.b_fc1 = bt.b_fc2 = v;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 3; ++rep) {
        cudaMemset(dbg, 0, 3 * 64 * 8);
        bt.dbg = rep == 2 ? dbg : nullptr;
        cudaEventRecord(e0);
        int rc = launch_block_tail_inst<false>(tmO, tmWp, tmW1, tmW2, tmXld, tmXst, tmXN, bt, 0);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("rep %d rc=%d %s %.1f us\n", rep, rc, cudaGetErrorString(e), ms * 1e3);
    }
    static long long h[3 * 64];
    cudaMemcpy(h, dbg, sizeof(h), cudaMemcpyDeviceToHost);
    long long t0 = h[0];
    const char* mn[16] = {"proj0", "projEnd", "ln2ok", "fc1(0)e", "fc2(0)b", "fc2(0)e", "", "fc1(1)e", "fc2(1)b", "fc2(1)e", "", "fc1(2)e", "fc2(2)b", "fc2(2)e", "", ""};
    const char* en[16] = {"epi1b", "pass1e", "ln2done", "e2(0)b", "e2(0)e", "e2(1)b", "e2(1)e", "e2(2)b", "e2(2)e", "epi3b", "Xdone", "epi3e", "", "", "", ""};
    for (int t = 0; t < 4; ++t) {
        printf("tile %d  MMA:", t);
        for (int i = 0; i < 16; ++i) if (h[t * 16 + i]) printf(" %s=%lld", mn[i], h[t * 16 + i] - t0);
        printf("\n        EPI:");
        for (int i = 0; i < 16; ++i) if (h[64 + t * 16 + i]) printf(" %s=%lld", en[i], h[64 + t * 16 + i] - t0);
        printf("  reswait=%lld\n", h[128 + t]);
    }
    return 0;
}

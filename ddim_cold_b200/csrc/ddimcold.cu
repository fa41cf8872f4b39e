// C-ABI implementation (include/ddimcold.h): plan, weight packing, the step schedule, the
// per-step kernel chain and its CUDA-graph replay.  Device memory always belongs to the
// caller; this file owns only host structs, TMA descriptors and graph objects.
#include "../../include/ddimcold.h"

#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "attention.cuh"
#include "common.cuh"
#include "elementwise.cuh"
#include "gemm2.cuh"
#include "block_tail.cuh"
#include "attention_fa.cuh"
#include "fp32_path.cuh"
#include "qkv_attn.cuh"
#include "qkv_attn_tc.cuh"

using namespace ddc;

static_assert(sizeof(ddc_step) == sizeof(StepEntry), "host/device step entry mismatch");

// ------------------------------------------------------------------------------- errors
static thread_local std::string g_err;

static int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CK(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess)                                                                          \
            return fail(DDC_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

#define TRY(call)            \
    do {                     \
        int r_ = (call);     \
        if (r_ != DDC_OK) return r_; \
    } while (0)

extern "C" const char* ddc_last_error(void) { return g_err.c_str(); }
extern "C" int ddc_version(void) { return 100; }

// ------------------------------------------------------------------------------- TMA descriptors
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int get_encode_fn(EncodeTiledFn* out) {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
        if (p == nullptr || qres != cudaDriverEntryPointSuccess)
            return fail(DDC_E_CUDA, "cuTensorMapEncodeTiled not available from this driver");
        fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    *out = fn;
    return DDC_OK;
}

// [rows, cols] row-major tensor, 128-byte swizzle; the box is 128 bytes wide (64 bf16 or 32 fp32)
static int make_tmap(CUtensorMap* m, bool f32, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                     uint64_t row_stride_elems = 0) {
    EncodeTiledFn enc;
    TRY(get_encode_fn(&enc));
    const uint32_t box_cols = f32 ? 32 : 64;
    const uint32_t esz = f32 ? 4 : 2;
    if (cols % box_cols != 0 || rows == 0 || box_rows == 0 || box_rows > 256 || (reinterpret_cast<uintptr_t>(base) & 0xF) != 0)
        return fail(DDC_E_INVALID, "tensor map: rows=%llu cols=%llu box_rows=%u base=%p", (unsigned long long)rows,
                    (unsigned long long)cols, box_rows, base);
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {(row_stride_elems ? row_stride_elems : cols) * esz};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                     const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(DDC_E_CUDA, "cuTensorMapEncodeTiled -> %d", static_cast<int>(r));
    return DDC_OK;
}

// ------------------------------------------------------------------------------- GEMM launch
static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}
static int pick_from(int N, const int* cands, int n) {
    for (int i = 0; i < n; ++i)
        if (N % cands[i] == 0) return cands[i];
    return 0;
}
static int pick_bn_stream(int N) {
    const int forced = env_int("DDC_BN_STREAM", 0);      // tuning aid
    if (forced > 0 && N % forced == 0 && forced % 64 == 0) return forced;
    const int c[] = {256, 192, 128, 64};
    return pick_from(N, c, 4);
}
static int pick_bn_f32(int N) { const int c[] = {256, 192, 128, 96, 64, 48, 32, 16}; return pick_from(N, c, 8); }
static int pick_bn_row(int N) { const int c[] = {128, 64}; return pick_from(N, c, 2); }

static int pick_stages2(int K, int N, int bn, bool rowblock) {
    const long budget = 227 * 1024 - static_cast<long>(gemm2_smem_bytes(K, N, bn, 0, rowblock));
    long s = budget / (static_cast<long>(bn) * 128);
    if (s > 8) s = 8;
    const int cap = env_int("DDC_MAX_STAGES", 0);          // tuning aid
    if (cap >= 2 && s > cap) s = cap;
    return static_cast<int>(s);
}

static int num_sms() {
    static int sms[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (sms[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        sms[dev] = v;
    }
    return sms[dev];
}

template <int EPI, bool F16, int CS>
static int launch_gemm2_inst(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmO, const CUtensorMap& tmR,
                             const CUtensorMap& tmN, const Gemm2Params& p, size_t smem, cudaStream_t st) {
    static int attr_dev_mask = 0;   // bit per device where the smem attribute has been raised
    int dev = 0;
    CK(cudaGetDevice(&dev));
    auto kern = gemm2_kernel<EPI, F16, CS>;
    if (!(attr_dev_mask & (1 << dev))) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_dev_mask |= (1 << dev);
    }
    const int sms = num_sms();
    // one CTA per M tile at most -- or, with the balanced unit split, per (M tile, N tile) unit
    const long work = p.balanced ? static_cast<long>(p.num_tiles) * (p.N / p.bn) : p.num_tiles;
    int grid = work < sms ? static_cast<int>(work) : sms;
    grid = grid / CS * CS;
    CK(launch_chain(kern, dim3(grid), dim3(G2_THREADS), smem, st, CS, tmA, tmB, tmO, tmR, tmN, p));
    return DDC_OK;
}

// Each CTA of a pair stages its own bn/cs rows of every W tile (tmB's box holds bn/cs rows); the pair's MMAs read both.
static int gemm_cluster_size(int M) {
    const int forced = env_int("DDC_CLUSTER", 0);        // tuning aid
    if (forced == 1 || forced == 2) return forced;
    return (M + G2_BM - 1) / G2_BM >= 2 ? 2 : 1;
}

template <int EPI>
static int launch_gemm2(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmO, const CUtensorMap& tmR,
                        const CUtensorMap& tmN, Gemm2Params p, bool f16, int cs, cudaStream_t st) {
    constexpr bool rowblock = (EPI == R_RESID || EPI == R_EMBED);
    const bool bn_ok = (EPI == S_F32) ? (p.bn % 16 == 0) : (rowblock ? (p.bn % 32 == 0) : (p.bn % 64 == 0));
    if (p.K % G2_BK != 0 || p.K > 512 || !bn_ok || p.bn > 256 || p.bn <= 0 || p.N % p.bn != 0 || p.stages < 2 ||
        p.stages > 8 || (rowblock && (p.N > 512 || p.N / p.bn > 4 || p.N % 64 != 0)) || (cs != 1 && cs != 2) ||
        (p.bn / cs) % 8 != 0)
        return fail(DDC_E_INVALID, "gemm: unsupported M=%d N=%d K=%d bn=%d stages=%d epi=%d cs=%d", p.M, p.N, p.K, p.bn,
                    p.stages, EPI, cs);
    const size_t smem = gemm2_smem_bytes(p.K, p.N, p.bn / cs, p.stages, rowblock);
    if (smem > 227 * 1024) return fail(DDC_E_INVALID, "gemm: %zu B of shared memory needed", smem);
    p.num_tiles = ((p.M + G2_BM - 1) / G2_BM + cs - 1) / cs * cs;
    p.balanced = (!rowblock && env_int("DDC_BALANCED", 1)) ? 1 : 0;
    p.dbg_flags |= env_int("DDC_GEMM_FLAGS", 0);          // tuning: 16 = no L2 eviction hints on the operand loads
    if (f16) {
        if (cs == 2) return launch_gemm2_inst<EPI, true, 2>(tmA, tmB, tmO, tmR, tmN, p, smem, st);
        return launch_gemm2_inst<EPI, true, 1>(tmA, tmB, tmO, tmR, tmN, p, smem, st);
    }
    if (cs == 2) return launch_gemm2_inst<EPI, false, 2>(tmA, tmB, tmO, tmR, tmN, p, smem, st);
    return launch_gemm2_inst<EPI, false, 1>(tmA, tmB, tmO, tmR, tmN, p, smem, st);
}

// ------------------------------------------------------------------------------- fused block tail launch
static bool block_tail_usable(int M, int C, int bn_row) {
    return C % 128 == 0 && C <= 384 && bn_row == 128 && (M + G2_BM - 1) / G2_BM >= 1 && !env_int("DDC_NO_FUSE", 0);
}

template <bool F16, int NG>
static int launch_block_tail_ng(const CUtensorMap& tmO, const CUtensorMap& tmWp, const CUtensorMap& tmW1,
                                  const CUtensorMap& tmW2, const CUtensorMap& tmXld, const CUtensorMap& tmXst,
                                  const CUtensorMap& tmXN, BlockTailParams p, cudaStream_t st) {
    static int attr_dev_mask = 0;
    int dev = 0;
    CK(cudaGetDevice(&dev));
    auto kern = block_tail_kernel<F16, NG>;
    if (!(attr_dev_mask & (1 << dev))) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_dev_mask |= (1 << dev);
    }
    // Two K blocks per ring stage: the stage handshake (full barrier -> MMA -> pair commit -> empty barrier) costs about
    // 500 cycles whatever it carries (scripts/probes/pair_ring_probe.cu), so 16 KB stages stream W twice as fast as 8 KB.
    p.gk = env_int("DDC_TAIL_GK", 2) == 1 ? 1 : 2;
    p.flags = env_int("DDC_TAIL_FLAGS", 2);                  // L2 eviction hints on (bit 1), next-tile L2 prefetch off (bit 0)
    p.stagger = env_int("DDC_TAIL_STAGGER", 115 * p.C);      // ~3/4 of a tile in SM cycles (measured ~58k per tile at C = 384)
    int stages = static_cast<int>((227 * 1024 - static_cast<long>(block_tail_smem_bytes(p.C, 0, 1))) / (p.gk * 64 * 128));
    if (stages > 8) stages = 8;
    if (stages < 2) return fail(DDC_E_INVALID, "block tail: C=%d leaves no room for the weight ring", p.C);
    p.stages = stages;
    p.num_tiles = ((p.M + G2_BM - 1) / G2_BM + 1) / 2 * 2;
    const int sms = num_sms();
    int grid = p.num_tiles < sms ? p.num_tiles : sms;
    grid = grid / 2 * 2;
    CK(launch_chain(kern, dim3(grid), dim3(bt_threads(NG)), block_tail_smem_bytes(p.C, stages, p.gk), st, 2, tmO, tmWp, tmW1, tmW2,
                    tmXld, tmXst, tmXN, p));
    return DDC_OK;
}

template <bool F16>
static int launch_block_tail_inst(const CUtensorMap& tmO, const CUtensorMap& tmWp, const CUtensorMap& tmW1,
                                  const CUtensorMap& tmW2, const CUtensorMap& tmXld, const CUtensorMap& tmXst,
                                  const CUtensorMap& tmXN, BlockTailParams p, cudaStream_t st) {
    // Two epilogue groups of four warps.  Sixteen epilogue warps measured SLOWER (129 vs 117 us): the passes are paced by
    // TMEM reads and per-chunk hand-shakes, not by warp count, and 640 threads leave 96 registers each.
    return launch_block_tail_ng<F16, 2>(tmO, tmWp, tmW1, tmW2, tmXld, tmXst, tmXN, p, st);
}

// ------------------------------------------------------------------------------- attention launch
static int att_pick_kc(int L) {
    const int cands[] = {128, 96, 80, 64, 48, 32};
    int best = 0, best_pad = 1 << 30;
    for (int c : cands) {
        const int pad = (L + c - 1) / c * c;
        if (pad < best_pad) { best_pad = pad; best = c; }
    }
    return best;
}

// tcgen05 attention for long sequences (attention_fa.cuh) plus, when L = 128 n + r with a few remainder rows, the
// CUDA-core kernel for those rows.  DDC_ATT_FA: 0 = never; 1 (default) = where it measures at least as fast as the
// mma.sync kernel: head_dim 32 and L > 128 (S3: 617 vs 602 us per launch; S2, head_dim 64, is 386 vs 330 us and stays
// on the mma.sync kernel); 2 = every shape; 3 = every L > 128.
static bool attention_fa_wanted(int L, int C, int D) {
    const int mode = env_int("DDC_ATT_FA", 1);
    if (!((D == 32 || D == 64) && C % 64 == 0)) return false;
    return mode == 2 || (mode == 3 && L > 128) || (mode == 1 && L > 128 && D == 32);
}

template <int D, bool F16>
static int launch_att_fa(const bf16* qkv, bf16* out, int B, int L, int C, int heads, cudaStream_t st) {
    static int attr_dev_mask = 0;
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (!(attr_dev_mask & (1 << dev))) {
        CK(cudaFuncSetAttribute(attention_fa_kernel<D, F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, FA_SMEM_BYTES));
        attr_dev_mask |= (1 << dev);
    }
    FaParams p{};
    p.L = L; p.C = C; p.n_hs = C / 64;
    fa_split_rows(L, &p.n_qt, &p.rows_main);
    p.n_qi = D == 64 ? (p.n_qt + 1) / 2 : p.n_qt;       // head_dim 64: two query tiles per iteration; 32: two heads
    p.n_items = B * p.n_hs * p.n_qi;
    p.dbg = env_int("DDC_FA_DBG", 0);
    p.n_kc = (L + FA_KCH - 1) / FA_KCH;
    p.scale_log2e = (1.0f / sqrtf(static_cast<float>(D))) * 1.44269504088896340736f;
    if (p.rows_main < L && L > FA_ROWS_MAXJ * 32) return fail(DDC_E_INVALID, "attention: L=%d too long for the remainder-row kernel", L);
    CUtensorMap tm;
    TRY(make_tmap(&tm, false, qkv, static_cast<uint64_t>(B) * L, static_cast<uint64_t>(3) * C, 128));
    int grid = num_sms();
    if (grid > p.n_items) grid = p.n_items;
    CK(launch_chain(attention_fa_kernel<D, F16>, dim3(grid), dim3(FA_THREADS), static_cast<size_t>(FA_SMEM_BYTES), st, 1, tm, out, p));
    if (p.rows_main < L) {
        const int n_tasks = B * heads * (L - p.rows_main);
        int blocks = (n_tasks + 7) / 8;
        if (blocks > 4 * num_sms()) blocks = 4 * num_sms();
        CK(launch_chain(attention_rows_kernel<D, F16>, dim3(blocks), dim3(256), static_cast<size_t>(0), st, 1, qkv, out, B, L, C, heads,
                        p.rows_main, p.scale_log2e));
    }
    CK(cudaGetLastError());
    return DDC_OK;
}

template <int D, int KC, bool F16>
static int launch_att(const bf16* qkv, bf16* out, int B, int L, int C, int heads, cudaStream_t st) {
    static int attr_dev_mask = 0;
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (!(attr_dev_mask & (1 << dev))) {
        CK(cudaFuncSetAttribute(attention_head_kernel<D, KC, F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        CK(cudaFuncSetAttribute(attention_stream_kernel<D, KC, F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_dev_mask |= (1 << dev);
    }
    const float scale_log2e = (1.0f / sqrtf(static_cast<float>(D))) * 1.44269504088896340736f;
    const int qtiles = (L + 15) / 16;
    const int lk_pad = (L + KC - 1) / KC * KC;
    const size_t ssmem = att_stream_smem_bytes(L, KC);
    if (C % 64 == 0 && L <= KC && lk_pad <= 256 && ssmem <= 75 * 1024) {
        // persistent CTAs, next item's tiles in flight while this one is computed
        CUtensorMap tm;
        TRY(make_tmap(&tm, false, qkv, static_cast<uint64_t>(B) * L, static_cast<uint64_t>(3) * C, lk_pad));
        int warps = 1;
        for (int w = 1; w <= ATT_GROUP_WARPS; ++w)
            if (qtiles % w == 0) warps = w;                // every warp the same number of q-tiles
        const int n_hg = C / 64;
        const int n_items = B * n_hg;
        const uint32_t n_hg_magic = n_hg > 1 ? static_cast<uint32_t>((1ull << 32) / static_cast<uint32_t>(n_hg)) + 1u : 0u;
        const int per_sm = static_cast<int>((227 * 1024) / (ssmem + 1024));
        int grid = num_sms() * (per_sm < 3 ? per_sm : 3);
        if (grid > n_items) grid = n_items;
        CK(launch_chain(attention_stream_kernel<D, KC, F16>, dim3(grid), dim3(warps * 32), ssmem, st, 1, tm, out, L, C, n_items,
                        n_hg_magic, scale_log2e));
    } else {
        const size_t smem = att_head_smem_bytes(L, D, KC);
        if (smem > 227 * 1024) return fail(DDC_E_INVALID, "attention: L=%d needs %zu B of shared memory", L, smem);
        const int warps = qtiles < ATT_MAX_WARPS ? qtiles : ATT_MAX_WARPS;
        CK(launch_chain(attention_head_kernel<D, KC, F16>, dim3(heads, B), dim3(warps * 32), smem, st, 1, qkv, out, L, C,
                        scale_log2e));
    }
    CK(cudaGetLastError());
    return DDC_OK;
}

static int run_attention(const bf16* qkv, bf16* out, int B, int L, int C, int heads, bool f16, cudaStream_t st) {
    if (heads <= 0 || C % heads != 0) return fail(DDC_E_INVALID, "attention: C=%d heads=%d", C, heads);
    const int D = C / heads;
    if (attention_fa_wanted(L, C, D)) {
        if (D == 32) return f16 ? launch_att_fa<32, true>(qkv, out, B, L, C, heads, st) : launch_att_fa<32, false>(qkv, out, B, L, C, heads, st);
        return f16 ? launch_att_fa<64, true>(qkv, out, B, L, C, heads, st) : launch_att_fa<64, false>(qkv, out, B, L, C, heads, st);
    }
    const int kc = att_pick_kc(L);
#define ATT_CASE(DD, KK)                                                                \
    if (D == DD && kc == KK)                                                            \
        return f16 ? launch_att<DD, KK, true>(qkv, out, B, L, C, heads, st)             \
                   : launch_att<DD, KK, false>(qkv, out, B, L, C, heads, st);
    ATT_CASE(32, 128) ATT_CASE(32, 96) ATT_CASE(32, 80) ATT_CASE(32, 64) ATT_CASE(32, 48) ATT_CASE(32, 32)
    ATT_CASE(64, 128) ATT_CASE(64, 96) ATT_CASE(64, 80) ATT_CASE(64, 64) ATT_CASE(64, 48) ATT_CASE(64, 32)
#undef ATT_CASE
    return fail(DDC_E_INVALID, "attention: head_dim %d unsupported (32 or 64)", D);
}

// ------------------------------------------------------------------------------- fused qkv + attention launch
// DDC_QKV_ATTN=0 falls back to the separate qkv GEMM and attention kernels.
static bool qkv_attn_usable(int L, int C, int heads) {
    return L == QA_L && heads > 0 && C % heads == 0 && C / heads == 32 && C % 64 == 0 && C <= 384 && env_int("DDC_QKV_ATTN", 1) != 0;
}

static long long* g_qa_trace = nullptr;      // tuning aid: ddc_dbg_qkv_attn_trace

template <bool F16>
static int launch_qkv_attn(const CUtensorMap& tmA, const CUtensorMap& tmW, QkvAttnParams p, cudaStream_t st) {
    static int attr_dev_mask = 0;
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (!(attr_dev_mask & (1 << dev))) {
        CK(cudaFuncSetAttribute(qkv_attn_tc_kernel<F16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        CK(cudaFuncSetAttribute(qkv_attn_tc_kernel<F16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_dev_mask |= (1 << dev);
    }
    int stages = static_cast<int>((227 * 1024 - static_cast<long>(qkv_attn_tc_smem_bytes(p.C, 0))) / (96 * 128));
    if (stages > 8) stages = 8;
    if (stages < 2) return fail(DDC_E_INVALID, "qkv+attention: C=%d leaves no room for the weight ring", p.C);
    p.stages = stages;
    p.n_hg = p.C / 64;
    p.num_tp = (p.B + 3) / 4;
    p.scale_log2e = (1.0f / sqrtf(32.0f)) * 1.44269504088896340736f;
    p.trace = g_qa_trace;
    // Work units per 4-image tile: the tile's head pairs are cut into n_split runs, each run a unit of its own (it reloads
    // the tile's activations, ~0.4 head-pair times).  More, shorter units shorten the last round of a large grid and
    // spread a small batch over more SMs; the result of an (image, head) does not depend on the cut.  Picked by cost:
    // rounds * (head pairs per unit + reload).
    p.n_split = 1;
    const int pairs = num_sms() / 2;
    float best = 1e30f;
    for (int d = 1; d <= p.n_hg; ++d) {
        if (p.n_hg % d) continue;
        const int rounds = (p.num_tp * d + pairs - 1) / pairs;
        const float cost = static_cast<float>(rounds) * (static_cast<float>(p.n_hg / d) + 0.4f);
        if (cost < best - 1e-3f) { best = cost; p.n_split = d; }
    }
    const int units = p.num_tp * p.n_split;
    int grid = 2 * units < num_sms() ? 2 * units : num_sms();
    grid = grid / 2 * 2;
    const size_t smem = qkv_attn_tc_smem_bytes(p.C, stages);
    if (g_qa_trace) CK(launch_chain(qkv_attn_tc_kernel<F16, true>, dim3(grid), dim3(QT_THREADS), smem, st, 2, tmA, tmW, p));
    else CK(launch_chain(qkv_attn_tc_kernel<F16, false>, dim3(grid), dim3(QT_THREADS), smem, st, 2, tmA, tmW, p));
    return DDC_OK;
}

// ------------------------------------------------------------------------------- plan
struct LayerW {
    float *ln1_g, *ln1_b, *ln2_g, *ln2_b, *b_qkv, *b_proj, *b_fc1, *b_fc2;
    bf16 *w_qkv, *w_proj, *w_fc1, *w_fc2;
    CUtensorMap tm_qkv[2], tm_proj[2], tm_fc1[2], tm_fc2[2];   // W tile boxes for cluster size 1 and 2
    CUtensorMap tm_fc1_tail;                                  // fc1 as 64-row half tiles (fused block tail)
    bf16* w_qkv_hp;                                           // qkv weight rows regrouped per head pair (fused qkv + attention)
    float* b_qkv_hp;
    CUtensorMap tm_qkv_hp;                                    // box 64 x 96
};

struct GraphKey {
    int B;
    const void *x, *x0_last, *x0_seq, *eps_seq;
    bool operator==(const GraphKey& o) const {
        return B == o.B && x == o.x && x0_last == o.x0_last && x0_seq == o.x0_seq && eps_seq == o.eps_seq;
    }
};

struct ddc_plan {
    ddc_shape sh;
    int maxB, device;
    int L, n, P, C, Kp, Mpad;
    int bn_qkv, bn_fc1, bn_row, bn_head;
    bool packed;
    bool f16;       // 16-bit operand format: false = bfloat16 (default), true = IEEE half
    bool separate_qkv_attn;   // ddc_plan_set_fused_qkv_attention(plan, 0): qkv GEMM and attention as two kernels for every shape
    bool f32;       // fp32 mode: every contraction in fp32 on the CUDA cores (fp32_path.cuh); overrides f16
    // fp32 mode reads the weight matrices where the caller keeps them (state_dict tensors given to ddc_pack_weights)
    const float* w32_patch;
    const float* w32_head;
    std::vector<const float*> w32_qkv, w32_proj, w32_fc1, w32_fc2;
    // weight arena
    bf16* w_patch;
    float* posb;
    float* time_tab;
    std::vector<LayerW> layers;
    float *lnf_g, *lnf_b, *b_head;
    bf16* w_head;
    CUtensorMap tm_wpatch[2], tm_whead[2];
    CUtensorMap tm_posb;                     // position table [L + 127, C] fp32, box 32 x 128 (the embed GEMM's "residual")
    // work arena
    float* X;
    bf16 *QKV, *O, *Hd, *Ap, *XN;
    float* Y;
    StepEntry* tbl;
    int* ctr;
    bf16* QKVcls;                            // [maxB, 3C] q | k | v of the class rows (fused qkv + attention)
    CUtensorMap tm_XN64, tm_XNcls, st_QKVcls; // XN as 64-row boxes; XN's class rows (row stride L*C); store map of QKVcls
    CUtensorMap tm_Ap, tm_O, tm_Hd, tm_XN;   // A-operand loads (box 64 x 128, padded row count)
    CUtensorMap tm_Xld;                      // residual loads (fp32, box 32 x 32, padded row count)
    // store maps are clipped to exactly batch*L rows, so they are rebuilt when the batch changes
    int maps_B;
    CUtensorMap st_QKV, st_Hd, st_XN, st_X, st_Y;      // st_Y only when P % 32 == 0 (fp32 boxes are 32 columns wide)
    // graph cache
    cudaGraphExec_t gexec;
    GraphKey gkey;
    // ddc_forward_fixed_scratch: activation pointers / descriptors into the library's built-in array (plain
    // members, so that entry point never touches the heap)
    struct WorkView {
        float* X;
        bf16 *QKV, *O, *Hd, *Ap, *XN;
        float* Y;
        StepEntry* tbl;
        int* ctr;
        bf16* QKVcls;
        CUtensorMap tm_Ap, tm_O, tm_Hd, tm_XN, tm_Xld, st_QKV, st_Hd, st_XN, st_X, st_Y;
        CUtensorMap tm_XN64, tm_XNcls, st_QKVcls;
        int maps_B, Mpad, maxB;
        cudaGraphExec_t gexec;
    };
    WorkView fixed_view;
    int fixed_view_cap;              // images the fixed view was laid out for (0 = not built yet)
};

struct Carver {
    char* base;
    size_t off;
    template <typename T>
    T* take(size_t n) {
        off = (off + 255) & ~static_cast<size_t>(255);
        T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
        off += n * sizeof(T);
        return p;
    }
};

static int check_shape(const ddc_shape* s, int max_batch) {
    if (s == nullptr) return fail(DDC_E_INVALID, "shape is NULL");
    if (s->in_chans != 3) return fail(DDC_E_INVALID, "in_chans must be 3 (got %d)", s->in_chans);
    if (s->patch <= 0 || s->patch % 2 != 0 || s->img_h % s->patch != 0 || s->img_w % s->patch != 0)
        return fail(DDC_E_INVALID, "patch %d must be even and divide the image %dx%d", s->patch, s->img_h, s->img_w);
    if (s->embed % 64 != 0 || s->embed > 512 || s->embed <= 0)
        return fail(DDC_E_INVALID, "embed_dim %d must be a multiple of 64 and <= 512", s->embed);
    if (s->heads <= 0 || s->embed % s->heads != 0 || (s->embed / s->heads != 32 && s->embed / s->heads != 64))
        return fail(DDC_E_INVALID, "head_dim = %d/%d must be 32 or 64", s->embed, s->heads);
    if (s->depth <= 0 || s->total_steps <= 0 || max_batch <= 0) return fail(DDC_E_INVALID, "depth/total_steps/max_batch");
    const int P = 3 * s->patch * s->patch;
    if (P % 16 != 0 || ((P + 63) / 64 * 64) > 512) return fail(DDC_E_INVALID, "patch dim %d unsupported", P);
    return DDC_OK;
}

static void carve(ddc_plan* pl, void* weight_arena, void* work_arena, size_t* wbytes, size_t* kbytes) {
    const ddc_shape& s = pl->sh;
    const size_t C = pl->C, L = pl->L, P = pl->P, Kp = pl->Kp, Mp = pl->Mpad;
    Carver w{static_cast<char*>(weight_arena), 0};
    pl->w_patch = w.take<bf16>(C * Kp);
    pl->posb = w.take<float>((L + 127) * C);        // position rows, extended so any tile's 128 rows are contiguous
    pl->time_tab = w.take<float>(static_cast<size_t>(s.total_steps) * C);
    pl->layers.resize(s.depth);
    for (auto& l : pl->layers) {
        l.ln1_g = w.take<float>(C);  l.ln1_b = w.take<float>(C);
        l.w_qkv = w.take<bf16>(3 * C * C);  l.b_qkv = w.take<float>(3 * C);
        l.w_proj = w.take<bf16>(C * C);  l.b_proj = w.take<float>(C);
        l.ln2_g = w.take<float>(C);  l.ln2_b = w.take<float>(C);
        l.w_fc1 = w.take<bf16>(C * C);  l.b_fc1 = w.take<float>(C);
        l.w_fc2 = w.take<bf16>(C * C);  l.b_fc2 = w.take<float>(C);
        l.w_qkv_hp = w.take<bf16>(3 * C * C);  l.b_qkv_hp = w.take<float>(3 * C);
    }
    pl->lnf_g = w.take<float>(C);
    pl->lnf_b = w.take<float>(C);
    pl->w_head = w.take<bf16>(P * C);
    pl->b_head = w.take<float>(P);
    *wbytes = (w.off + 255) & ~static_cast<size_t>(255);

    Carver k{static_cast<char*>(work_arena), 0};
    pl->X = k.take<float>(Mp * C);
    pl->QKV = k.take<bf16>(Mp * 3 * C);
    pl->O = k.take<bf16>(Mp * C);
    pl->Hd = k.take<bf16>(Mp * C);
    pl->Ap = k.take<bf16>(Mp * Kp);
    pl->XN = k.take<bf16>(Mp * C);
    pl->Y = k.take<float>(Mp * P);
    pl->tbl = k.take<StepEntry>(s.total_steps);
    pl->ctr = k.take<int>(64);
    pl->QKVcls = k.take<bf16>(static_cast<size_t>(pl->maxB) * 3 * C);
    *kbytes = (k.off + 255) & ~static_cast<size_t>(255);
}

static void derive(ddc_plan* pl, const ddc_shape* s, int max_batch) {
    pl->sh = *s;
    pl->maxB = max_batch;
    pl->n = (s->img_h / s->patch) * (s->img_w / s->patch);
    pl->L = pl->n + 1;
    pl->P = 3 * s->patch * s->patch;
    pl->C = s->embed;
    pl->Kp = (pl->P + 63) / 64 * 64;
    pl->Mpad = (max_batch * pl->L + G2_BM - 1) / G2_BM * G2_BM;
    pl->bn_qkv = pick_bn_stream(3 * pl->C);
    pl->bn_fc1 = pick_bn_stream(pl->C);
    pl->bn_row = pick_bn_row(pl->C);
    pl->bn_head = pick_bn_f32(pl->P);
}

extern "C" int ddc_arena_bytes(const ddc_shape* shape, int32_t max_batch, size_t* weight_bytes, size_t* work_bytes) {
    TRY(check_shape(shape, max_batch));
    ddc_plan tmp{};
    derive(&tmp, shape, max_batch);
    size_t w = 0, k = 0;
    carve(&tmp, nullptr, nullptr, &w, &k);
    if (weight_bytes) *weight_bytes = w;
    if (work_bytes) *work_bytes = k;
    return DDC_OK;
}

extern "C" int ddc_plan_create(const ddc_shape* shape, int32_t max_batch, int32_t device, void* weight_arena,
                               size_t weight_bytes, void* work_arena, size_t work_bytes, ddc_plan** out) {
    if (out == nullptr) return fail(DDC_E_INVALID, "out is NULL");
    *out = nullptr;
    TRY(check_shape(shape, max_batch));
    if (weight_arena == nullptr || work_arena == nullptr) return fail(DDC_E_INVALID, "arena pointer is NULL");
    if ((reinterpret_cast<uintptr_t>(weight_arena) | reinterpret_cast<uintptr_t>(work_arena)) & 255)
        return fail(DDC_E_INVALID, "arenas must be 256-byte aligned");
    CK(cudaSetDevice(device));
    int major = 0, minor = 0;
    CK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    CK(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
    if (major != 10)
        return fail(DDC_E_CUDA, "device %d is sm_%d%d; this library contains sm_100a code only", device, major, minor);
    ddc_plan* pl = new ddc_plan();
    derive(pl, shape, max_batch);
    pl->device = device;
    pl->packed = false;
    pl->f16 = false;
    pl->f32 = false;
    pl->separate_qkv_attn = false;
    pl->gexec = nullptr;
    pl->gkey = GraphKey{0, nullptr, nullptr, nullptr, nullptr};
    pl->fixed_view_cap = 0;
    size_t w = 0, k = 0;
    carve(pl, weight_arena, work_arena, &w, &k);
    if (w > weight_bytes || k > work_bytes) {
        delete pl;
        return fail(DDC_E_STATE, "arena too small: need %zu/%zu B, got %zu/%zu B", w, k, weight_bytes, work_bytes);
    }
    const uint64_t C = pl->C, Mp = pl->Mpad;
    int r = DDC_OK;
    auto mk = [&](CUtensorMap* m, bool f32, const void* base, uint64_t rows, uint64_t cols, uint32_t box) {
        if (r == DDC_OK) r = make_tmap(m, f32, base, rows, cols, box);
    };
    pl->maps_B = 0;
    mk(&pl->tm_Ap, false, pl->Ap, Mp, pl->Kp, G2_BM);
    mk(&pl->tm_O, false, pl->O, Mp, C, G2_BM);
    mk(&pl->tm_Hd, false, pl->Hd, Mp, C, G2_BM);
    mk(&pl->tm_XN, false, pl->XN, Mp, C, G2_BM);
    mk(&pl->tm_Xld, true, pl->X, Mp, C, G2_BM);
    mk(&pl->tm_posb, true, pl->posb, static_cast<uint64_t>(pl->L) + 127, C, G2_BM);
    mk(&pl->tm_XN64, false, pl->XN, Mp, C, 64);
    if (r == DDC_OK) r = make_tmap(&pl->tm_XNcls, false, pl->XN, static_cast<uint64_t>(pl->maxB), C, G2_BM, static_cast<uint64_t>(pl->L) * C);
    for (auto& l : pl->layers) mk(&l.tm_qkv_hp, false, l.w_qkv_hp, 3 * C, C, 96);
    for (int cs = 1; cs <= 2; ++cs) {
        mk(&pl->tm_wpatch[cs - 1], false, pl->w_patch, C, pl->Kp, pl->bn_row / cs);
        mk(&pl->tm_whead[cs - 1], false, pl->w_head, pl->P, C, pl->bn_head / cs);
        for (auto& l : pl->layers) {
            mk(&l.tm_qkv[cs - 1], false, l.w_qkv, 3 * C, C, pl->bn_qkv / cs);
            mk(&l.tm_proj[cs - 1], false, l.w_proj, C, C, pl->bn_row / cs);
            mk(&l.tm_fc1[cs - 1], false, l.w_fc1, C, C, pl->bn_fc1 / cs);
            mk(&l.tm_fc2[cs - 1], false, l.w_fc2, C, C, pl->bn_row / cs);
        }
    }
    for (auto& l : pl->layers) mk(&l.tm_fc1_tail, false, l.w_fc1, C, C, 64);
    if (r != DDC_OK) {
        delete pl;
        return r;
    }
    *out = pl;
    return DDC_OK;
}

extern "C" int ddc_plan_destroy(ddc_plan* plan) {
    if (plan == nullptr) return DDC_OK;
    if (plan->gexec) cudaGraphExecDestroy(plan->gexec);
    delete plan;
    return DDC_OK;
}

extern "C" int ddc_plan_set_operand_format(ddc_plan* plan, int32_t fmt) {
    if (plan == nullptr) return fail(DDC_E_INVALID, "plan is NULL");
    if (fmt < 0 || fmt > 2) return fail(DDC_E_INVALID, "operand format %d (0 = bf16, 1 = fp16, 2 = fp32)", fmt);
    if (plan->f16 != (fmt == 1) || plan->f32 != (fmt == 2)) {
        plan->f16 = (fmt == 1);
        plan->f32 = (fmt == 2);
        plan->packed = false;            // packed weights are in the old format
        if (plan->gexec) {
            cudaGraphExecDestroy(plan->gexec);
            plan->gexec = nullptr;
        }
    }
    return DDC_OK;
}

static int f32_max_batch(const ddc_plan* pl);

extern "C" int ddc_plan_max_batch(const ddc_plan* plan) {
    if (plan == nullptr) return 0;
    return plan->f32 ? f32_max_batch(plan) : plan->maxB;
}

extern "C" int ddc_launches_per_step(const ddc_plan* plan) {
    if (plan == nullptr) return 0;
    // patchify, embed GEMM, per block {qkv GEMM, attention, fused tail} or {qkv, attention, proj, fc1, fc2}, head GEMM,
    // head_ddim (the fused tail needs embed_dim % 128 == 0 and <= 384; other shapes use the separate kernels)
    if (plan->f32) return 2 + 7 * plan->sh.depth + 3;      // patchify, embed, depth x {LN, qkv, attention, proj, LN, fc1, fc2}, LN, head, head_ddim
    const bool fused = block_tail_usable(plan->maxB * plan->L, plan->C, plan->bn_row);
    int att = 1;                     // the tcgen05 attention hands a few remainder rows to a second, CUDA-core launch
    if (attention_fa_wanted(plan->L, plan->C, plan->C / plan->sh.heads)) {
        int n_qt, rows_main;
        fa_split_rows(plan->L, &n_qt, &rows_main);
        if (rows_main < plan->L) att = 2;
    }
    // inside the sampling loop patchify runs once per loop, not per step, when head_ddim can emit the next patch matrix
    HeadDdimParams q{};
    q.ldy = plan->P; q.H = plan->sh.img_h; q.W = plan->sh.img_w; q.p = plan->sh.patch; q.L = plan->L; q.Kp = plan->Kp;
    const int patchify = (env_int("DDC_STEP_PATCHIFY", 0) || head_ddim_rows(q, true) == 0) ? 1 : 0;
    return patchify + 1 + ((fused ? 2 : 4) + att) * plan->sh.depth + 2;
}

extern "C" int ddc_plan_fused_qkv_attention(const ddc_plan* plan) {
    if (plan == nullptr || plan->f32) return 0;
    return (!plan->separate_qkv_attn && qkv_attn_usable(plan->L, plan->C, plan->sh.heads)) ? 1 : 0;
}

extern "C" int ddc_plan_set_fused_qkv_attention(ddc_plan* plan, int32_t enable) {
    if (plan == nullptr) return fail(DDC_E_INVALID, "plan is NULL");
    const bool sep = enable == 0;
    if (plan->separate_qkv_attn != sep && plan->gexec) {      // the captured step graph holds the other kernel sequence
        cudaGraphExecDestroy(plan->gexec);
        plan->gexec = nullptr;
    }
    plan->separate_qkv_attn = sep;
    return DDC_OK;
}

// ------------------------------------------------------------------------------- packing
extern "C" int ddc_pack_weights(ddc_plan* pl, const float* const* t, int32_t n_tensors, void* stream) {
    if (pl == nullptr || t == nullptr) return fail(DDC_E_INVALID, "NULL argument");
    const int expect = 9 + 12 * pl->sh.depth;
    if (n_tensors != expect) return fail(DDC_E_INVALID, "expected %d state_dict tensors, got %d", expect, n_tensors);
    for (int i = 0; i < n_tensors; ++i)
        if (t[i] == nullptr) return fail(DDC_E_INVALID, "state_dict tensor %d is NULL", i);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(pl->device));
    const int C = pl->C, L = pl->L, P = pl->P;
    auto pack = [&](const float* src, bf16* dst, int rows, int cols, int cols_pad) {
        if (pl->f32) return;                          // fp32 mode multiplies by the caller's fp32 matrices directly
        if (pl->f16) pack_16bit_kernel<true><<<296, 256, 0, st>>>(src, dst, rows, cols, cols_pad);
        else pack_16bit_kernel<false><<<296, 256, 0, st>>>(src, dst, rows, cols, cols_pad);
    };
    auto copy = [&](const float* src, float* dst, size_t n) { copy_f32_kernel<<<148, 256, 0, st>>>(src, dst, n); };
    const float* cls = t[0];
    const float* pos = t[1];
    pl->w32_patch = t[2];
    pl->w32_head = t[5 + 12 * pl->sh.depth + 2];
    pl->w32_qkv.assign(pl->sh.depth, nullptr); pl->w32_proj.assign(pl->sh.depth, nullptr);
    pl->w32_fc1.assign(pl->sh.depth, nullptr); pl->w32_fc2.assign(pl->sh.depth, nullptr);
    for (int i = 0; i < pl->sh.depth; ++i) {
        const float* const* b = t + 5 + 12 * i;
        pl->w32_qkv[i] = b[2]; pl->w32_proj[i] = b[4]; pl->w32_fc1[i] = b[8]; pl->w32_fc2[i] = b[10];
    }
    pack(t[2], pl->w_patch, C, P, pl->Kp);
    pack_posb_kernel<<<148, 256, 0, st>>>(pos, cls, t[3], pl->posb, L, C, L + 127);
    copy(t[4], pl->time_tab, static_cast<size_t>(pl->sh.total_steps) * C);
    // 16-bit formats: the affine part of every LayerNorm goes into the Linear that consumes it (W' = W o gamma,
    // b' = b + W beta; norm1 -> qkv, norm2 -> fc1, final norm -> head), and the LayerNorm vectors in the arena become
    // ones / zeros: the kernels that fuse a LayerNorm into their epilogue then only normalise (the fused block tail does not
    // even load the vectors).  fp32 mode keeps the reference's form (its kernels read the caller's fp32 matrices).
    auto fold = [&](const float* w, const float* g, const float* be, const float* bias, bf16* wout, float* bout, int n, int k, int kpad) {
        if (pl->f16) fold_ln_linear_kernel<true><<<n, 128, 0, st>>>(w, g, be, bias, wout, bout, k, kpad);
        else fold_ln_linear_kernel<false><<<n, 128, 0, st>>>(w, g, be, bias, wout, bout, k, kpad);
    };
    auto ln_vectors = [&](const float* g, const float* be, float* g_dst, float* b_dst) {
        if (pl->f32) { copy(g, g_dst, C); copy(be, b_dst, C); return; }
        fill_f32_kernel<<<4, 256, 0, st>>>(g_dst, 1.f, C);
        fill_f32_kernel<<<4, 256, 0, st>>>(b_dst, 0.f, C);
    };
    for (int i = 0; i < pl->sh.depth; ++i) {
        const float* const* b = t + 5 + 12 * i;
        LayerW& l = pl->layers[i];
        ln_vectors(b[0], b[1], l.ln1_g, l.ln1_b);
        if (pl->f32) copy(b[3], l.b_qkv, 3 * C);
        else fold(b[2], b[0], b[1], b[3], l.w_qkv, l.b_qkv, 3 * C, C, C);
        if (!pl->f32) repack_qkv_hp_kernel<<<296, 64, 0, st>>>(l.w_qkv, l.b_qkv, l.w_qkv_hp, l.b_qkv_hp, C);
        pack(b[4], l.w_proj, C, C, C);  copy(b[5], l.b_proj, C);
        ln_vectors(b[6], b[7], l.ln2_g, l.ln2_b);
        if (pl->f32) copy(b[9], l.b_fc1, C);
        else fold(b[8], b[6], b[7], b[9], l.w_fc1, l.b_fc1, C, C, C);
        pack(b[10], l.w_fc2, C, C, C);  copy(b[11], l.b_fc2, C);
    }
    const float* const* e = t + 5 + 12 * pl->sh.depth;
    ln_vectors(e[0], e[1], pl->lnf_g, pl->lnf_b);
    if (pl->f32) copy(e[3], pl->b_head, P);
    else fold(e[2], e[0], e[1], e[3], pl->w_head, pl->b_head, P, C, C);
    CK(cudaGetLastError());
    // The step-path kernels are launched with programmatic dependent launch and read weights (bias, LayerNorm affine)
    // in their prologue, BEFORE they wait for their predecessor.  Packing is rare, so make it a full stop: once this
    // returns, every packed byte is visible to anything launched later, on any stream.
    CK(cudaStreamSynchronize(st));
    pl->packed = true;
    return DDC_OK;
}

// ------------------------------------------------------------------------------- per-class timing
enum { KC_PATCHIFY = 0, KC_EMBED, KC_QKV, KC_ATTN, KC_PROJ, KC_FC1, KC_FC2, KC_HEAD, KC_DDIM, KC_COUNT };

struct Prof {
    cudaStream_t st;
    std::vector<cudaEvent_t> ev;
    std::vector<int> cls;
    int begin() {
        cudaEvent_t e;
        CK(cudaEventCreate(&e));
        ev.push_back(e);
        CK(cudaEventRecord(e, st));
        return DDC_OK;
    }
    int mark(int c) {
        cudaEvent_t e;
        CK(cudaEventCreate(&e));
        ev.push_back(e);
        cls.push_back(c);
        CK(cudaEventRecord(e, st));
        return DDC_OK;
    }
    void release() {
        for (cudaEvent_t e : ev) cudaEventDestroy(e);
        ev.clear();
        cls.clear();
    }
};
#define MARK(c)                          \
    do {                                 \
        if (prof) TRY(prof->mark(c));    \
    } while (0)

// ------------------------------------------------------------------------------- one predictor evaluation
// Enqueues: patchify -> embed GEMM -> depth x [LN1+qkv GEMM, attention, proj GEMM(+res),
// LN2+fc1 GEMM(+GELU), fc2 GEMM(+res)] -> LN+head GEMM.  Leaves the head output in pl->Y.
static int ensure_store_maps(ddc_plan* pl, int B) {
    if (pl->maps_B == B) return DDC_OK;
    const uint64_t M = static_cast<uint64_t>(B) * pl->L, C = pl->C;
    TRY(make_tmap(&pl->st_QKV, false, pl->QKV, M, 3 * C, G2_BM));
    TRY(make_tmap(&pl->st_Hd, false, pl->Hd, M, C, G2_BM));
    TRY(make_tmap(&pl->st_XN, false, pl->XN, M, C, G2_BM));
    TRY(make_tmap(&pl->st_X, true, pl->X, M, C, G2_BM));
    if (pl->P % 32 == 0) TRY(make_tmap(&pl->st_Y, true, pl->Y, M, pl->P, G2_BM));
    TRY(make_tmap(&pl->st_QKVcls, false, pl->QKVcls, static_cast<uint64_t>(B), 3 * C, G2_BM));
    pl->maps_B = B;
    if (pl->gexec) {                 // the captured graph holds the old maps by value
        cudaGraphExecDestroy(pl->gexec);
        pl->gexec = nullptr;
    }
    return DDC_OK;
}

// ------------------------------------------------------------------------------- fp32 mode
// The fp32 activations (LayerNorm output, qkv, attention output, MLP hidden, im2col) overlay the region of the work arena
// that the 16-bit path uses for its 16-bit activations; X, Y, the step table and the counter are shared.  fp32 rows are
// twice as wide, so a plan created for max_batch images serves about half as many in this mode.
struct F32View {
    float *XN, *QKV, *O, *H, *Ap;
};
static bool f32_layout(const ddc_plan* pl, int B, F32View* v) {
    const size_t rows = static_cast<size_t>(B) * pl->L, C = pl->C, P = pl->P;
    Carver k{reinterpret_cast<char*>(pl->QKV), 0};
    F32View t;
    t.XN = k.take<float>(rows * C);
    t.QKV = k.take<float>(rows * 3 * C);
    t.O = k.take<float>(rows * C);
    t.H = k.take<float>(rows * C);
    t.Ap = k.take<float>(rows * P);
    if (v != nullptr) *v = t;
    return k.off <= static_cast<size_t>(reinterpret_cast<char*>(pl->Y) - reinterpret_cast<char*>(pl->QKV));
}
static int f32_max_batch(const ddc_plan* pl) {
    int b = pl->maxB;
    while (b > 0 && !f32_layout(pl, b, nullptr)) --b;
    return b;
}

template <int EPI>
static int launch_sgemm(const float* A, const float* Wt, const float* bias, float* out, int M, int N, int K, const SgemmEmbed& e,
                        cudaStream_t st) {
    if (K % 8 != 0) return fail(DDC_E_INVALID, "fp32 GEMM: K=%d must be a multiple of 8", K);
    dim3 grid((N + 127) / 128, (M + 127) / 128);
    sgemm_f32_kernel<EPI><<<grid, 256, 0, st>>>(A, Wt, bias, out, M, N, K, e);
    CK(cudaGetLastError());
    return DDC_OK;
}

static int enqueue_predictor_f32(ddc_plan* pl, const float* x, int B, const int* t_img, bool use_step_ctr, cudaStream_t st, Prof* prof) {
    const ddc_shape& s = pl->sh;
    const int C = pl->C, L = pl->L, P = pl->P, M = B * L;
    F32View v;
    if (!f32_layout(pl, B, &v)) return fail(DDC_E_INVALID, "fp32 mode: batch %d does not fit this plan (at most %d)", B, f32_max_batch(pl));
    const int D = C / s.heads;
    static int attr_dev_mask = 0;
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (!(attr_dev_mask & (1 << dev))) {
        CK(cudaFuncSetAttribute(attention_f32_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        CK(cudaFuncSetAttribute(attention_f32_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_dev_mask |= (1 << dev);
    }
    const size_t att_smem = attention_f32_smem(L, D);
    if (att_smem > 227 * 1024) return fail(DDC_E_INVALID, "fp32 attention: L=%d needs %zu B of shared memory", L, att_smem);
    SgemmEmbed none{};
    patchify_f32_kernel<<<dim3(s.img_h / s.patch, B), 256, 0, st>>>(x, v.Ap, s.img_h, s.img_w, s.patch, L, P, use_step_ctr ? pl->ctr : nullptr);
    CK(cudaGetLastError());
    MARK(KC_PATCHIFY);
    SgemmEmbed e{};
    e.posb = pl->posb; e.time_tab = pl->time_tab; e.t_img = t_img; e.tbl = pl->tbl; e.cur = pl->ctr; e.L = L; e.T = s.total_steps;
    TRY(launch_sgemm<F_EMBED>(v.Ap, pl->w32_patch, nullptr, pl->X, M, C, P, e, st));
    MARK(KC_EMBED);
    const int ln_blocks = (M + 7) / 8;
    for (int i = 0; i < s.depth; ++i) {
        const LayerW& l = pl->layers[i];
        layernorm_f32_kernel<<<ln_blocks, 256, 0, st>>>(pl->X, l.ln1_g, l.ln1_b, v.XN, M, C, 1e-5f);
        TRY(launch_sgemm<F_BIAS>(v.XN, pl->w32_qkv[i], l.b_qkv, v.QKV, M, 3 * C, C, none, st));
        MARK(KC_QKV);
        const float scale = 1.0f / sqrtf(static_cast<float>(D));
        if (D == 32) attention_f32_kernel<32><<<dim3(s.heads, B), 256, att_smem, st>>>(v.QKV, v.O, L, C, scale);
        else attention_f32_kernel<64><<<dim3(s.heads, B), 256, att_smem, st>>>(v.QKV, v.O, L, C, scale);
        CK(cudaGetLastError());
        MARK(KC_ATTN);
        TRY(launch_sgemm<F_RESID>(v.O, pl->w32_proj[i], l.b_proj, pl->X, M, C, C, none, st));
        MARK(KC_PROJ);
        layernorm_f32_kernel<<<ln_blocks, 256, 0, st>>>(pl->X, l.ln2_g, l.ln2_b, v.XN, M, C, 1e-5f);
        TRY(launch_sgemm<F_GELU>(v.XN, pl->w32_fc1[i], l.b_fc1, v.H, M, C, C, none, st));
        MARK(KC_FC1);
        TRY(launch_sgemm<F_RESID>(v.H, pl->w32_fc2[i], l.b_fc2, pl->X, M, C, C, none, st));
        MARK(KC_FC2);
    }
    layernorm_f32_kernel<<<ln_blocks, 256, 0, st>>>(pl->X, pl->lnf_g, pl->lnf_b, v.XN, M, C, 1e-5f);
    TRY(launch_sgemm<F_BIAS>(v.XN, pl->w32_head, pl->b_head, pl->Y, M, P, C, none, st));
    CK(cudaGetLastError());
    MARK(KC_HEAD);
    return DDC_OK;
}

// Enqueues: patchify -> patch-embed GEMM (+cls/pos/time, emits LN1(0)) -> depth x [qkv GEMM,
// attention, proj GEMM (+residual, emits LN2), fc1 GEMM (+GELU), fc2 GEMM (+residual, emits the
// next LayerNorm)] -> head GEMM.  Leaves the head output in pl->Y.
static int enqueue_predictor(ddc_plan* pl, const float* x, int B, const int* t_img, bool use_step_ctr, cudaStream_t st,
                             Prof* prof = nullptr, bool skip_patchify = false) {
    if (pl->f32) return enqueue_predictor_f32(pl, x, B, t_img, use_step_ctr, st, prof);
    const ddc_shape& s = pl->sh;
    const int C = pl->C, L = pl->L, M = B * L;
    TRY(ensure_store_maps(pl, B));
    const bool f16 = pl->f16;
    const int cs = gemm_cluster_size(M);
    if (!skip_patchify) {      // inside the sampling loop the previous step's head_ddim has already written Ap (see enqueue_step)
        CK(f16 ? launch_patchify<true>(x, pl->Ap, B, s.img_h, s.img_w, s.patch, L, pl->Kp, use_step_ctr ? pl->ctr : nullptr, st)
               : launch_patchify<false>(x, pl->Ap, B, s.img_h, s.img_w, s.patch, L, pl->Kp, use_step_ctr ? pl->ctr : nullptr, st));
        MARK(KC_PATCHIFY);
    }
    const LayerW& l0 = pl->layers[0];
    Gemm2Params g{};
    g.M = M; g.N = C; g.K = pl->Kp; g.bn = pl->bn_row; g.stages = pick_stages2(g.K, g.N, g.bn / cs, true);
    g.ln_g = l0.ln1_g; g.ln_b = l0.ln1_b; g.ln_eps = 1e-5f;
    g.posb = pl->posb; g.L = L; g.time_tab = pl->time_tab; g.t_img = t_img; g.tbl = pl->tbl; g.cur = pl->ctr; g.B = B;
    g.T = s.total_steps;
    TRY((launch_gemm2<R_EMBED>(pl->tm_Ap, pl->tm_wpatch[cs - 1], pl->st_X, pl->tm_posb, pl->st_XN, g, f16, cs, st)));
    MARK(KC_EMBED);
    for (int i = 0; i < s.depth; ++i) {
        LayerW& l = pl->layers[i];
        Gemm2Params q{};
        q.N = 3 * C; q.K = C; q.bn = pl->bn_qkv; q.bias = l.b_qkv;
        if (!pl->separate_qkv_attn && qkv_attn_usable(L, C, s.heads)) {
            // q | k | v of the class rows (one per image: a [B x 3C] GEMM over XN rows L apart), then qkv + attention of
            // the patch rows in one kernel that reads those class rows from L2
            const int ccs = gemm_cluster_size(B);
            q.M = B; q.stages = pick_stages2(q.K, q.N, q.bn / ccs, false);
            TRY((launch_gemm2<S_BF16>(pl->tm_XNcls, l.tm_qkv[ccs - 1], pl->st_QKVcls, pl->tm_Xld, pl->st_XN, q, f16, ccs, st)));
            MARK(KC_QKV);
            QkvAttnParams qa{};
            qa.B = B; qa.C = C; qa.bias_hp = l.b_qkv_hp; qa.qkv_cls = pl->QKVcls; qa.out = pl->O;
            if (f16) TRY(launch_qkv_attn<true>(pl->tm_XN64, l.tm_qkv_hp, qa, st));
            else TRY(launch_qkv_attn<false>(pl->tm_XN64, l.tm_qkv_hp, qa, st));
            MARK(KC_ATTN);
        } else {
            q.M = M; q.stages = pick_stages2(q.K, q.N, q.bn / cs, false);
            TRY((launch_gemm2<S_BF16>(pl->tm_XN, l.tm_qkv[cs - 1], pl->st_QKV, pl->tm_Xld, pl->st_XN, q, f16, cs, st)));
            MARK(KC_QKV);
            TRY(run_attention(pl->QKV, pl->O, B, L, C, s.heads, f16, st));
            MARK(KC_ATTN);
        }
        const float* lnn_g = (i + 1 < s.depth) ? pl->layers[i + 1].ln1_g : pl->lnf_g;
        const float* lnn_b = (i + 1 < s.depth) ? pl->layers[i + 1].ln1_b : pl->lnf_b;
        if (block_tail_usable(M, C, pl->bn_row)) {
            BlockTailParams bt{};
            bt.M = M; bt.C = C; bt.ln_eps = 1e-5f;
            bt.b_proj = l.b_proj; bt.b_fc1 = l.b_fc1; bt.b_fc2 = l.b_fc2;
            if (f16) TRY((launch_block_tail_inst<true>(pl->tm_O, l.tm_proj[1], l.tm_fc1_tail, l.tm_fc2[1], pl->tm_Xld, pl->st_X, pl->st_XN, bt, st)));
            else TRY((launch_block_tail_inst<false>(pl->tm_O, l.tm_proj[1], l.tm_fc1_tail, l.tm_fc2[1], pl->tm_Xld, pl->st_X, pl->st_XN, bt, st)));
            MARK(KC_PROJ);
            continue;
        }
        Gemm2Params r{};
        r.M = M; r.N = C; r.K = C; r.bn = pl->bn_row; r.stages = pick_stages2(r.K, r.N, r.bn / cs, true);
        r.bias = l.b_proj; r.ln_g = l.ln2_g; r.ln_b = l.ln2_b; r.ln_eps = 1e-5f;
        TRY((launch_gemm2<R_RESID>(pl->tm_O, l.tm_proj[cs - 1], pl->st_X, pl->tm_Xld, pl->st_XN, r, f16, cs, st)));
        MARK(KC_PROJ);
        Gemm2Params f{};
        f.M = M; f.N = C; f.K = C; f.bn = pl->bn_fc1; f.stages = pick_stages2(f.K, f.N, f.bn / cs, false);
        f.bias = l.b_fc1;
        TRY((launch_gemm2<S_BF16_GELU>(pl->tm_XN, l.tm_fc1[cs - 1], pl->st_Hd, pl->tm_Xld, pl->st_XN, f, f16, cs, st)));
        MARK(KC_FC1);
        r.bias = l.b_fc2;
        r.ln_g = lnn_g;
        r.ln_b = lnn_b;
        TRY((launch_gemm2<R_RESID>(pl->tm_Hd, l.tm_fc2[cs - 1], pl->st_X, pl->tm_Xld, pl->st_XN, r, f16, cs, st)));
        MARK(KC_FC2);
    }
    Gemm2Params h{};
    h.M = M; h.N = pl->P; h.K = C; h.bn = pl->bn_head; h.stages = pick_stages2(h.K, h.N, h.bn / cs, false);
    h.bias = pl->b_head; h.out_f32 = pl->Y; h.ld_f32 = pl->P;
    h.f32_tma = (pl->P % 32 == 0 && h.bn % 32 == 0) ? 1 : 0;      // TMA-staged fp32 stores when the boxes fit
    TRY((launch_gemm2<S_F32>(pl->tm_XN, pl->tm_whead[cs - 1], h.f32_tma ? pl->st_Y : pl->st_XN, pl->tm_Xld, pl->st_XN, h, f16, cs, st)));
    MARK(KC_HEAD);
    return DDC_OK;
}


static int check_call(ddc_plan* pl, int batch) {
    if (pl == nullptr) return fail(DDC_E_INVALID, "plan is NULL");
    if (!pl->packed) return fail(DDC_E_STATE, "ddc_pack_weights has not been called on this plan");
    if (batch <= 0 || batch > pl->maxB) return fail(DDC_E_INVALID, "batch %d outside 1..%d", batch, pl->maxB);
    if (pl->f32 && !f32_layout(pl, batch, nullptr))
        return fail(DDC_E_INVALID, "fp32 mode: batch %d does not fit this plan (fp32 activations are twice as wide: at most %d)", batch,
                    f32_max_batch(pl));
    return DDC_OK;
}

extern "C" int ddc_forward(ddc_plan* pl, const float* x, const int32_t* t, int32_t batch, float* out, void* stream) {
    TRY(check_call(pl, batch));
    if (x == nullptr || t == nullptr || out == nullptr) return fail(DDC_E_INVALID, "NULL tensor");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(pl->device));
    TRY(enqueue_predictor(pl, x, batch, t, false, st));
    HeadDdimParams q{};
    q.y = pl->Y; q.ldy = pl->P; q.x = out;
    q.H = pl->sh.img_h; q.W = pl->sh.img_w; q.p = pl->sh.patch; q.L = pl->L; q.B = batch;
    CK(launch_head_ddim<MODE_FORWARD>(q, st));
    return DDC_OK;
}

// ------------------------------------------------------------------------------- forward without any allocation
// Some hosts may not allocate device memory at the moment they need a prediction: a serving loop that has
// handed its whole memory budget to another framework's pool, or a latency-critical thread for which
// cudaMalloc's implicit device synchronisation is unacceptable.  For them the activations of one forward pass can live in this
// fixed array instead of a caller-provided work arena.  It is part of the library image: every GPU that
// loads the library carries it until the process exits, which is why it is kept modest.
#define DDC_FIXED_SCRATCH_BYTES (32u << 20)
static __device__ __align__(256) unsigned char g_fixed_scratch[DDC_FIXED_SCRATCH_BYTES];
static void* g_fixed_scratch_dev[64] = {nullptr};   // per device: the array's DEVICE address (the symbol itself is a
                                                    // host-side stand-in and must never be handed to a kernel)

static size_t fixed_scratch_layout(const ddc_plan* pl, int B, ddc_plan::WorkView* v, char* origin) {
    const size_t C = pl->C, P = pl->P, Kp = pl->Kp;
    const size_t Mp = (static_cast<size_t>(B) * pl->L + G2_BM - 1) / G2_BM * G2_BM;
    Carver k{origin, 0};
    float* X = k.take<float>(Mp * C);
    bf16* QKV = k.take<bf16>(Mp * 3 * C);
    bf16* O = k.take<bf16>(Mp * C);
    bf16* Hd = k.take<bf16>(Mp * C);
    bf16* Ap = k.take<bf16>(Mp * Kp);
    bf16* XN = k.take<bf16>(Mp * C);
    float* Y = k.take<float>(Mp * P);
    bf16* QKVcls = k.take<bf16>(static_cast<size_t>(B) * 3 * C);
    if (v != nullptr) {
        v->X = X; v->QKV = QKV; v->O = O; v->Hd = Hd; v->Ap = Ap; v->XN = XN; v->Y = Y; v->QKVcls = QKVcls;
        v->tbl = nullptr; v->ctr = nullptr;
        v->Mpad = static_cast<int>(Mp);
        v->maxB = B;
        v->maps_B = 0;
        v->gexec = nullptr;
    }
    return (k.off + 255) & ~static_cast<size_t>(255);
}

static void view_save(const ddc_plan* pl, ddc_plan::WorkView* v) {
    v->X = pl->X; v->QKV = pl->QKV; v->O = pl->O; v->Hd = pl->Hd; v->Ap = pl->Ap; v->XN = pl->XN; v->Y = pl->Y;
    v->tbl = pl->tbl; v->ctr = pl->ctr; v->QKVcls = pl->QKVcls;
    v->tm_XN64 = pl->tm_XN64; v->tm_XNcls = pl->tm_XNcls; v->st_QKVcls = pl->st_QKVcls;
    v->tm_Ap = pl->tm_Ap; v->tm_O = pl->tm_O; v->tm_Hd = pl->tm_Hd; v->tm_XN = pl->tm_XN; v->tm_Xld = pl->tm_Xld;
    v->st_QKV = pl->st_QKV; v->st_Hd = pl->st_Hd; v->st_XN = pl->st_XN; v->st_X = pl->st_X; v->st_Y = pl->st_Y;
    v->maps_B = pl->maps_B; v->Mpad = pl->Mpad; v->maxB = pl->maxB; v->gexec = pl->gexec;
}

static void view_install(ddc_plan* pl, const ddc_plan::WorkView* v) {
    pl->X = v->X; pl->QKV = v->QKV; pl->O = v->O; pl->Hd = v->Hd; pl->Ap = v->Ap; pl->XN = v->XN; pl->Y = v->Y;
    pl->tbl = v->tbl; pl->ctr = v->ctr; pl->QKVcls = v->QKVcls;
    pl->tm_XN64 = v->tm_XN64; pl->tm_XNcls = v->tm_XNcls; pl->st_QKVcls = v->st_QKVcls;
    pl->tm_Ap = v->tm_Ap; pl->tm_O = v->tm_O; pl->tm_Hd = v->tm_Hd; pl->tm_XN = v->tm_XN; pl->tm_Xld = v->tm_Xld;
    pl->st_QKV = v->st_QKV; pl->st_Hd = v->st_Hd; pl->st_XN = v->st_XN; pl->st_X = v->st_X; pl->st_Y = v->st_Y;
    pl->maps_B = v->maps_B; pl->Mpad = v->Mpad; pl->maxB = v->maxB; pl->gexec = v->gexec;
}

extern "C" int ddc_forward_fixed_scratch(ddc_plan* pl, const float* x, const int32_t* t, int32_t batch, float* out,
                                         void* stream) {
    if (pl == nullptr) return fail(DDC_E_INVALID, "plan is NULL");
    if (!pl->packed) return fail(DDC_E_STATE, "ddc_pack_weights has not been called on this plan");
    if (x == nullptr || t == nullptr || out == nullptr) return fail(DDC_E_INVALID, "NULL tensor");
    if (batch <= 0) return fail(DDC_E_INVALID, "batch %d", batch);
    if (pl->f32) return fail(DDC_E_STATE, "ddc_forward_fixed_scratch serves the 16-bit operand formats only (the plan is in fp32 mode)");
    const size_t need = fixed_scratch_layout(pl, batch, nullptr, nullptr);
    if (need > DDC_FIXED_SCRATCH_BYTES) {
        int fit = batch;
        while (fit > 0 && fixed_scratch_layout(pl, fit, nullptr, nullptr) > DDC_FIXED_SCRATCH_BYTES) fit -= (fit > 64 ? fit / 16 : 1);
        while (fixed_scratch_layout(pl, fit + 1, nullptr, nullptr) <= DDC_FIXED_SCRATCH_BYTES) ++fit;
        return fail(DDC_E_INVALID,
                    "ddc_forward_fixed_scratch: batch %d needs %zu B of scratch but the built-in array holds %u B "
                    "(at most %d images of this shape); use ddc_forward with a caller-provided work arena",
                    batch, need, DDC_FIXED_SCRATCH_BYTES, fit);
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int dev = pl->device;
    if (dev < 0 || dev >= 64) return fail(DDC_E_INVALID, "device %d", dev);
    CK(cudaSetDevice(dev));
    if (g_fixed_scratch_dev[dev] == nullptr) {
        // first call on this device: resolve the symbol to its address ON THIS DEVICE (must happen with the
        // device current, and outside stream capture)
        void* ptr = nullptr;
        CK(cudaGetSymbolAddress(&ptr, g_fixed_scratch));
        if (ptr == nullptr) return fail(DDC_E_CUDA, "cudaGetSymbolAddress returned a null address");
        g_fixed_scratch_dev[dev] = ptr;
    }
    if (pl->fixed_view_cap < batch) {
        // lay the activations of as many images as fit out in the built-in array and describe them to the TMA
        // unit: host-side bookkeeping inside the plan struct, no heap, no device memory
        int fit = batch;
        while (fixed_scratch_layout(pl, fit * 2, nullptr, nullptr) <= DDC_FIXED_SCRATCH_BYTES && fit < (1 << 20)) fit *= 2;
        ddc_plan::WorkView v;
        fixed_scratch_layout(pl, fit, &v, static_cast<char*>(g_fixed_scratch_dev[dev]));
        const uint64_t C = pl->C, Mp = v.Mpad;
        TRY(make_tmap(&v.tm_Ap, false, v.Ap, Mp, pl->Kp, G2_BM));
        TRY(make_tmap(&v.tm_O, false, v.O, Mp, C, G2_BM));
        TRY(make_tmap(&v.tm_Hd, false, v.Hd, Mp, C, G2_BM));
        TRY(make_tmap(&v.tm_XN, false, v.XN, Mp, C, G2_BM));
        TRY(make_tmap(&v.tm_Xld, true, v.X, Mp, C, G2_BM));
        TRY(make_tmap(&v.tm_XN64, false, v.XN, Mp, C, 64));
        TRY(make_tmap(&v.tm_XNcls, false, v.XN, static_cast<uint64_t>(fit), C, G2_BM, static_cast<uint64_t>(pl->L) * C));
        pl->fixed_view = v;
        pl->fixed_view_cap = fit;
    }
    // Run the ordinary kernel chain with the plan temporarily looking at the built-in array.  Nothing in the
    // array has to start at zero: every row below batch*L of every buffer is written by the kernel that precedes
    // its first reader, and rows above it are computed on but never stored.
    ddc_plan::WorkView own;
    view_save(pl, &own);
    view_install(pl, &pl->fixed_view);
    int rc = enqueue_predictor(pl, x, batch, t, false, st);
    if (rc == DDC_OK) {
        HeadDdimParams q{};
        q.y = pl->Y; q.ldy = pl->P; q.x = out;
        q.H = pl->sh.img_h; q.W = pl->sh.img_w; q.p = pl->sh.patch; q.L = pl->L; q.B = batch;
        const cudaError_t e = launch_head_ddim<MODE_FORWARD>(q, st);
        if (e != cudaSuccess) rc = fail(DDC_E_CUDA, "head_ddim launch: %s", cudaGetErrorString(e));
    }
    view_save(pl, &pl->fixed_view);      // keeps the store descriptors built for this batch
    view_install(pl, &own);
    return rc;
}

static HeadDdimParams step_head_params(ddc_plan* pl, float* x, int B, float* x0_last, float* x0_seq, float* eps_seq) {
    HeadDdimParams q{};
    q.y = pl->Y; q.ldy = pl->P; q.x = x;
    q.H = pl->sh.img_h; q.W = pl->sh.img_w; q.p = pl->sh.patch; q.L = pl->L; q.B = B;
    q.tbl = pl->tbl; q.cur = pl->ctr;
    q.x0_last = x0_last; q.x0_seq = x0_seq; q.eps_seq = eps_seq;
    q.Kp = pl->Kp; q.f16 = pl->f16 ? 1 : 0;
    return q;
}

// In the sampling loop the unpatchify + DDIM pass of a step can also write the NEXT step's patch matrix (it holds the updated
// pixels anyway), which takes patchify -- a 50 MB read and a launch -- out of every step but the first.  Needs the vector
// head_ddim kernel with room for the patch rows, and a 16-bit operand format.
static bool step_emits_patches(ddc_plan* pl, float* x, int B, float* x0_last, float* x0_seq, float* eps_seq) {
    if (pl->f32 || env_int("DDC_STEP_PATCHIFY", 0)) return false;
    HeadDdimParams q = step_head_params(pl, x, B, x0_last, x0_seq, eps_seq);
    q.ap_next = pl->Ap;
    return head_ddim_rows(q, true) > 0;
}

// One denoise step.  emit_patches: head_ddim writes the next step's patch matrix and moves the step counter on, and
// patchify is skipped here (the caller ran it once before the first step).
static int enqueue_step(ddc_plan* pl, float* x, int B, float* x0_last, float* x0_seq, float* eps_seq, cudaStream_t st,
                        Prof* prof = nullptr, bool emit_patches = false) {
    TRY(enqueue_predictor(pl, x, B, nullptr, true, st, prof, emit_patches));
    HeadDdimParams q = step_head_params(pl, x, B, x0_last, x0_seq, eps_seq);
    if (emit_patches) { q.ap_next = pl->Ap; q.ctr_adv = pl->ctr; }
    CK(launch_head_ddim<MODE_STEP>(q, st));
    MARK(KC_DDIM);
    return DDC_OK;
}

// The loop's one patchify (step counter: {0, 1}) when the steps emit their successors' patch matrices
static int enqueue_first_patchify(ddc_plan* pl, const float* x, int B, cudaStream_t st, Prof* prof = nullptr) {
    const ddc_shape& s = pl->sh;
    CK(pl->f16 ? launch_patchify<true>(x, pl->Ap, B, s.img_h, s.img_w, s.patch, pl->L, pl->Kp, pl->ctr, st)
               : launch_patchify<false>(x, pl->Ap, B, s.img_h, s.img_w, s.patch, pl->L, pl->Kp, pl->ctr, st));
    MARK(KC_PATCHIFY);
    return DDC_OK;
}

extern "C" int ddc_profile_steps(ddc_plan* pl, float* x, int32_t batch, const ddc_step* host_table, int32_t n_steps,
                                 float* class_ms, int32_t* class_launches, void* stream) {
    TRY(check_call(pl, batch));
    if (x == nullptr || host_table == nullptr || class_ms == nullptr || class_launches == nullptr)
        return fail(DDC_E_INVALID, "NULL argument");
    if (n_steps <= 0 || n_steps > pl->sh.total_steps) return fail(DDC_E_INVALID, "n_steps %d", n_steps);
    for (int i = 0; i < n_steps; ++i)
        if (host_table[i].t < 0 || host_table[i].t >= pl->sh.total_steps)
            return fail(DDC_E_INVALID, "step %d: t=%d outside the time_embed table", i, host_table[i].t);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(pl->device));
    CK(cudaMemcpyAsync(pl->tbl, host_table, sizeof(StepEntry) * n_steps, cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(pl->ctr, 0, 3 * sizeof(int), st));
    for (int c = 0; c < KC_COUNT; ++c) { class_ms[c] = 0.f; class_launches[c] = 0; }
    Prof prof;
    prof.st = st;
    int rc = DDC_OK;
    const bool emit = step_emits_patches(pl, x, batch, nullptr, nullptr, nullptr);
    for (int i = 0; i < n_steps && rc == DDC_OK; ++i) {
        rc = prof.begin();
        if (rc == DDC_OK && emit && i == 0) rc = enqueue_first_patchify(pl, x, batch, st, &prof);
        if (rc == DDC_OK) rc = enqueue_step(pl, x, batch, nullptr, nullptr, nullptr, st, &prof, emit);
        if (rc != DDC_OK) break;
        const cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { rc = fail(DDC_E_CUDA, "profile: %s", cudaGetErrorString(e)); break; }
        for (size_t j = 0; j < prof.cls.size(); ++j) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, prof.ev[j], prof.ev[j + 1]);
            class_ms[prof.cls[j]] += ms;
            class_launches[prof.cls[j]] += 1;
        }
        prof.release();
    }
    prof.release();
    return rc;
}

extern "C" int ddc_sample(ddc_plan* pl, float* x, int32_t batch, const ddc_step* host_table, int32_t n_steps,
                          float* x0_last, float* x0_seq, float* eps_seq, int32_t use_graph, void* stream) {
    TRY(check_call(pl, batch));
    if (x == nullptr || host_table == nullptr) return fail(DDC_E_INVALID, "NULL tensor");
    if (n_steps <= 0 || n_steps > pl->sh.total_steps)
        return fail(DDC_E_INVALID, "n_steps %d outside 1..%d", n_steps, pl->sh.total_steps);
    for (int i = 0; i < n_steps; ++i)
        if (host_table[i].t < 0 || host_table[i].t >= pl->sh.total_steps)
            return fail(DDC_E_INVALID, "step %d: t=%d outside the time_embed table", i, host_table[i].t);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(pl->device));
    // pageable-source async copies are staged before returning, so host_table may be freed by the caller
    CK(cudaMemcpyAsync(pl->tbl, host_table, sizeof(StepEntry) * n_steps, cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(pl->ctr, 0, 3 * sizeof(int), st));
    const bool emit = step_emits_patches(pl, x, batch, x0_last, x0_seq, eps_seq);
    if (!use_graph) {
        if (emit) TRY(enqueue_first_patchify(pl, x, batch, st));
        for (int i = 0; i < n_steps; ++i) TRY(enqueue_step(pl, x, batch, x0_last, x0_seq, eps_seq, st, nullptr, emit));
        return DDC_OK;
    }
    if (st == nullptr) return fail(DDC_E_INVALID, "graph replay needs a non-default stream (capture is illegal on stream 0)");
    const GraphKey key{batch, x, x0_last, x0_seq, eps_seq};
    if (pl->gexec == nullptr || !(pl->gkey == key)) {
        if (pl->gexec) {
            cudaGraphExecDestroy(pl->gexec);
            pl->gexec = nullptr;
        }
        cudaGraph_t graph = nullptr;
        CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        const int r = enqueue_step(pl, x, batch, x0_last, x0_seq, eps_seq, st, nullptr, emit);
        const cudaError_t e = cudaStreamEndCapture(st, &graph);
        if (r != DDC_OK) {
            if (graph) cudaGraphDestroy(graph);
            return r;
        }
        if (e != cudaSuccess) return fail(DDC_E_CUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(e));
        const cudaError_t ei = cudaGraphInstantiate(&pl->gexec, graph, 0);
        cudaGraphDestroy(graph);
        if (ei != cudaSuccess) {
            pl->gexec = nullptr;
            return fail(DDC_E_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(ei));
        }
        pl->gkey = key;
    }
    if (emit) TRY(enqueue_first_patchify(pl, x, batch, st));      // the captured step has no patchify of its own
    for (int i = 0; i < n_steps; ++i) CK(cudaGraphLaunch(pl->gexec, st));
    return DDC_OK;
}

// ------------------------------------------------------------------------------- schedules (host)
extern "C" int ddc_ddim_table(int32_t k, int32_t total_steps, int32_t t_start, ddc_step* tab, int32_t capacity,
                              int32_t* n_steps) {
    if (tab == nullptr || n_steps == nullptr) return fail(DDC_E_INVALID, "NULL argument");
    if (k <= 0 || total_steps <= 0 || t_start < 0 || t_start >= total_steps)
        return fail(DDC_E_INVALID, "k=%d total_steps=%d t_start=%d", k, total_steps, t_start);
    const double T = static_cast<double>(total_steps);
    int n = 0;
    for (int t = t_start; t > 0; t -= k) {
        if (t + 1 - k < 0) return fail(DDC_E_DOMAIN, "math domain error");          // sqrt(<0) at ViT.py:231
        const double a_tk = 1.0 - sqrt((t + 1 - k) / T);
        const double a_t = 1.0 - sqrt((t + 1) / T) + 1e-5;
        if (a_t <= 0.0 || a_t >= 1.0) return fail(DDC_E_DOMAIN, "math domain error");
        if (a_tk == 0.0) return fail(DDC_E_DOMAIN, "float division by zero");        // ViT.py:234 (1-a_tk)/a_tk
        if (a_tk < 0.0) return fail(DDC_E_DOMAIN, "math domain error");
        if (n >= capacity) return fail(DDC_E_INVALID, "table capacity %d too small", capacity);
        const double A = sqrt(1.0 - a_tk) / sqrt(1.0 - a_t);
        const double Bc = sqrt(a_tk) - A * sqrt(a_t);
        ddc_step e;
        memset(&e, 0, sizeof(e));
        e.t = t;
        e.a = static_cast<float>(A);
        e.b = static_cast<float>(Bc);
        e.sqrt_at = static_cast<float>(sqrt(a_t));
        e.inv_sqrt_1mat = static_cast<float>(1.0 / sqrt(1.0 - a_t));
        tab[n++] = e;
    }
    if (n == 0) return fail(DDC_E_INVALID, "empty schedule");
    tab[n - 1].flags |= 1;
    *n_steps = n;
    return DDC_OK;
}

extern "C" int ddc_cold_table(ddc_step* tab, int32_t capacity, int32_t* n_steps) {
    if (tab == nullptr || n_steps == nullptr) return fail(DDC_E_INVALID, "NULL argument");
    if (capacity < 6) return fail(DDC_E_INVALID, "table capacity %d too small", capacity);
    for (int i = 0; i < 6; ++i) {
        memset(&tab[i], 0, sizeof(ddc_step));
        tab[i].t = 6 - i;
        tab[i].a = 0.f;
        tab[i].b = 1.f;                 // x <- clamp(f(x, t))
        tab[i].sqrt_at = 1.f;
        tab[i].inv_sqrt_1mat = 1.f;
    }
    tab[5].flags |= 1;
    *n_steps = 6;
    return DDC_OK;
}

// ------------------------------------------------------------------------------- kernel-level entries
extern "C" int ddc_k_gemm(int32_t fmt, int32_t epi, int32_t M, int32_t N, int32_t K, const void* A, const void* W,
                          const float* bias, void* out, int32_t ld_out, void* stream) {
    const bool f16 = fmt == 1;
    const int cs = gemm_cluster_size(M);
    if (A == nullptr || W == nullptr || bias == nullptr || out == nullptr) return fail(DDC_E_INVALID, "NULL tensor");
    if (M <= 0 || N <= 0 || K <= 0 || K % 64 != 0 || ld_out != N) return fail(DDC_E_INVALID, "bad sizes");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const uint64_t Mp = (static_cast<uint64_t>(M) + G2_BM - 1) / G2_BM * G2_BM;
    const int bn = (epi == S_F32) ? pick_bn_f32(N) : pick_bn_stream(N);
    if (bn == 0) return fail(DDC_E_INVALID, "N=%d must be a multiple of %d", N, epi == S_F32 ? 16 : 64);
    CUtensorMap tmA, tmB, tmO;
    TRY(make_tmap(&tmA, false, A, Mp, K, G2_BM));
    TRY(make_tmap(&tmB, false, W, N, K, bn / cs));
    Gemm2Params p{};
    p.M = M; p.N = N; p.K = K; p.bn = bn; p.stages = pick_stages2(K, N, bn / cs, false); p.bias = bias;
    switch (epi) {
        case S_BF16:
            TRY(make_tmap(&tmO, false, out, M, N, G2_BM));
            return launch_gemm2<S_BF16>(tmA, tmB, tmO, tmA, tmA, p, f16, cs, st);
        case S_BF16_GELU:
            TRY(make_tmap(&tmO, false, out, M, N, G2_BM));
            return launch_gemm2<S_BF16_GELU>(tmA, tmB, tmO, tmA, tmA, p, f16, cs, st);
        case S_F32:
            p.out_f32 = static_cast<float*>(out); p.ld_f32 = ld_out;
            if (N % 32 == 0 && bn % 32 == 0) {
                p.f32_tma = 1;
                TRY(make_tmap(&tmO, true, out, M, N, G2_BM));
                return launch_gemm2<S_F32>(tmA, tmB, tmO, tmA, tmA, p, f16, cs, st);
            }
            return launch_gemm2<S_F32>(tmA, tmB, tmA, tmA, tmA, p, f16, cs, st);
    }
    return fail(DDC_E_INVALID, "epi=%d", epi);
}

extern "C" int ddc_k_gemm_resid_ln(int32_t fmt, int32_t M, int32_t C, int32_t K, const void* A, const void* W, const float* bias,
                                   float* x, const float* ln_g, const float* ln_b, void* xn, void* stream) {
    if (A == nullptr || W == nullptr || bias == nullptr || x == nullptr || ln_g == nullptr || ln_b == nullptr ||
        xn == nullptr)
        return fail(DDC_E_INVALID, "NULL tensor");
    if (M <= 0 || C <= 0 || K <= 0 || K % 64 != 0 || C % 64 != 0 || C > 512) return fail(DDC_E_INVALID, "bad sizes");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const uint64_t Mp = (static_cast<uint64_t>(M) + G2_BM - 1) / G2_BM * G2_BM;
    const int bn = pick_bn_row(C);
    CUtensorMap tmA, tmB, tmO, tmR, tmN;
    TRY(make_tmap(&tmA, false, A, Mp, K, G2_BM));
    const int cs = gemm_cluster_size(M);
    TRY(make_tmap(&tmB, false, W, C, K, bn / cs));
    TRY(make_tmap(&tmO, true, x, M, C, G2_BM));
    TRY(make_tmap(&tmR, true, x, Mp, C, G2_BM));
    TRY(make_tmap(&tmN, false, xn, M, C, G2_BM));
    Gemm2Params p{};
    p.M = M; p.N = C; p.K = K; p.bn = bn; p.stages = pick_stages2(K, C, bn / cs, true); p.bias = bias;
    p.ln_g = ln_g; p.ln_b = ln_b; p.ln_eps = 1e-5f;
    return launch_gemm2<R_RESID>(tmA, tmB, tmO, tmR, tmN, p, fmt == 1, cs, st);
}

extern "C" int ddc_k_embed(int32_t fmt, int32_t batch, int32_t L, int32_t C, int32_t Kp, int32_t total_steps, const void* ap,
                           const void* w_patch, const float* posb, const float* time_tab, const int32_t* t_img,
                           const float* ln_g, const float* ln_b, float* x, void* xn, void* stream) {
    if (!ap || !w_patch || !posb || !time_tab || !t_img || !ln_g || !ln_b || !x || !xn) return fail(DDC_E_INVALID, "NULL tensor");
    if (batch <= 0 || L <= 1 || C <= 0 || C % 64 != 0 || C > 512 || Kp <= 0 || Kp % 64 != 0 || Kp > 512 || total_steps <= 0)
        return fail(DDC_E_INVALID, "embed: bad sizes (batch=%d L=%d C=%d Kp=%d)", batch, L, C, Kp);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int M = batch * L;
    const uint64_t Mp = (static_cast<uint64_t>(M) + G2_BM - 1) / G2_BM * G2_BM;
    const int bn = pick_bn_row(C);
    const int cs = gemm_cluster_size(M);
    CUtensorMap tmA, tmB, tmO, tmR, tmN;
    TRY(make_tmap(&tmA, false, ap, Mp, Kp, G2_BM));
    TRY(make_tmap(&tmB, false, w_patch, C, Kp, bn / cs));
    TRY(make_tmap(&tmO, true, x, M, C, G2_BM));
    TRY(make_tmap(&tmR, true, posb, static_cast<uint64_t>(L) + 127, C, G2_BM));
    TRY(make_tmap(&tmN, false, xn, M, C, G2_BM));
    Gemm2Params g{};
    g.M = M; g.N = C; g.K = Kp; g.bn = bn; g.stages = pick_stages2(Kp, C, bn / cs, true);
    g.ln_g = ln_g; g.ln_b = ln_b; g.ln_eps = 1e-5f;
    g.posb = posb; g.L = L; g.time_tab = time_tab; g.t_img = t_img; g.B = batch; g.T = total_steps;
    return launch_gemm2<R_EMBED>(tmA, tmB, tmO, tmR, tmN, g, fmt == 1, cs, st);
}

extern "C" int ddc_k_block_tail(int32_t fmt, int32_t M, int32_t C, const void* O, const void* Wproj, const float* b_proj,
                                const void* Wfc1, const float* b_fc1, const void* Wfc2, const float* b_fc2, float* x, void* xn,
                                void* stream) {
    if (!O || !Wproj || !b_proj || !Wfc1 || !b_fc1 || !Wfc2 || !b_fc2 || !x || !xn) return fail(DDC_E_INVALID, "NULL tensor");
    if (M < 129 || C % 128 != 0 || C > 384 || C <= 0)
        return fail(DDC_E_INVALID, "block tail needs M > 128 and embed_dim in {128, 256, 384} (M=%d C=%d)", M, C);
    const uint64_t Mp = (static_cast<uint64_t>(M) + G2_BM - 1) / G2_BM * G2_BM;
    CUtensorMap tmO, tmWp, tmW1, tmW2, tmXld, tmXst, tmXN;
    TRY(make_tmap(&tmO, false, O, Mp, C, G2_BM));
    TRY(make_tmap(&tmWp, false, Wproj, C, C, 64));
    TRY(make_tmap(&tmW1, false, Wfc1, C, C, 64));
    TRY(make_tmap(&tmW2, false, Wfc2, C, C, 64));
    TRY(make_tmap(&tmXld, true, x, Mp, C, G2_BM));
    TRY(make_tmap(&tmXst, true, x, M, C, G2_BM));
    TRY(make_tmap(&tmXN, false, xn, M, C, G2_BM));
    BlockTailParams bt{};
    bt.M = M; bt.C = C; bt.ln_eps = 1e-5f;
    bt.b_proj = b_proj; bt.b_fc1 = b_fc1; bt.b_fc2 = b_fc2;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (fmt == 1) return launch_block_tail_inst<true>(tmO, tmWp, tmW1, tmW2, tmXld, tmXst, tmXN, bt, st);
    return launch_block_tail_inst<false>(tmO, tmWp, tmW1, tmW2, tmXld, tmXst, tmXN, bt, st);
}

extern "C" int ddc_k_attention(int32_t fmt, const void* qkv, void* out, int32_t batch, int32_t L, int32_t C,
                               int32_t heads, void* stream) {
    if (qkv == nullptr || out == nullptr || batch <= 0 || L <= 0) return fail(DDC_E_INVALID, "bad argument");
    return run_attention(static_cast<const bf16*>(qkv), static_cast<bf16*>(out), batch, L, C, heads, fmt == 1,
                         static_cast<cudaStream_t>(stream));
}

extern "C" int ddc_k_qkv_attn(int32_t fmt, const void* xn, const void* w_qkv, const float* b_qkv, void* out, int32_t batch, int32_t C,
                              int32_t heads, void* w_hp, float* b_hp, void* qkv_cls, void* stream) {
    if (!xn || !w_qkv || !b_qkv || !out || !w_hp || !b_hp || !qkv_cls || batch <= 0) return fail(DDC_E_INVALID, "bad argument");
    if (!qkv_attn_usable(QA_L, C, heads)) return fail(DDC_E_INVALID, "qkv+attention: needs 65 tokens, head_dim 32, embed_dim %% 64 == 0 and <= 384 (C=%d heads=%d)", C, heads);
    const bool f16 = fmt == 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int L = QA_L, M = batch * L;
    const uint64_t Mp = (static_cast<uint64_t>(M) + G2_BM - 1) / G2_BM * G2_BM;
    repack_qkv_hp_kernel<<<296, 64, 0, st>>>(static_cast<const bf16*>(w_qkv), b_qkv, static_cast<bf16*>(w_hp), b_hp, C);
    CK(cudaGetLastError());
    CUtensorMap tmCls, tmW, tmO, tmA64, tmWhp;
    const int cs = gemm_cluster_size(batch);
    const int bn = pick_bn_stream(3 * C);
    TRY(make_tmap(&tmCls, false, xn, static_cast<uint64_t>(batch), C, G2_BM, static_cast<uint64_t>(L) * C));
    TRY(make_tmap(&tmW, false, w_qkv, 3 * static_cast<uint64_t>(C), C, bn / cs));
    TRY(make_tmap(&tmO, false, qkv_cls, static_cast<uint64_t>(batch), 3 * static_cast<uint64_t>(C), G2_BM));
    Gemm2Params q{};
    q.M = batch; q.N = 3 * C; q.K = C; q.bn = bn; q.stages = pick_stages2(C, 3 * C, bn / cs, false); q.bias = b_qkv;
    TRY((launch_gemm2<S_BF16>(tmCls, tmW, tmO, tmCls, tmCls, q, f16, cs, st)));
    TRY(make_tmap(&tmA64, false, xn, Mp, C, 64));
    TRY(make_tmap(&tmWhp, false, w_hp, 3 * static_cast<uint64_t>(C), C, 96));
    QkvAttnParams qa{};
    qa.B = batch; qa.C = C; qa.bias_hp = b_hp; qa.qkv_cls = static_cast<const bf16*>(qkv_cls); qa.out = static_cast<bf16*>(out);
    return f16 ? launch_qkv_attn<true>(tmA64, tmWhp, qa, st) : launch_qkv_attn<false>(tmA64, tmWhp, qa, st);
}

extern "C" int ddc_dbg_qkv_attn_trace(void* dev_buffer) {
    g_qa_trace = static_cast<long long*>(dev_buffer);
    return DDC_OK;
}

extern "C" int ddc_k_patchify(int32_t fmt, const float* x, void* ap, int32_t batch, int32_t H, int32_t W, int32_t patch, int32_t Kp,
                              void* stream) {
    if (x == nullptr || ap == nullptr || batch <= 0 || patch <= 0 || patch % 2 || H % patch || W % patch ||
        Kp < 3 * patch * patch || Kp % 2)
        return fail(DDC_E_INVALID, "bad argument");
    const int L = (H / patch) * (W / patch) + 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CK(fmt == 1 ? launch_patchify<true>(x, static_cast<bf16*>(ap), batch, H, W, patch, L, Kp, nullptr, st)
                : launch_patchify<false>(x, static_cast<bf16*>(ap), batch, H, W, patch, L, Kp, nullptr, st));
    return DDC_OK;
}

extern "C" int ddc_k_smooth_l1(const float* pred, const float* target, size_t n, float* scratch, float* out, void* stream) {
    if (pred == nullptr || target == nullptr || scratch == nullptr || out == nullptr) return fail(DDC_E_INVALID, "NULL tensor");
    if ((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(target)) & 15) return fail(DDC_E_INVALID, "smooth_l1: 16-byte alignment");
    if (n == 0) return fail(DDC_E_INVALID, "smooth_l1: empty tensors (the reference's mean over nothing is NaN)");
    size_t blocks = (n / 4 + 255) / 256;
    const size_t cap = static_cast<size_t>(num_sms()) * 4;          // <= 1023 block sums fit the scratch
    if (blocks > cap) blocks = cap;
    if (blocks > 1023) blocks = 1023;
    if (blocks == 0) blocks = 1;
    smooth_l1_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(pred, target, n, scratch, out);
    CK(cudaGetLastError());
    return DDC_OK;
}

extern "C" int ddc_k_affine(const float* in, float* out, size_t n, float add, float mul, void* stream) {
    if (in == nullptr || out == nullptr) return fail(DDC_E_INVALID, "NULL tensor");
    if ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) return fail(DDC_E_INVALID, "affine: 16-byte alignment");
    if (n == 0) return DDC_OK;
    size_t blocks = (n / 4 + 255) / 256;
    const size_t cap = static_cast<size_t>(num_sms()) * 8;
    if (blocks > cap) blocks = cap;
    if (blocks == 0) blocks = 1;
    affine_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(in, out, n, add, mul);
    CK(cudaGetLastError());
    return DDC_OK;
}

extern "C" int ddc_k_mix(const float* z, const float* draft, float* out, size_t n, size_t period, float a, float b, void* stream) {
    if (z == nullptr || draft == nullptr || out == nullptr || period == 0) return fail(DDC_E_INVALID, "bad argument");
    if (n == 0) return DDC_OK;
    size_t blocks = (n + 255) / 256;
    const size_t cap = static_cast<size_t>(num_sms()) * 16;
    if (blocks > cap) blocks = cap;
    mix_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(z, draft, out, n, period, a, b);
    CK(cudaGetLastError());
    return DDC_OK;
}

extern "C" int ddc_k_head_ddim(const float* y, int32_t ldy, float* x, int32_t batch, int32_t H, int32_t W, int32_t patch,
                               const ddc_step* host_step, float* x0_out, float* eps_out, void* stream) {
    if (y == nullptr || x == nullptr || host_step == nullptr || batch <= 0 || patch <= 0 || H % patch || W % patch)
        return fail(DDC_E_INVALID, "bad argument");
    HeadDdimParams q{};
    q.y = y; q.ldy = ldy; q.x = x; q.H = H; q.W = W; q.p = patch; q.L = (H / patch) * (W / patch) + 1; q.B = batch;
    q.tbl = nullptr; q.cur = nullptr;
    memcpy(&q.one, host_step, sizeof(StepEntry));
    q.one.flags |= 1;
    q.x0_last = x0_out; q.x0_seq = nullptr; q.eps_seq = eps_out;
    CK(launch_head_ddim<MODE_STEP>(q, static_cast<cudaStream_t>(stream)));
    return DDC_OK;
}

// Shared device helpers for the sm_100a kernels: mbarrier, TMA, tcgen05/TMEM and
// mma.sync wrappers written as inline PTX.  Nothing here is portable to other
// architectures on purpose (compile with -gencode arch=compute_100a,code=sm_100a).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <utility>

namespace ddc {

typedef __nv_bfloat16 bf16;

// One row of the per-step table the sampling loop reads on the device.  The host fills
// it in double precision from the reference's schedule (ViT.py:231-232) and the kernels
// index it with the device-side step counter, so one captured graph replays every step.
struct StepEntry {
    int t;           // timestep fed to time_embed (ViT.py:227)
    int flags;       // bit0: last step -> also write x0 to the result buffer
    float a;         // x_{t-k} = a * x_t + b * x0   (collapsed ViT.py:233-234)
    float b;
    float sqrt_at;   // eps = (x_t - sqrt_at * x0) * inv_sqrt_1mat
    float inv_sqrt_1mat;
    float pad0, pad1;
};

#ifndef DDC_SPIN_LIMIT
#define DDC_SPIN_LIMIT (1u << 24)   // bounded waits: a protocol bug traps instead of hanging the GPU
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ----------------------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > DDC_SPIN_LIMIT) {
            printf("ddimcold: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
            __trap();
        }
    }
}

// The same wait with a suspend-time hint: the hardware parks the thread until the phase completes (or the hint, in
// nanoseconds, runs out) instead of returning at once, so a waiting warp does not compete for issue slots with the
// warps that work.  For kernels in which many warps of an SM wait on each other most of the time.
__device__ __forceinline__ void mbar_wait_parked(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0, ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.b32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity), "r"(20000u)
            : "memory");
        if (!ok && ++spins > (DDC_SPIN_LIMIT >> 6)) {
            printf("ddimcold: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
            __trap();
        }
    } while (!ok);
}

// ----------------------------------------------------------------------------- proxies / fences
__device__ __forceinline__ void fence_proxy_async_smem() {
    // generic-proxy smem writes -> visible to the async proxy (TMA / tcgen05.mma operand reads)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ----------------------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 2-D tiled load: coordinates are {c0 = innermost (K) element index, c1 = row index}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}

// L2 eviction-priority hints for TMA traffic: activations that are touched once stream through (evict_first) so that
// they do not push the weights -- re-read by every CTA all the time -- out of L2 (evict_last).
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_normal() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void tma_load_2d_hint(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, uint64_t pol) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(pol)
        : "memory");
}

// 3-D tiled load {c0 = element within a 64-wide K block, c1 = row, c2 = K block}: one instruction fetches several
// K blocks (a TMA instruction occupies its issuing thread for ~300 cycles whatever its size, so few large loads
// beat many small ones).
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// Bring a box into L2 only (no shared memory, no barrier): used to turn a later DRAM-latency TMA load into an
// L2 hit when the address is known long before the data is needed.
__device__ __forceinline__ void tma_prefetch_l2_2d(const CUtensorMap* m, int c0, int c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global [%0, {%1, %2}];"
                 ::"l"(reinterpret_cast<uint64_t>(m)), "r"(c0), "r"(c1)
                 : "memory");
}
// CTA-pair load: the data lands in THIS CTA's shared memory, the transaction bytes are credited to the
// barrier at the same offset in the pair's leader CTA (even rank), which is the one that issues the MMAs.
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;      // clears the CTA-rank bit of a shared::cluster address
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair_hint(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, uint64_t pol) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "l"(pol)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d_pair(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1),
          "r"(c2)
        : "memory");
}
// Arrive on the barrier at this offset in the pair's leader CTA (works from either CTA of the pair).
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
// Semantics: the default (.release at CTA scope).  A cluster-scope release also has to publish this thread's prior
// GLOBAL writes to the peer SM, i.e. a round trip to L2: measured ~750 cycles per arrive for the issuing thread
// (scripts/probes/stream_timeline.cu), which made the epilogue the bottleneck of every pair kernel.  Nothing handed
// over through these barriers lives in global memory: TMEM hand-backs are ordered by tcgen05.fence, shared-memory
// operands by fence.proxy.async in the writing CTA before the arrive.
#ifndef DDC_ARRIVE_SEM
#define DDC_ARRIVE_SEM ""
#endif
    asm volatile("mbarrier.arrive" DDC_ARRIVE_SEM ".shared::cluster.b64 _, [%0];" ::"r"(smem_u32(bar) & kPeerBitMask) : "memory");
}

// 2-D tiled store smem -> global (bulk async group); rows / columns outside the tensor map are clipped
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_store_2d_hint(const CUtensorMap* m, const void* smem_src, int c0, int c1, uint64_t pol) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group.L2::cache_hint [%0, {%2, %3}], [%1], %4;"
                 ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "l"(pol)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {   // at most N committed groups still reading their smem source
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ----------------------------------------------------------------------------- TMEM
// CG = 1: this CTA's tensor memory.  CG = 2: the same columns in both CTAs of a pair (issued by the
// same warp id in each CTA, writing the same shared-memory slot offset).
template <int CG = 1>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
    if (CG == 2)
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)),
                     "r"(ncols) : "memory");
    else
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)),
                     "r"(ncols) : "memory");
}
template <int CG = 1>
__device__ __forceinline__ void tmem_relinquish() {
    if (CG == 2) asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    else asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int CG = 1>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    if (CG == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// Shared-memory matrix descriptor for a K-major bf16 operand tile stored as rows of 128 B
// with the 128-byte swizzle (what TMA SWIZZLE_128B writes): 8-row groups are 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);   // start address, bits [0,14)
    d |= static_cast<uint64_t>(1) << 16;                        // leading byte offset (unused for SW128 K-major)
    d |= static_cast<uint64_t>(1024 >> 4) << 32;                // stride byte offset: 8 rows * 128 B
    d |= static_cast<uint64_t>(1) << 46;                        // descriptor version (Blackwell)
    d |= static_cast<uint64_t>(2) << 61;                        // layout: SWIZZLE_128B
    return d;
}

// Instruction descriptor for kind::f16 with fp32 accumulation, both operands K-major.
// `f16` selects IEEE half operands (format 0) instead of bfloat16 (format 1).
__host__ __device__ __forceinline__ uint32_t umma_idesc_16bit(uint32_t m, uint32_t n, bool f16) {
    const uint32_t fmt = f16 ? 0u : 1u;
    return (1u << 4)            // D format: f32
           | (fmt << 7)         // A format
           | (fmt << 10)        // B format
           | ((n >> 3) << 17)   // N, 3 LSBs dropped
           | ((m >> 4) << 24);  // M, 4 LSBs dropped
}

__device__ __forceinline__ void umma_bf16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// One K block (64 elements = four K=16 instructions) of D[tmem] (+)= A * B.  The descriptors address the
// first 32-byte K slice; consecutive slices are +32 B = +2 in the descriptor's 16-byte address units.
// A single asm statement with warp-uniform operands lets ptxas keep everything in uniform registers
// and issue the four UTCHMMA back to back (no per-instruction elect/broadcast sequence).
// CG = 2: one instruction drives both SMs of a CTA pair (M = 256): each CTA supplies its own 128 rows of
// A and half of the N rows of B from the same shared-memory offsets, and receives its 128 rows of D.
#define DDC_UMMA_KBLOCK_ASM(CGS)                                                           \
    asm volatile(                                                                          \
        "{\n\t.reg .pred p, q;\n\t.reg .b64 a1, b1;\n\t"                                   \
        "setp.ne.b32 p, %4, 0;\n\t"                                                        \
        "setp.eq.b32 q, 0, 0;\n\t"                                                         \
        "tcgen05.mma.cta_group::" CGS ".kind::f16 [%0], %1, %2, %3, p;\n\t"                \
        "add.u64 a1, %1, 2;\n\tadd.u64 b1, %2, 2;\n\t"                                     \
        "tcgen05.mma.cta_group::" CGS ".kind::f16 [%0], a1, b1, %3, q;\n\t"                \
        "add.u64 a1, %1, 4;\n\tadd.u64 b1, %2, 4;\n\t"                                     \
        "tcgen05.mma.cta_group::" CGS ".kind::f16 [%0], a1, b1, %3, q;\n\t"                \
        "add.u64 a1, %1, 6;\n\tadd.u64 b1, %2, 6;\n\t"                                     \
        "tcgen05.mma.cta_group::" CGS ".kind::f16 [%0], a1, b1, %3, q;\n\t}"               \
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate_first)           \
        : "memory")
template <int CG = 1>
__device__ __forceinline__ void umma_kblock(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate_first) {
    if (CG == 2) DDC_UMMA_KBLOCK_ASM("2");
    else DDC_UMMA_KBLOCK_ASM("1");
}
// Attention P*V over one K block of 64 keys (four K = 16 steps), two accumulators per step: O (+)= P * V with V an
// MN-major operand whose 16-key slices lie 2048 B apart (+128 descriptor units), and the row sums (+)= P * ones.  One asm
// statement, uniform operands: the eight UTCHMMA issue back to back (a loop of single-MMA statements cost ~100 cycles
// per instruction in descriptor arithmetic and register-to-uniform moves).
__device__ __forceinline__ void umma_pv_kblock(uint32_t tmem_o, uint32_t tmem_sum, uint64_t pdesc, uint64_t vdesc, uint64_t onesdesc,
                                               uint32_t idesc_o, uint32_t idesc_1, uint32_t accumulate_first) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t.reg .b64 a1, b1;\n\t"
        "setp.ne.b32 p, %7, 0;\n\t"
        "setp.eq.b32 q, 0, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %2, %3, %5, p;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%1], %2, %4, %6, p;\n\t"
        "add.u64 a1, %2, 2;\n\tadd.u64 b1, %3, 128;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %5, q;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%1], a1, %4, %6, q;\n\t"
        "add.u64 a1, %2, 4;\n\tadd.u64 b1, %3, 256;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %5, q;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%1], a1, %4, %6, q;\n\t"
        "add.u64 a1, %2, 6;\n\tadd.u64 b1, %3, 384;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %5, q;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%1], a1, %4, %6, q;\n\t}"
        ::"r"(tmem_o), "r"(tmem_sum), "l"(pdesc), "l"(vdesc), "l"(onesdesc), "r"(idesc_o), "r"(idesc_1), "r"(accumulate_first)
        : "memory");
}
// Q K^T over a head of 32 dims: two K = 16 steps, fresh accumulator
__device__ __forceinline__ void umma_k32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t.reg .b64 a1, b1;\n\t"
        "setp.ne.b32 p, 0, 0;\n\t"
        "setp.eq.b32 q, 0, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "add.u64 a1, %1, 2;\n\tadd.u64 b1, %2, 2;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, q;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc)
        : "memory");
}
// True in exactly one lane of the (fully active) warp; ptxas recognises the pattern and needs no
// per-instruction leader election for the single-thread tcgen05 / TMA instructions that follow.
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
// Arrive on an mbarrier once every tcgen05 op this thread issued so far has completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

// CTA-pair commit: arrives on the barrier at this offset in BOTH CTAs once every pair-wide tcgen05 op
// issued so far by this thread has completed.
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
        ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3))
        : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// TMEM -> registers: this warp's 32 lanes x 16 consecutive fp32 columns (thread i = lane i).
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// 32 columns per thread; the wait sits in the same asm statement so no use can be hoisted above it
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
// The same load without the wait, writing straight into the caller's floats: several can be in flight, then ONE
// tmem_ld_wait, whose "+f" operands tie every destination register to the wait so no use is scheduled above it.
__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, float* v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]),
          "=f"(v[8]), "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15]),
          "=f"(v[16]), "=f"(v[17]), "=f"(v[18]), "=f"(v[19]), "=f"(v[20]), "=f"(v[21]), "=f"(v[22]), "=f"(v[23]),
          "=f"(v[24]), "=f"(v[25]), "=f"(v[26]), "=f"(v[27]), "=f"(v[28]), "=f"(v[29]), "=f"(v[30]), "=f"(v[31])
        : "r"(taddr)
        : "memory");
}
// waits for all outstanding tcgen05.ld of this thread; v = the 128 floats they target
__device__ __forceinline__ void tmem_ld_wait(float* v) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 128; i += 8)
        asm volatile("" : "+f"(v[i]), "+f"(v[i + 1]), "+f"(v[i + 2]), "+f"(v[i + 3]), "+f"(v[i + 4]), "+f"(v[i + 5]), "+f"(v[i + 6]), "+f"(v[i + 7]));
}
__device__ __forceinline__ void tmem_ld_wait32(float* v) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; i += 8)
        asm volatile("" : "+f"(v[i]), "+f"(v[i + 1]), "+f"(v[i + 2]), "+f"(v[i + 3]), "+f"(v[i + 4]), "+f"(v[i + 5]), "+f"(v[i + 6]), "+f"(v[i + 7]));
}
__device__ __forceinline__ void tmem_ld_wait64(float* v) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 64; i += 8)
        asm volatile("" : "+f"(v[i]), "+f"(v[i + 1]), "+f"(v[i + 2]), "+f"(v[i + 3]), "+f"(v[i + 4]), "+f"(v[i + 5]), "+f"(v[i + 6]), "+f"(v[i + 7]));
}
// registers -> TMEM: this warp's 32 lanes x 32 consecutive fp32 columns
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float* v) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
        "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};\n\t"
        "tcgen05.wait::st.sync.aligned;"
        ::"r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
          "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
          "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])), "r"(__float_as_uint(v[8])),
          "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
          "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])),
          "r"(__float_as_uint(v[15])), "r"(__float_as_uint(v[16])), "r"(__float_as_uint(v[17])),
          "r"(__float_as_uint(v[18])), "r"(__float_as_uint(v[19])), "r"(__float_as_uint(v[20])),
          "r"(__float_as_uint(v[21])), "r"(__float_as_uint(v[22])), "r"(__float_as_uint(v[23])),
          "r"(__float_as_uint(v[24])), "r"(__float_as_uint(v[25])), "r"(__float_as_uint(v[26])),
          "r"(__float_as_uint(v[27])), "r"(__float_as_uint(v[28])), "r"(__float_as_uint(v[29])),
          "r"(__float_as_uint(v[30])), "r"(__float_as_uint(v[31]))
        : "memory");
}

// ----------------------------------------------------------------------------- legacy tensor path (attention)
__device__ __forceinline__ void ldmatrix_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3,
                                                  uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
// D(16x8, fp32) += A(16x16, row) * B(16x8, col), operands bf16 or fp16
template <bool F16>
__device__ __forceinline__ void mma_16816(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
    if (F16) {
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    } else {
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }
}
__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// two fp32 -> one packed pair of 16-bit operands (round to nearest even; fp16 saturates to finite)
template <bool F16>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    uint32_t r;
    if (F16) {
        asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    } else {
        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    }
    return r;
}

// four 8x8 b16 matrices from registers to shared memory (lane l supplies the address of row l%8 of matrix l/8);
// after ldmatrix.trans this writes the transposed tiles
__device__ __forceinline__ void stmatrix_x4(uint32_t addr, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
    asm volatile("stmatrix.sync.aligned.m8n8.x4.shared.b16 [%0], {%1, %2, %3, %4};"
                 ::"r"(addr), "r"(r0), "r"(r1), "r"(r2), "r"(r3) : "memory");
}

// 16-bit pair / scalar -> fp32 (bf16: the value IS the upper half of the fp32 word)
template <bool F16>
__device__ __forceinline__ float2 unpack2(uint32_t w) {
    if (F16) return __half22float2(*reinterpret_cast<const __half2*>(&w));
    return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xFFFF0000u));
}
template <bool F16>
__device__ __forceinline__ float unpack1(uint16_t v) {
    if (F16) return __half2float(*reinterpret_cast<const __half*>(&v));
    return __uint_as_float(static_cast<uint32_t>(v) << 16);
}

// Exact-erf GELU (nn.GELU() default, approximate='none'; ViT.py:75,80) -- the erf form, not the tanh one:
//     gelu(x) = x Phi(x) = relu(x) - |x| Phi(-|x|),        Phi(-u) = erfc(u / sqrt 2) / 2 = 2^q(u)
// with q a quartic fitted (scripts/fit_gelu.py, minimax on [0, 5.5]) so that |u (2^q(u) - Phi(-u))| <= 6.1e-6, i.e. the
// GELU error is <= 6.2e-6 absolute everywhere (fp32 evaluation included), far below the 16-bit rounding of the result;
// beyond u = 5.5 the correction term is < 1.1e-7 and u is clamped.  Seven FP32 instructions and ONE MUFU per element:
// the MLP epilogue of the block tail is bound by instruction issue and by the MUFU pipe (16 results / clk / SM) alike, and
// the usual erf recipes (Abramowitz-Stegun 7.1.25/26: a reciprocal and an exponential) cost two MUFU.
__device__ __forceinline__ float gelu_erf(float x) {
    float u;                                       // min(|x|, 5.5); the .NaN form so that a NaN input stays a NaN
    asm("min.NaN.f32 %0, %1, %2;" : "=f"(u) : "f"(fabsf(x)), "f"(5.5f));
    float q = fmaf(0.0038656180f, u, -0.044076372f);
    q = fmaf(q, u, -0.46801907f);
    q = fmaf(q, u, -1.1473700f);
    q = fmaf(q, u, -1.0004799f);
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(q));
    return fmaf(-u, e, fmaxf(x, 0.f));
}


// ---------------------------------------------------------------------------- programmatic dependent launch
// The 25 kernels of a denoise step form a strict chain.  Launched plainly, each one starts only after its predecessor
// has fully drained, and then spends its first microseconds on set-up (barrier init, TMEM allocation, descriptor
// prefetch, parameters into shared memory) with the GPU otherwise idle.  With the programmatic-serialization launch
// attribute a kernel's CTAs are scheduled as soon as SMs free up in the predecessor's tail; they run their set-up and
// then block in pdl_wait() until the predecessor has completed and its writes are visible.  Rules kept by every
// step-path kernel: (1) pdl_launch_dependents() first thing; (2) nothing produced on the step path is read, and
// nothing at all is written to global memory, before pdl_wait(); (3) EVERY kernel calls pdl_wait(), so completion
// is transitive along the chain (a kernel that passed its wait knows all earlier kernels are done).
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

inline bool pdl_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("DDC_PDL");       // tuning aid: DDC_PDL=0 launches every kernel fully serialised
        v = (e != nullptr && atoi(e) == 0) ? 0 : 1;
    }
    return v != 0;
}

// cudaLaunchKernelEx with an optional cluster dimension and the programmatic-serialization attribute
template <typename... KArgs, typename... Args>
inline cudaError_t launch_chain(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster,
                                Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    unsigned n = 0;
    if (cluster > 1) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = static_cast<unsigned>(cluster);
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    if (pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

}  // namespace ddc

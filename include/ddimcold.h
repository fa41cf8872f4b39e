/* ddimcold.h -- C ABI of the B200 (sm_100a) DDIM sampling path of DDIM-COLD.
 *
 * The reference has no FFI: the path sits behind a Python class,
 * DiffusionVisionTransformer (reference ViT.py:158-256, ViT_draft2drawing.py:158-329).
 * This header is the boundary a host binds instead of that class's tensor arithmetic; the
 * Python drop-in in ddim_cold_b200/vit.py binds it with ctypes.
 *
 * Conventions
 *   - every function returns 0 on success, a negative DDC_E_* code otherwise, and never
 *     throws; ddc_last_error() returns a thread-local description of the last failure;
 *   - all tensor pointers are DEVICE pointers unless the name says host;
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream);
 *   - the library performs no device allocation: the caller owns the weight arena and the
 *     work arena (sizes from ddc_arena_bytes) and passes them to ddc_plan_create;
 *   - a plan is bound to one device, one shape and a maximum batch; plans are independent,
 *     so one process can drive eight GPUs with eight plans, or eight processes one each.
 */
#ifndef DDIMCOLD_H_
#define DDIMCOLD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDC_OK 0
#define DDC_E_INVALID (-1)    /* bad argument / unsupported shape */
#define DDC_E_CUDA (-2)       /* a CUDA runtime or driver call failed */
#define DDC_E_DOMAIN (-3)     /* schedule undefined for this k (reference: math domain error) */
#define DDC_E_STATE (-4)      /* weights not packed, arena too small, ... */

/* Constructor arguments that fix every tensor shape (reference ViT.py:160-162). */
typedef struct ddc_shape {
    int32_t img_h, img_w;   /* img_size */
    int32_t patch;          /* patch_size (even) */
    int32_t embed;          /* embed_dim, multiple of 64, <= 512 */
    int32_t depth;
    int32_t heads;          /* num_heads; embed/heads must be 32 or 64 */
    int32_t in_chans;       /* 3 */
    int32_t total_steps;    /* rows of time_embed (2000) */
} ddc_shape;

/* One entry of the per-step table (mirrors the device struct).  x' = a*x + b*x0 is the
 * collapsed form of reference ViT.py:233-234; eps = (x - sqrt_at*x0) * inv_sqrt_1mat is
 * ViT.py:233.  flags bit0 marks the step whose x0 is the sampler's return value. */
typedef struct ddc_step {
    int32_t t;
    int32_t flags;
    float a, b, sqrt_at, inv_sqrt_1mat;
    float pad0, pad1;
} ddc_step;

typedef struct ddc_plan ddc_plan;

const char* ddc_last_error(void);
int ddc_version(void);

/* Schedule of the DDIM loop `for t in range(t_start, 0, -k)` (reference ViT.py:226,231-232;
 * t_start = total_steps-1 for sampler(), the draft's t_start for ViT_draft2drawing.py:399).
 * Scalars are computed in double precision exactly as the reference's math.sqrt expressions
 * and rounded to float once.  Writes at most `capacity` entries to host memory, returns the
 * count in *n_steps.  DDC_E_DOMAIN when some step needs sqrt of a negative number or divides
 * by alpha_tk = 0 -- the cases where the reference raises mid-loop (ValueError: math domain
 * error / ZeroDivisionError). */
int ddc_ddim_table(int32_t k, int32_t total_steps, int32_t t_start, ddc_step* host_table, int32_t capacity,
                   int32_t* n_steps);

/* Cold-diffusion schedule: t = 6..1, x <- clamp(f(x, t))  (ViT_draft2drawing.py:272-288). */
int ddc_cold_table(ddc_step* host_table, int32_t capacity, int32_t* n_steps);

/* Arena sizes in bytes for `shape` at up to `max_batch` images per call. */
int ddc_arena_bytes(const ddc_shape* shape, int32_t max_batch, size_t* weight_bytes, size_t* work_bytes);

/* Create a plan over caller-owned device arenas (256-byte aligned).  The work arena needs no
 * particular initial contents: every row below batch*L of every buffer is written by the kernel
 * that precedes its first reader, rows above it are computed on but never stored (zero-filling it
 * once keeps sanitizers quiet about the padding rows and is what the Python class does). */
int ddc_plan_create(const ddc_shape* shape, int32_t max_batch, int32_t device, void* weight_arena, size_t weight_bytes,
                    void* work_arena, size_t work_bytes, ddc_plan** out);
int ddc_plan_destroy(ddc_plan* plan);

/* Choose the arithmetic format: 0 = bfloat16 tensor-core operands (the C default), 1 = IEEE half
 * tensor-core operands (what the Python class selects by default: same rate, three more mantissa
 * bits, meets the 1e-2 per-step eps bound that bfloat16 misses), 2 = fp32 mode: every contraction
 * in fp32 on the CUDA cores (the reference itself computes in fp32, ViT.py:208-218 has no autocast
 * on this path; per-step eps within 1e-4; a precision mode at a few per cent of the 16-bit
 * throughput).  Formats 0 and 1 accumulate in fp32.  Must be followed by ddc_pack_weights.  In fp32
 * mode ddc_pack_weights copies nothing but the small fp32 tables: the weight matrices are read from
 * the state_dict tensors it was given, which must stay alive and unchanged; and the fp32
 * activations are twice as wide, so the plan serves about half of max_batch (ddc_plan_max_batch). */
int ddc_plan_set_operand_format(ddc_plan* plan, int32_t fmt);

/* Largest batch the plan accepts in its current format (max_batch for the 16-bit formats). */
int ddc_plan_max_batch(const ddc_plan* plan);

/* 1 when this plan's blocks run the qkv Linear and attention as ONE kernel (65-token images with heads of 32 in a
 * 16-bit operand format; reference ViT.py:105-114), 0 when they run as a qkv GEMM followed by an attention kernel.
 * For reports: the per-class timings of ddc_profile_steps then mean {class-row qkv GEMM, fused qkv + attention}. */
int ddc_plan_fused_qkv_attention(const ddc_plan* plan);

/* enable = 0: run the qkv Linear and attention as two kernels for every shape (what shapes other than 65 tokens / heads of
 * 32 do anyway).  The fused kernel (default, enable = 1) is the faster one from about 256 images per call up; below that
 * it pays one more dependent launch per block (0.32 vs 0.28 ms per denoise step at one image).  The two paths round
 * differently, so use ONE setting for all calls whose samples must be bit-identical (e.g. all ranks of a sharded run). */
int ddc_plan_set_fused_qkv_attention(ddc_plan* plan, int32_t enable);

/* Pack an fp32 state dict into the plan's weight arena (bf16 K-major GEMM operands, folded
 * cls/pos/conv-bias rows, fp32 time table).  `tensors` is a HOST array of `n_tensors` device
 * pointers in the reference's state_dict order: cls_token, pos_embed, patch_embed.proj.weight,
 * patch_embed.proj.bias, time_embed.weight, then per block norm1.{weight,bias},
 * attn.qkv.{weight,bias}, attn.proj.{weight,bias}, norm2.{weight,bias}, mlp.fc1.{weight,bias},
 * mlp.fc2.{weight,bias}, then norm.{weight,bias}, head.{weight,bias}
 * (9 + 12*depth entries; reference ViT.py:169-183).  Synchronises `stream` before returning, so the
 * packed weights are visible to every later launch (the step kernels read weights in their
 * prologue, ahead of the programmatic-dependent-launch wait). */
int ddc_pack_weights(ddc_plan* plan, const float* const* tensors, int32_t n_tensors, void* stream);

/* One evaluation of the x0 predictor: out = forward(x, t), unclamped
 * (reference ViT.py:208-218).  x, out: [B,3,H,W] fp32; t: [B] int32 on the device, each in
 * [0, total_steps).  The reference's nn.Embedding raises IndexError for a t outside the table
 * (ViT.py:204); the Python class does the same before calling in.  The library cannot see device
 * values from the host, so a raw caller's out-of-range t is clamped into the table by the kernel. */
int ddc_forward(ddc_plan* plan, const float* x, const int32_t* t, int32_t batch, float* out, void* stream);

/* ddc_forward for callers that cannot allocate device memory (no counterpart in the reference).
 * Same inputs, outputs and result as ddc_forward, but the intermediate activations live in a fixed
 * 32 MiB device array that is part of this library instead of the plan's work arena, so the call
 * performs no cudaMalloc / cudaFree and needs no extra arena from the host.  Meant for hosts whose
 * device memory is owned by another allocator (a serving process that gave its whole budget to a
 * framework pool), for latency-critical threads that must not hit cudaMalloc's implicit device
 * synchronisation, and for code that is not allowed to allocate at that point of its lifetime.
 *   - Limit: the activations of `batch` images must fit in 32 MiB (about 64 images of the S1 shape,
 *     a few thousand of a 16x16 toy shape); larger batches fail with DDC_E_INVALID and a message
 *     naming the largest batch that fits.
 *   - The first call on each device resolves the array's device address (cudaGetSymbolAddress) and
 *     builds host-side descriptors; make that first call before any stream capture begins.
 *   - The array is shared by every plan and every call on that device and keeps the previous call's
 *     contents.  A forward pass overwrites everything it later reads, so nothing needs clearing,
 *     but two calls that overlap on the GPU (different streams, or another process thread) race on
 *     it: serialise them.
 *   - Every GPU that loads the library holds the 32 MiB until the process exits. */
int ddc_forward_fixed_scratch(ddc_plan* plan, const float* x, const int32_t* t, int32_t batch, float* out,
                              void* stream);

/* The whole sampling loop (reference ViT.py:226-236): for each table entry run the predictor
 * at entry.t, clamp to [-1,1], apply the DDIM update in place on `x`.
 *   x        [B,3,H,W] fp32, in: x_T (or the partially noised draft), out: the final state
 *   x0_last  [B,3,H,W] fp32, receives the clamped x0 of the flagged (last) step -- the tensor
 *            the reference's sampler returns as (x0+1)/2; may be NULL
 *   x0_seq   [n_steps,B,3,H,W] fp32 or NULL: every step's clamped x0 (diffusion_sequence,
 *            ViT.py:254, and parity taps)
 *   eps_seq  [n_steps,B,3,H,W] fp32 or NULL: every step's derived eps (ViT.py:233)
 *   use_graph != 0 captures one step as a CUDA graph and replays it n_steps times. */
int ddc_sample(ddc_plan* plan, float* x, int32_t batch, const ddc_step* host_table, int32_t n_steps, float* x0_last,
               float* x0_seq, float* eps_seq, int32_t use_graph, void* stream);

/* Measurement aid (no reference counterpart): run `n_steps` steps of the loop with direct
 * launches and a CUDA event between every kernel; accumulates device milliseconds and launch
 * counts per kernel class into the HOST arrays class_ms[9] / class_launches[9], indexed
 * 0 patchify, 1 patch-embed GEMM, 2 LN+qkv GEMM, 3 attention, 4 proj GEMM, 5 LN+fc1 GEMM,
 * 6 fc2 GEMM, 7 LN+head GEMM, 8 unpatchify+clamp+DDIM.  Synchronises the stream. */
int ddc_profile_steps(ddc_plan* plan, float* x, int32_t batch, const ddc_step* host_table, int32_t n_steps,
                      float* class_ms, int32_t* class_launches, void* stream);

/* Number of kernel launches one step of ddc_sample issues (inside the loop patchify runs once per loop, not per step,
 * whenever the unpatchify + DDIM pass can write the next step's patch matrix itself; ddc_forward issues one more). */
int ddc_launches_per_step(const ddc_plan* plan);

/* ---- kernel-level entry points (unit tests; hosts that schedule kernels themselves) ----
 * `fmt` is the 16-bit operand format: 0 = bf16, 1 = fp16. */

/* Streaming GEMM, out = epilogue(A * W^T + bias).  A: bf16 [m_pad, K] (rows allocated up to
 * a multiple of 128), W: bf16 [N, K], K % 64 == 0, ld_out == N.  epi 0: bf16 out (qkv,
 * reference ViT.py:107); 1: bf16 gelu_erf out (fc1, ViT.py:85-86); 2: fp32 out (head,
 * ViT.py:213).  N % 64 == 0 for epi 0/1, N % 16 == 0 for epi 2.  Only rows < M are written. */
int ddc_k_gemm(int32_t fmt, int32_t epi, int32_t M, int32_t N, int32_t K, const void* A, const void* W, const float* bias, void* out,
               int32_t ld_out, void* stream);

/* Patch-embedding GEMM with the token assembly fused in (reference ViT.py:150,154,199-206) and the first LayerNorm
 * (ViT.py:133):  x[b*L + l, :] = ap[b*L + l, :] * w_patch^T + posb[l, :] + time_tab[t_img[b], :]  (fp32), and
 * xn = 16-bit LayerNorm(x; ln_g, ln_b, eps 1e-5).  ap: 16-bit [m_pad, Kp] from ddc_k_patchify (row b*L is the all-zero
 * cls slot); w_patch: 16-bit [C, Kp]; posb: fp32 [L + 127, C] with row i = pos_embed[i % L] + (i % L == 0 ? cls_token :
 * conv bias), i.e. the position table repeated so the 128 token rows of any tile are contiguous in it; time_tab: fp32
 * [total_steps, C]; t_img: [batch] int32 on the device; x: fp32 [m_pad, C]; xn: 16-bit [batch*L, C]. */
int ddc_k_embed(int32_t fmt, int32_t batch, int32_t L, int32_t C, int32_t Kp, int32_t total_steps, const void* ap,
                const void* w_patch, const float* posb, const float* time_tab, const int32_t* t_img, const float* ln_g,
                const float* ln_b, float* x, void* xn, void* stream);

/* Residual GEMM with the following LayerNorm fused in (proj / fc2, reference ViT.py:136-137
 * then :133/:137/:212):  x[M, C] += A * W^T + bias  (fp32, in place; x allocated to m_pad
 * rows), and xn[M, C] = bf16(LayerNorm(x_new; ln_g, ln_b, eps 1e-5)).  A: bf16 [m_pad, K],
 * W: bf16 [C, K]; C % 64 == 0, C <= 512, K % 64 == 0. */
int ddc_k_gemm_resid_ln(int32_t fmt, int32_t M, int32_t C, int32_t K, const void* A, const void* W, const float* bias, float* x,
                        const float* ln_g, const float* ln_b, void* xn, void* stream);

/* Fused tail of a transformer block (reference ViT.py:136-137 with Mlp.forward :84-90 and the LayerNorm that
 * follows, :133 / :212):  x += O Wproj^T + b_proj;  h = gelu_erf(N(x) Wfc1^T + b_fc1);  x += h Wfc2^T + b_fc2 (fp32, in
 * place);  xn = 16-bit N(x), where N is LayerNorm WITHOUT its affine part (eps 1e-5): the caller folds norm2's gamma / beta
 * into Wfc1 / b_fc1 (W o gamma, b + W beta) and the following LayerNorm's into the Linear that consumes xn, as
 * ddc_pack_weights does for a whole model.  O: 16-bit [m_pad, C]; W*: 16-bit [C, C]; x: fp32 [m_pad, C]; xn: 16-bit [M, C];
 * C in {128, 256, 384}; M > 128. */
int ddc_k_block_tail(int32_t fmt, int32_t M, int32_t C, const void* O, const void* Wproj, const float* b_proj, const void* Wfc1,
                     const float* b_fc1, const void* Wfc2, const float* b_fc2, float* x, void* xn, void* stream);

/* softmax(q k^T / sqrt(d)) v per (image, head) (reference ViT.py:107-114).
 * qkv: bf16 [B*L, 3C] (column s*C + head*d + e); out: bf16 [B*L, C]. */
int ddc_k_attention(int32_t fmt, const void* qkv, void* out, int32_t batch, int32_t L, int32_t C, int32_t heads, void* stream);

/* qkv Linear and attention of a block in one kernel, for 65-token images with heads of 32 (reference ViT.py:105-114 up to
 * the projection): out[b*65 + l, :] = concat_h softmax(q_h k_h^T / sqrt(32)) v_h with q | k | v = xn * w_qkv^T + b_qkv, the
 * q / k / v values never written to global memory (except the class rows).  xn: 16-bit [m_pad, C] (LayerNorm output, rows
 * allocated to a multiple of 128); w_qkv: 16-bit [3C, C]; b_qkv: fp32 [3C]; out: 16-bit [batch*65, C].  Caller-provided
 * scratch: w_hp 16-bit [3C, C] and b_hp fp32 [3C] (the weight regrouped per head pair), qkv_cls 16-bit [batch, 3C].
 * C % 64 == 0, C <= 384, C / heads == 32. */
int ddc_k_qkv_attn(int32_t fmt, const void* xn, const void* w_qkv, const float* b_qkv, void* out, int32_t batch, int32_t C,
                   int32_t heads, void* w_hp, float* b_hp, void* qkv_cls, void* stream);

/* Patchify (im2col of the stride-p conv, reference ViT.py:150,154): x fp32 [B,3,H,W] ->
 * ap bf16 [B*L, Kp] with a zero cls row per image and zero K padding. */
int ddc_k_patchify(int32_t fmt, const float* x, void* ap, int32_t batch, int32_t H, int32_t W, int32_t patch, int32_t Kp,
                   void* stream);

/* Unpatchify + clamp + DDIM update (reference ViT.py:214-217,229,233-234) for one step given
 * by value.  y: fp32 [B*L, ldy] head output; x updated in place; x0_out / eps_out may be NULL. */
int ddc_k_head_ddim(const float* y, int32_t ldy, float* x, int32_t batch, int32_t H, int32_t W, int32_t patch,
                    const ddc_step* host_step, float* x0_out, float* eps_out, void* stream);

/* out[0] = mean_i smooth_l1(pred[i] - target[i]) with beta = 1 -- F.smooth_l1_loss(denoised_img, img) of the trainer's
 * evaluate() (reference multi_gpu_trainer.py:42), so that a validation pass stays on the device.  pred, target: fp32,
 * 16-byte aligned, n > 0 elements.  scratch: 1024 floats of device memory, ZERO before the first call (the last one is a
 * ticket counter that every call leaves at zero again); calls that share a scratch must not overlap.  The sum is formed
 * in a fixed order (block sums in index order, in double), so the value is reproducible run to run. */
int ddc_k_smooth_l1(const float* pred, const float* target, size_t n, float* scratch, float* out, void* stream);

/* out[i] = (in[i] + add) * mul over n floats (16-byte aligned pointers; in == out allowed).  The samplers' return
 * value is (x0 + 1) / 2 (reference ViT.py:237, ViT_draft2drawing.py:288): applied on the device before the copy back. */
int ddc_k_affine(const float* in, float* out, size_t n, float add, float mul, void* stream);

/* out[i] = a * z[i] + b * draft[i % period] over n floats: the draft->drawing start state
 * x = N(0,1) * sqrt(1 - alpha) + draft * sqrt(alpha) (reference ViT_draft2drawing.py:395-396), one draft for the whole
 * batch (period = 3*H*W) or one per image (period = n); each product and the sum rounded separately like the reference's
 * tensor expression.  in-place (out == z) allowed. */
int ddc_k_mix(const float* z, const float* draft, float* out, size_t n, size_t period, float a, float b, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DDIMCOLD_H_ */

/*
 * chirrup_b200 — C ABI of the B200-native batched RWKV-7 forward step.
 *
 * Plain pointers and sizes only (no torch types).  All pointers are DEVICE pointers unless a parameter says "host".
 * Every entry point takes an explicit CUDA stream (a cudaStream_t passed as void*) and returns 0 on success, a
 * positive cudaError_t, or a negative B200_ERR_* code.  Nothing here ever falls back to a CPU path.
 *
 * Reference interfaces replaced (paths under /root/reference):
 *   b200_wkv7_forward_seq   <- cuda_forward_seq   Albatross/cuda/rwkv7_state_fwd_fp16.cu:313-316
 *                              (torch op rwkv7_state_fwd_fp16::forward_seq, Albatross/cuda/rwkv7_state_fwd_fp16.cpp:10-12,21-25)
 *   b200_wkv7_forward_one   <- cuda_forward_one   Albatross/cuda/rwkv7_state_fwd_fp16.cu:318-322 (.cpp:14-16)
 *   b200_spmv_forward       <- cuda_spmv_forward  Albatross/cuda/rwkv7_state_fwd_fp16.cu:324-330 (.cpp:17-19)
 *                              and rwkv_mm_sparsity, Albatross/rwkv_mm_op_triton.py:39-61
 *   b200_linear_f16         <- the F.linear / `@` call sites of Albatross/rwkv7.py:625-630,637,649,678-679,551 (cuBLAS)
 *   b200_model_*            <- RWKV_x070.forward_seq_batch_seperate and friends, Albatross/rwkv7.py:503-563
 */
#ifndef CHIRRUP_B200_H
#define CHIRRUP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ERR_BAD_ARG (-1)
#define B200_ERR_NO_DRIVER (-2)   /* cuTensorMapEncodeTiled could not be resolved (no CUDA driver) */
#define B200_ERR_WORKSPACE (-3)   /* workspace too small for this call */
#define B200_ERR_UNSUPPORTED (-4)
#define B200_ERR_TOO_LARGE (-5)   /* a size beyond the stated maximum of a fixed-scratch entry point (b200_linear_f16_static) */

/* ABI version of this header (bumped on any signature change). */
int b200_abi_version(void);

/*
 * WKV-7 recurrence over T tokens for B sequences x H heads (head size 64), fp16, state updated IN PLACE.
 *   state              fp16, tile of slot s / head h at  state + s*state_slot_stride + h*64*64, laid out [i = value row][j = key col]
 *   state_slot_stride  elements between consecutive slots (C*64 for the reference's contiguous (B,H,64,64) layout)
 *   n_slots            number of addressable slots (ignored, = B, when slot_idx is NULL)
 *   r,w,k,v,a,b,y      fp16 (B,T,C) contiguous; w is the PRE-activation decay (the kernel applies exp(-e^-0.5 sigmoid(w)) - 1 + dither)
 *   elapsed_t          int32 (B): tokens consumed before this call; token t uses elapsed_t[b] + t for the dither
 *   slot_idx           int32 (B) or NULL: sequence b uses state slot slot_idx[b] (NULL: slot b)
 * Bit-identical to the reference kernel on the same inputs.
 */
int b200_wkv7_forward_seq(int B, int T, int C, int H, void* state, long state_slot_stride, int n_slots, const void* r,
                          const void* w, const void* k, const void* v, const void* a, const void* b, void* y,
                          const int* elapsed_t, const int* slot_idx, void* stream);

/* T = 1 form (inputs (B,C)). */
int b200_wkv7_forward_one(int B, int C, int H, void* state, long state_slot_stride, int n_slots, const void* r,
                          const void* w, const void* k, const void* v, const void* a, const void* b, void* y,
                          const int* elapsed_t, const int* slot_idx, void* stream);

/*
 * Fused TMix core: everything between the projection GEMMs and the output GEMM of RWKV_x070_TMix_seq_batch
 * (Albatross/rwkv7.py:632-649) in ONE kernel per (sequence, head) tile:
 *   kk = normalize(k*k_k) per head; k' = k*(1+(a-1)*k_a); v' = v + (v_first - v)*vgate (skipped when v_first == NULL: layer 0);
 *   y = WKV7(state, r, w, k', v', -kk, kk*a)  [as b200_wkv7_forward_seq];  out = (GroupNorm_head(y; eps 64e-5)*ln_w + ln_b
 *   + (sum_head r*k'*r_k) v') * g.
 * r,w,k,v,a,g,vgate,v_first,out: fp16 (B,T,C); a and vgate already sigmoid-ed; k_k,k_a,r_k,ln_w,ln_b: fp16 (C).
 * State arguments as in b200_wkv7_forward_seq.  This is the HBM-roofline kernel of the decode step.
 */
int b200_tmix_core_fused(int B, int T, int C, int H, void* state, long state_slot_stride, int n_slots, const void* r,
                         const void* w, const void* k, const void* v, const void* a, const void* g, const void* vgate,
                         const void* v_first, const void* k_k, const void* k_a, const void* r_k, const void* ln_w,
                         const void* ln_b, void* out, const int* elapsed_t, const int* slot_idx, void* stream);

/*
 * The same fused TMix core for PREFILL (T >= 16 pays off): the WKV-7 recurrence evaluated 16 tokens at a time in its chunked
 * (WY / UT-transform) form on the tensor cores — fp16 operands rescaled around the middle of each chunk by the per-channel
 * cumulative decays (the elapsed_t dither folded into every token's decay), fp32 accumulation, state kept in fp32 for the
 * whole launch and rounded to fp16 once (csrc/wkv7_chunk.cu).  The reference has only the sequential T loop
 * (Albatross/cuda/rwkv7_state_fwd_fp16.cu:56-82); agreement with it is to its own fp16 noise, not bit for bit — the
 * sequential kernels (b200_tmix_core_fused) remain the bit-exact mode.  Same arguments as b200_tmix_core_fused.
 */
int b200_tmix_core_chunked(int B, int T, int C, int H, void* state, long state_slot_stride, int n_slots, const void* r,
                           const void* w, const void* k, const void* v, const void* a, const void* g, const void* vgate,
                           const void* v_first, const void* k_k, const void* k_a, const void* r_k, const void* ln_w,
                           const void* ln_b, void* out, const int* elapsed_t, const int* slot_idx, void* stream);

/*
 * out[c] = sum_d vec[d] * mat[d*C + c], skipping rows with vec[d] == 0 (ReLU^2 sparsity of the FFN key).
 * vec (D) fp16, mat (D,C) fp16 row-major, out (C) fp16 — overwritten (no pre-zeroing needed, unlike the reference).
 * fp32 accumulation in a fixed order: deterministic, one rounding.
 */
int b200_spmv_forward(int D, int C, const void* vec, const void* mat, void* out, void* stream);

/* Epilogues of b200_linear_f16. */
enum {
    B200_EPI_NONE = 0,     /* out = f16(acc + bias)                                   */
    B200_EPI_TANH = 1,     /* out = f16(tanh(f16(acc + bias)))                        */
    B200_EPI_SIGMOID = 2,  /* out = f16(sigmoid(f16(acc + bias)))                     */
    B200_EPI_RELU_SQ = 3,  /* out = f16(relu(f16(acc))^2)                             */
    B200_EPI_RESADD = 4    /* out = f16(residual + f16(acc)) ; residual may alias out */
};

/*
 * out (M,N) = epilogue( A (M,K) . W (N,K)^T ), fp16 in, fp32 accumulate on tcgen05 tensor cores, fp16 out.
 *   lda, ldw, ldo   row pitches in elements (multiples of 8); K % 8 == 0
 *   bias            fp16 (N) or NULL ; residual fp16 (M, ldo) or NULL (B200_EPI_RESADD only)
 *   workspace       device scratch for split-K partials + tile counters, b200_linear_workspace_bytes() bytes, zero-initialised
 *                   once by the caller (the kernel restores the counters to zero)
 */
int b200_linear_f16(int M, int N, int K, const void* A, long lda, const void* W, long ldw, void* out, long ldo,
                    const void* bias, const void* residual, int epilogue, void* workspace, long workspace_bytes,
                    void* stream);
long b200_linear_workspace_bytes(int max_M, int max_N);

/*
 * b200_linear_f16 for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO counterpart in the reference.
 * Same arguments and results as b200_linear_f16 minus the workspace: the split-K scratch is a file-scope __device__ array of
 * B200_LINEAR_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it until the process exits).
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 *   limits   M <= B200_LINEAR_STATIC_MAX_M, N <= B200_LINEAR_STATIC_MAX_N, K <= B200_LINEAR_STATIC_MAX_K, else
 *            B200_ERR_TOO_LARGE; a plan whose partials would not fit the scratch is not chosen, or the call returns
 *            B200_ERR_WORKSPACE.
 *   sharing  every call on a device uses the SAME array (it keeps the last call's contents; nothing in it has to start at zero:
 *            each launch writes the partials it reads).  Calls that can overlap on the GPU — different streams, concurrent graph
 *            replays — race on it: issue them on one stream.
 */
#define B200_LINEAR_STATIC_SCRATCH_BYTES (32 * 1024 * 1024)
#define B200_LINEAR_STATIC_MAX_M 4096
#define B200_LINEAR_STATIC_MAX_N 65536
#define B200_LINEAR_STATIC_MAX_K 16384
int b200_linear_f16_static(int M, int N, int K, const void* A, long lda, const void* W, long ldw, void* out, long ldo,
                           const void* bias, const void* residual, int epilogue, void* stream);
long b200_linear_static_scratch_bytes(void);

/*
 * Up to 8 independent linear layers that share M in ONE launch (e.g. the r, k, v projections and the four LoRA
 * down-projections of a TMix layer, Albatross/rwkv7.py:625-630, :637 — they all read the six token-shift mixes).
 * `problems` is a HOST array of n descriptors; fields as in b200_linear_f16.  For 128 < M each output tile is computed by a
 * CTA pair (tcgen05.mma.cta_group::2); other M, or more than 8 problems, run as one launch per problem.
 */
typedef struct b200_linear_desc {
    int N, K;
    const void* A;
    long lda;
    const void* W;
    long ldw;
    void* out;
    long ldo;
    const void* bias;
    const void* residual;
    int epilogue;
} b200_linear_desc;
int b200_linear_group_f16(int n, const b200_linear_desc* problems, int M, void* workspace, long workspace_bytes, void* stream);

/* out[i,:] = LayerNorm(x[i,:]) * w + b over `rows` contiguous rows of C fp16 (eps 1e-5, fp32 statistics); used at load time
 * to fold ln0 into the embedding table (Albatross/rwkv7.py:206).  out may alias x. */
int b200_layer_norm_f16(int rows, int C, const void* x, const void* w, const void* b, void* out, void* stream);


/* ------------------------------------------------------------------------------------------------------------------
 * The whole batched step: RWKV_x070.forward_seq_batch_seperate (Albatross/rwkv7.py:555-563 -> :503-553), i.e.
 * embedding gather, L x [LN1, TMix (:618-649), residual, LN2, CMix (:673-679), residual], last-token select, ln_out, head,
 * state[2] += T — launched as one sequence of sm_100a kernels on `stream`.
 *
 * Weights stay owned by the caller (device memory); the model object only records pointers.  Layouts are those of the
 * reference's transformed `z` dict (Albatross/rwkv7.py:211-221) except ffn.value.weight, kept in checkpoint layout (C, 4C)
 * = (N, K) so that every dense weight is K-major.
 *   global_w[4]                : emb.weight (V,C) [ln0 already folded in, rwkv7.py:206], ln_out.weight, ln_out.bias, head.weight (V,C)
 *   layer_w[n_layer * B200_LW_COUNT] : per layer, indexed by the B200_LW_* enum below
 */
enum {
    B200_LW_LN1_W = 0, B200_LW_LN1_B, B200_LW_LN2_W, B200_LW_LN2_B,
    B200_LW_X_R, B200_LW_X_W, B200_LW_X_K, B200_LW_X_V, B200_LW_X_A, B200_LW_X_G,
    B200_LW_W0, B200_LW_W1 /* (D_w,C) */, B200_LW_W2 /* (C,D_w) */,
    B200_LW_A0, B200_LW_A1, B200_LW_A2,
    B200_LW_V0, B200_LW_V1, B200_LW_V2,      /* ignored on layer 0 (rwkv7.py:636) */
    B200_LW_G1, B200_LW_G2,
    B200_LW_K_K, B200_LW_K_A, B200_LW_R_K,
    B200_LW_RECEPTANCE, B200_LW_KEY, B200_LW_VALUE, B200_LW_OUTPUT, /* (C,C) */
    B200_LW_LNX_W, B200_LW_LNX_B,
    B200_LW_FFN_X_K, B200_LW_FFN_KEY /* (4C,C) */, B200_LW_FFN_VALUE /* (C,4C) */,
    B200_LW_COUNT
};

typedef struct b200_model b200_model;

int b200_model_create(int n_layer, int n_embd, int vocab_size, int d_w, int d_a, int d_v, int d_g,
                      const void* const* global_w, const void* const* layer_w, b200_model** out);
void b200_model_destroy(b200_model* m);
/* Scratch needed by b200_model_forward for up to max_tokens = B*T tokens per call (includes the linear workspace). */
long b200_model_workspace_bytes(const b200_model* m, int max_tokens);

/*
 * One forward over B sequences x T tokens (decode: T = 1).  State tensors are the caller's, updated in place, and may be
 * B-axis views of larger allocations (chirrup/worker.py:697-701), hence the explicit strides (in elements):
 *   tokens      int32 (B*T) device, row-major (B,T)
 *   state0      fp16 shift states; row of (layer l, which w in {0: TMix, 1: CMix}, slot s) at
 *               state0 + l*s0_layer_stride + w*s0_which_stride + s*C          [reference shape (L,2,B,C), rwkv7.py:228]
 *   state1      fp16 WKV states; tile (l, s, h) at state1 + l*s1_layer_stride + s*C*64 + h*4096   [(L,B,H,64,64), rwkv7.py:229]
 *   elapsed     int32 (B) = state[2]; read by every layer, then += T (rwkv7.py:552)
 *   slot_idx    int32 (B) or NULL; sequence b lives in slot slot_idx[b] of state0/state1 (elapsed stays indexed by b)
 *   logits      fp16 (B, V), or (B*T, V) when full_output != 0; NULL skips ln_out + head (prefill that discards logits,
 *               chirrup/worker.py:776)
 *   workspace   b200_model_workspace_bytes(m, >= B*T) bytes, zero-initialised once by the caller
 */
int b200_model_forward(b200_model* m, int B, int T, const int* tokens, void* state0, long s0_layer_stride,
                       long s0_which_stride, void* state1, long s1_layer_stride, int n_slots, int* elapsed,
                       const int* slot_idx, void* logits, int full_output, void* workspace, long workspace_bytes,
                       void* stream);
/* Optional: hand the model ffn.value.weight of every layer in the reference's transformed layout (F, C) row-major
 * (Albatross/rwkv7.py:217-218 transposes it at load) — n_layer device pointers, or NULL to clear.  With it, a B = T = 1 step
 * (forward / forward_one, RWKV_x070_CMix_one, rwkv7.py:653-662) runs its FFN value product as the ReLU^2-sparse
 * vector x matrix product (b200_spmv_forward semantics: only rows with a non-zero key activation are read). */
int b200_model_set_ffn_value_rows(b200_model* m, const void* const* ffn_value_fc);

/* Greedy sampling of `rows` logit rows (fp16, V columns, row pitch ld): out[i] = lowest index of the row maximum (int32).
 * The top_k == 1 case of the worker's sampler (chirrup/worker.py:733-740, chirrup/utils/samplers.py:171-255). */
int b200_argmax_f16(int rows, int V, const void* logits, long ld, int* out, void* stream);

/*
 * One decode SAMPLING step for `rows` logit rows in one kernel (one CTA per row, no sort): what chirrup/worker.py:713-740 does
 * to the decode logits (forbidden tokens -= 1e10; occurrence *= penalty_decay; logits -= alpha_presence + occurrence *
 * frequency_penalty — in place, fp32 math, one fp16 rounding per op), then chirrup/utils/samplers.py:98-168
 * (sample_logits_real_batch: softmax(logits / temperature), keep sorted rank < top_k (top_k <= 0: no limit) and cumulative
 * probability <= top_p (top_p >= 1: no limit), always at least one token, draw proportionally), then
 * chirrup/worker.py:527-547 for the drawn token (occurrence += 1 unless the token is in no_penalty; alpha_presence =
 * presence_penalty).  Rows with top_k == 1, top_p <= 0 or temperature <= 0 take the arg-max (lowest index).
 * All pointers are device pointers; the struct itself is read on the host.  Per-slot arrays are indexed by slot[i] (NULL: i).
 * The draw is a pure function of (seed, step, row index): Philox4x32-10.  V <= 65536, V % 8 == 0.
 */
typedef struct b200_sample_params {
    const int* slot;                 /* int32 (rows) or NULL */
    float* occurrence;               /* fp32 (n_slots, V) or NULL: no penalties */
    float* presence;                 /* fp32 (n_slots, V) alpha_presence_vector; NULL iff occurrence is NULL */
    const float* penalty_decay;      /* fp32 (n_slots) */
    const float* frequency_penalty;  /* fp32 (n_slots) */
    const float* presence_penalty;   /* fp32 (n_slots) */
    const float* temperature;        /* fp32 (n_slots) or NULL (1.0) */
    const float* top_p;              /* fp32 (n_slots) or NULL (1.0) */
    const int* top_k;                /* int32 (n_slots) or NULL (0) */
    const int* forbidden;            /* int32 (rows, max_forbidden), padded with -1, or NULL */
    int max_forbidden;               /* <= 512 */
    const int* no_penalty;           /* int32 (n_no_penalty) token ids whose occurrence is not incremented, or NULL */
    int n_no_penalty;
    unsigned long long seed, step;
    int* kept;                       /* optional int32 (rows): size of the kept set (tests / debugging) */
    int mode;                        /* 0: samplers.py:98-168 as described above.  1: the sampler chirrup/worker.py:12 imports,
                                      * sample_logits_rwkv_pip_compatible (samplers.py:171-255): softmax WITHOUT temperature;
                                      * top-p keeps every token whose probability is >= the probability at the first sorted
                                      * position where the cumulative sum reaches top_p (ties at the cutoff all stay); top-k on
                                      * the sort rank; then probs ** fp16(1 / temperature); draw proportionally */
} b200_sample_params;
int b200_sample_f16(int rows, int V, void* logits, long ld, const b200_sample_params* params, int* out, void* stream);

/* One request's state rows in or out of a slot table (admission, cache extraction, migration between workers): `rows`
 * pieces of `width_bytes` at pitches dst_pitch / src_pitch bytes, as ONE strided DMA copy on `stream` — device to device on
 * one GPU, between peer GPUs (NVLink when peer access is enabled), or to / from pinned host memory.  No SM work, so it runs
 * under a decode step.  Replaces the host round trips and the three-copy slot swap of chirrup/worker.py:427-429, :468-470
 * (slot -> CPU), :586-597 (CPU -> slot), :304-360 (swap).  state[1] of slot s: dst = base + s*H*64*64*2, pitch =
 * n_slots*H*64*64*2, width = H*64*64*2, rows = L;  state[0]: pitch = n_slots*C*2, width = C*2, rows = 2L. */
int b200_slot_copy(void* dst, long dst_pitch, const void* src, long src_pitch, long width_bytes, long rows, void* stream);
/* A whole request in one call: b200_slot_copy of its state[1] rows (…1), of its state[0] rows (…0) and the 4-byte copy of its
 * elapsed counter (skipped when either pointer is NULL).  chirrup/worker.py:591-597 (admission) / :427-429 (extraction) for
 * all three state tensors at once; between two workers' GPUs this is the NVLink migration of BASELINE configs[4]. */
int b200_slot_move(void* dst1, long dst1_pitch, const void* src1, long src1_pitch, long width1, long rows1, void* dst0,
                   long dst0_pitch, const void* src0, long src0_pitch, long width0, long rows0, int* dst_elapsed,
                   const int* src_elapsed, void* stream);

/* Number of kernels the last b200_model_forward on this model launched (bench.py reports it as gpu_launches). */
int b200_model_last_launch_count(const b200_model* m);

#ifdef __cplusplus
}
#endif
#endif

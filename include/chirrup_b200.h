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

#ifdef __cplusplus
}
#endif
#endif

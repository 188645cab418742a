// Internal launch interface of the glue kernels and of the linear / WKV launchers used by the model step (model.cu).
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

namespace b200 {

template <int NMIX>
struct MixPtrs {
    const __half* mix[NMIX];  // lerp coefficients (C)
    __half* out[NMIX];        // outputs (M, C)
};

int launch_embed_gather(int M, int C, int V, const int* tokens, const __half* emb, __half* x, cudaStream_t s, bool pdl = false);
// part != nullptr (T == 1 only): x is first completed in place from the producer's n_part deferred K-split partials
// (fp32, split stride B*T*C floats; see linear_f16_launch) — x <- fp16(x + fp16(sum of partials in split order)).
int launch_ln_shift_mix6(int B, int T, int C, __half* x, const __half* ln_w, const __half* ln_b, __half* shift,
                         long shift_stride, const int* slot_idx, const MixPtrs<6>& mp, cudaStream_t s, bool pdl = false,
                         const float* part = nullptr, int n_part = 0);
int launch_ln_shift_mix1(int B, int T, int C, __half* x, const __half* ln_w, const __half* ln_b, __half* shift,
                         long shift_stride, const int* slot_idx, const MixPtrs<1>& mp, cudaStream_t s, bool pdl = false,
                         const float* part = nullptr, int n_part = 0);
int launch_ln_rows(int rows, int C, const __half* x, long row_step, long row_off, const __half* ln_w, const __half* ln_b,
                   __half* out, cudaStream_t s, bool pdl = false);
int launch_add_elapsed(int B, int T, int* elapsed, cudaStream_t s, bool pdl = false);
int launch_argmax_rows(int rows, int V, const __half* logits, long ld, int* out, cudaStream_t s);

struct TmixPrepArgs {
    int M, C;
    const __half *k, *v, *a, *vgate, *v_first;  // vgate/v_first == nullptr on layer 0
    const __half *k_k, *k_a;
    __half *k_out, *v_out, *a_out, *b_out;
};
int launch_tmix_prep(const TmixPrepArgs& a, cudaStream_t s, bool pdl = false);

struct TmixPostArgs {
    int M, C;
    const __half *y, *r, *k, *v, *g;
    const __half *ln_w, *ln_b, *r_k;
    __half* out;
};
int launch_tmix_post(const TmixPostArgs& a, cudaStream_t s, bool pdl = false);

// wkv7.cu: fused TMix core (prep + WKV-7 + GroupNorm/bonus/gate), inputs (B*T, C) fp16
struct WkvFusedArgs {
    int B, T, C, H, rows_per_slot;
    const __half *r, *w, *k, *v, *a, *g, *vgate, *v_first;  // vgate / v_first == nullptr on layer 0
    const __half *k_k, *k_a, *r_k, *ln_w, *ln_b;
    __half* out;
    const int* elapsed_t;
    const int* slot_idx;
};
int launch_wkv7_fused(const WkvFusedArgs& a, void* state, long state_slot_stride, int n_slots, cudaStream_t stream,
                      bool pdl = false);
// wkv7_chunk.cu: the same fused TMix core with the recurrence in its chunked tensor-core form (prefill; tolerance-level parity)
int launch_wkv7_chunk(const WkvFusedArgs& a, void* state, long state_slot_stride, int n_slots, cudaStream_t stream,
                      bool pdl = false);
int launch_wkv7_seq(int B, int T, int C, int H, void* state, long state_slot_stride, int n_slots, const void* r, const void* w,
                    const void* k, const void* v, const void* a, const void* b, void* y, const int* elapsed_t,
                    const int* slot_idx, cudaStream_t stream, bool pdl = false);

// spmv.cu: out[c] = (res ? res[c] + : ) sum_{d : vec[d] != 0} vec[d] * mat[d*C + c]   (bsz-1 FFN value product, rwkv7.py:659)
int launch_spmv(int D, int C, const __half* vec, const __half* mat, const __half* res, __half* out, cudaStream_t stream,
                bool pdl = false);

// gemm.cu
int linear_f16_launch(int M, int N, int K, const void* A, long lda, const void* W, long ldw, void* out, long ldo,
                      const void* bias, const void* residual, int epilogue, void* workspace, long workspace_bytes,
                      cudaStream_t stream, bool pdl = false, int* defer_splits = nullptr);
// defer_splits != nullptr (B200_EPI_RESADD only): the launch MAY split K over S = *defer_splits > 0 independent CTA pairs and
// leave the reduction to the consumer — `out` is then NOT written; split z's fp32 partial of out[row][col] is at
// ((float*)workspace)[(z * M + row) * N + col] and the consumer computes fp16(residual + round_h(p_0 + p_1 + ... in that order)),
// exactly what the in-kernel reduction produces.  *defer_splits == 0: the launch completed `out` itself.

struct LinearDesc {
    int N, K;
    const void* A;
    long lda;
    const void* W;
    long ldw;
    void* out;
    long ldo;
    const void* bias;
    const void* residual;
    int epi;
};
// Up to four linear layers sharing M in one launch (gemm.cu); larger groups / M fall back to one launch each.
int linear_group_launch(int n, const LinearDesc* d, int M, void* workspace, long workspace_bytes, cudaStream_t stream,
                        bool pdl = false);

// gemm_pair.cu: the same, on cta_group::2 CTA pairs (128 < M; up to 8 problems).  Returns B200_ERR_UNSUPPORTED when the group
// does not fit that kernel.
int pair_group_launch(int n, const LinearDesc* d, int M, void* workspace, long workspace_bytes, cudaStream_t stream, bool pdl,
                      unsigned long long* dbg = nullptr, int* defer_splits = nullptr);

// gemm_pair.cu: FFN key + FFN value of a decode step (128 < M <= 256) in one persistent launch; the value product is left as
// *n_split fp32 partials (layout of linear_f16_launch's deferred reduction) for the next ln_shift_mix.  `flags`: 16 ints, zero
// before the first call (the kernel leaves them zero).  B200_ERR_UNSUPPORTED when the shapes do not fit: use two launches.
int ffn_chain_launch(int M, int C, int F, const void* kx, const void* Wk, void* hidden, const void* Wv, float* partials,
                     long partials_bytes, int* flags, int* n_split, cudaStream_t stream, bool pdl);

}  // namespace b200

/*
 * TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.  The product path (chirrup_b200/) never does.
 *
 * CPU restatement of the reference's WKV-7 fp16 state-forward kernels and of its sparse vec x mat kernel:
 *
 *   wkv7_oracle_forward_seq  <-  kernel_forward_w0_fp16_dither_seq   Albatross/cuda/rwkv7_state_fwd_fp16.cu:26-97
 *                                (and ..._dither_one, :99-167, which is the T == 1 unrolling of the same math)
 *   wkv7_oracle_spmv         <-  spvecmatmul_noindices               Albatross/cuda/rwkv7_state_fwd_fp16.cu:222-310
 *
 * The reference has no CPU implementation of these ops, so this restates the arithmetic of the CUDA kernel
 * *as compiled for sm_100a* (SASS read from oracle/_ref/libref_wkv7.so, see DESIGN.md "fp16 op order"):
 *
 *   decay  w~_j = f16( fmaf( float(int32(0x9E3779B9 * t)), 2^-41,  ex2(rcp(1 + ex2(-log2e * w_j)) * -0.875038..) - 1 ) )
 *                                                                                           (.cu:20-23, :59)
 *   sa_i   = hadd( lo, hi )  of a half2 chain   acc = fma(a_j, S_ij, acc),  j = 0..63 (even j in lo, odd j in hi) (.cu:65-70)
 *   S_ij  <- hadd( S_ij, fma( sa_i, b_j, fma( S_ij, w~_j, hmul(k_j, v_i) ) ) )                                 (.cu:78)
 *   y_i    = hadd( lo, hi )  of a half2 chain   acc = fma(S_ij(new), r_j, acc)                                  (.cu:79-81)
 *
 * Every h* op is one IEEE binary16 round-to-nearest-even operation (denormals kept), emulated here by an exact
 * double computation followed by a single rounding.  The only non-bit-faithful pieces are the GPU's MUFU.EX2 /
 * MUFU.RCP approximations inside the decay (<= 2 ulp fp32 before the f16 rounding): w~ can differ from the GPU
 * value by one f16 ulp on a small fraction of elements.  Parity status: pinned against the reference's real CUDA
 * kernel on the GPU box (tests/test_wkv_gpu.py, through oracle/_ref), not by fixtures in the reference (it has none).
 *
 * Layouts follow the reference: state (B, H, 64, 64) f16, [b][h][i = value row][j = key col] (.cu:37, :52);
 * r,w,k,v,a,b,y (B, T, C) f16 contiguous, index b*T*C + t*C + h*64 + i (.cu:57); elapsed_t int32 (B,).
 */
#include <stdint.h>
#include <string.h>
#include <math.h>
#include <stdlib.h>

#define HS 64

typedef uint16_t f16;

static inline float h2f(f16 h) { _Float16 x; memcpy(&x, &h, 2); return (float)x; }
static inline f16 f2h(float f) { _Float16 x = (_Float16)f; f16 h; memcpy(&h, &x, 2); return h; }
static inline f16 d2h(double d) { _Float16 x = (_Float16)d; f16 h; memcpy(&h, &x, 2); return h; }

/* one-rounding binary16 ops (products of two binary16 values are exact in double) */
static inline f16 hmul(f16 a, f16 b) { return d2h((double)h2f(a) * (double)h2f(b)); }
static inline f16 hadd(f16 a, f16 b) { return d2h((double)h2f(a) + (double)h2f(b)); }
static inline f16 hfma(f16 a, f16 b, f16 c) { return d2h((double)h2f(a) * (double)h2f(b) + (double)h2f(c)); }

/* reference .cu:20-23 + :59 (fast-math form as compiled: FMUL, EX2, FADD, RCP, FMUL, EX2, FADD, FFMA, F2F) */
static inline f16 decay_dither(f16 w_pre, int32_t t)
{
    const float nlog2_e = -1.4426950408889634f;
    const float nexp_half_log2_e = -0.8750387749145276f;
    const float two_to_neg_41 = 4.547473508864641e-13f;
    float x = h2f(w_pre) * nlog2_e;
    float d = exp2f(x) + 1.0f;
    float q = (1.0f / d) * nexp_half_log2_e;
    float e = exp2f(q) - 1.0f;
    int32_t rot = (int32_t)(2654435769u * (uint32_t)t);   /* int32 wrap-around multiply */
    return f2h(fmaf((float)rot, two_to_neg_41, e));
}

void wkv7_oracle_decay(int n, const f16 *w_pre, int32_t t, f16 *out)
{
    for (int i = 0; i < n; i++) out[i] = decay_dither(w_pre[i], t);
}

/* One (b,h) tile for T tokens; `S` is the 64x64 tile, in place. */
static void tile_forward(int T, int C, f16 *S, const f16 *r, const f16 *w, const f16 *k, const f16 *v,
                         const f16 *a, const f16 *b, f16 *y, int32_t t0)
{
    f16 wd[HS];
    for (int t = 0; t < T; t++) {
        const f16 *rt = r + (size_t)t * C, *wt = w + (size_t)t * C, *kt = k + (size_t)t * C;
        const f16 *vt = v + (size_t)t * C, *at = a + (size_t)t * C, *bt = b + (size_t)t * C;
        for (int j = 0; j < HS; j++) wd[j] = decay_dither(wt[j], t0 + t);
        for (int i = 0; i < HS; i++) {
            f16 *row = S + i * HS;
            f16 lo = 0, hi = 0;
            for (int j = 0; j < HS; j += 2) {
                lo = hfma(at[j], row[j], lo);
                hi = hfma(at[j + 1], row[j + 1], hi);
            }
            f16 sa = hadd(lo, hi);
            f16 vi = vt[i];
            lo = 0; hi = 0;
            for (int j = 0; j < HS; j += 2) {
                f16 s0 = row[j], s1 = row[j + 1];
                s0 = hadd(s0, hfma(sa, bt[j], hfma(s0, wd[j], hmul(kt[j], vi))));
                s1 = hadd(s1, hfma(sa, bt[j + 1], hfma(s1, wd[j + 1], hmul(kt[j + 1], vi))));
                row[j] = s0; row[j + 1] = s1;
                lo = hfma(s0, rt[j], lo);
                hi = hfma(s1, rt[j + 1], hi);
            }
            (y + (size_t)t * C)[i] = hadd(lo, hi);
        }
    }
}

/* Same argument meaning as the reference's cuda_forward_seq (.cu:313-316).  state_batch_stride is in
 * elements (C*64 when the state is contiguous, as the reference assumes). */
void wkv7_oracle_forward_seq(int B, int T, int C, int H, f16 *state, long state_batch_stride,
                             const f16 *r, const f16 *w, const f16 *k, const f16 *v, const f16 *a, const f16 *b,
                             f16 *y, const int32_t *elapsed_t)
{
    #pragma omp parallel for collapse(2) schedule(static)
    for (int bb = 0; bb < B; bb++) {
        for (int h = 0; h < H; h++) {
            size_t off = (size_t)bb * T * C + (size_t)h * HS;
            tile_forward(T, C, state + (size_t)bb * state_batch_stride + (size_t)h * HS * HS,
                         r + off, w + off, k + off, v + off, a + off, b + off, y + off, elapsed_t[bb]);
        }
    }
}

/* out[c] = sum_{d : vec[d] != +-0} vec[d] * mat[d, c].  The reference accumulates with f16 hfma2 inside a
 * 64-row slab and merges slabs with f16 atomicAdd in arbitrary order (.cu:296-309): its result is not
 * run-to-run reproducible, so the oracle states the *exact* product in double, rounded once; callers compare
 * with a tolerance that covers the reference's f16 accumulation. */
void wkv7_oracle_spmv(int D, int C, const f16 *vec, const f16 *mat, f16 *out)
{
    double *acc = (double *)calloc((size_t)C, sizeof(double));
    for (int d = 0; d < D; d++) {
        if ((vec[d] & 0x7fff) == 0) continue;
        double x = (double)h2f(vec[d]);
        const f16 *row = mat + (size_t)d * C;
        for (int c = 0; c < C; c++) acc[c] += x * (double)h2f(row[c]);
    }
    for (int c = 0; c < C; c++) out[c] = d2h(acc[c]);
    free(acc);
}

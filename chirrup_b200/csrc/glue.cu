// Elementwise / small-reduction kernels of the RWKV-7 step ("the glue"): each one replaces a chain of ATen launches of the
// reference model (/root/reference/Albatross/rwkv7.py) and keeps its per-op fp16 rounding (ATen computes each half op in fp32
// and rounds the result to fp16 — so does every step below; `hr()` is that rounding).
//
//   embed_gather      emb[idx]                                                        rwkv7.py:507
//   ln_shift_mix<N>   LayerNorm -> token shift -> N lerps (+ new shift state)         rwkv7.py:523, :621-623 / :533, :675-677
//   tmix_prep         kk normalise, k/v mixing, a = -kk, b = kk*a                     rwkv7.py:632-637
//   tmix_post         GroupNorm(eps 64e-5) + bonus (r.k.r_k) v + gate                 rwkv7.py:647-649
//   ln_rows           ln_out on the selected rows                                     rwkv7.py:546-548
//   add_elapsed       state[2] += T                                                   rwkv7.py:552
#include "common.cuh"
#include "glue.h"

namespace b200 {
namespace {

DEVI float hr(float x) { return __half2float(__float2half_rn(x)); }

struct H8 {
    float f[8];
};
DEVI H8 ld8(const __half* p) {
    int4 q = *reinterpret_cast<const int4*>(p);
    const __half* h = reinterpret_cast<const __half*>(&q);
    H8 o;
#pragma unroll
    for (int e = 0; e < 8; e++) o.f[e] = __half2float(h[e]);
    return o;
}
DEVI H8 cvt8(const int4& q) {
    const __half* h = reinterpret_cast<const __half*>(&q);
    H8 o;
#pragma unroll
    for (int e = 0; e < 8; e++) o.f[e] = __half2float(h[e]);
    return o;
}
DEVI void st8(__half* p, const H8& v) {
    int4 q;
    __half* h = reinterpret_cast<__half*>(&q);
#pragma unroll
    for (int e = 0; e < 8; e++) h[e] = __float2half_rn(v.f[e]);
    *reinterpret_cast<int4*>(p) = q;
}

// Block-wide sum of one float per thread (blockDim.x multiple of 32, <= 1024). `red` holds 33 floats.
DEVI float block_sum(float v, float* red) {
    v = warp_sum(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    __syncthreads();  // protect `red` from the previous use
    if (lane == 0) red[warp] = v;
    __syncthreads();
    if (warp == 0) {
        float t = lane < nw ? red[lane] : 0.f;
        t = warp_sum(t);
        if (lane == 0) red[32] = t;
    }
    __syncthreads();
    return red[32];
}

// LayerNorm of one row held as 8 values per active thread; fp32 statistics (two-pass), output rounded to fp16.
DEVI H8 layer_norm8(const H8& x, bool active, int C, const __half* w, const __half* b, int c0, float eps, float* red) {
    float s = 0.f;
    if (active) {
#pragma unroll
        for (int e = 0; e < 8; e++) s += x.f[e];
    }
    const float mean = block_sum(s, red) / (float)C;
    float q = 0.f;
    if (active) {
#pragma unroll
        for (int e = 0; e < 8; e++) {
            float d = x.f[e] - mean;
            q += d * d;
        }
    }
    const float var = block_sum(q, red) / (float)C;
    const float rstd = 1.0f / sqrtf(var + eps);
    H8 o;
    if (active) {
        H8 ww = ld8(w + c0), bb = ld8(b + c0);
#pragma unroll
        for (int e = 0; e < 8; e++) o.f[e] = hr((x.f[e] - mean) * rstd * ww.f[e] + bb.f[e]);
    }
    return o;
}

__global__ void embed_gather_kernel(int M, int C, int V, const int* __restrict__ tokens, const __half* __restrict__ emb,
                                    __half* __restrict__ x) {
    pdl_launch_dependents();
    pdl_wait();
    const int m = blockIdx.x;
    const int tok = min(max(tokens[m], 0), V - 1);  // an id outside the vocabulary must not become a wild read
    const __half* src = emb + (size_t)tok * C;
    __half* dst = x + (size_t)m * C;
    for (int c = threadIdx.x * 8; c < C; c += blockDim.x * 8)
        *reinterpret_cast<int4*>(dst + c) = __ldg(reinterpret_cast<const int4*>(src + c));
}

// LayerNorm -> token shift -> NMIX lerps for a STRIP of consecutive tokens of one sequence per CTA:
//   cur = LN(x[b,t]);  prev = t ? LN(x[b,t-1]) : shift_state[b];  out_n[b,t] = cur + (prev - cur) * mix_n;  shift_state[b] <- LN(x[b,T-1])
// 128 threads per CTA, up to four 8-channel chunks per thread (C <= 4096), 64 registers: eight CTAs = eight independent rows
// in flight per SM, each reduction a barrier among four warps only.  (Round 1 ran one 512-thread CTA per row, two per SM, and
// for T > 1 normalised every row twice — once as `cur`, once as the next token's `prev`: 333 + 231 us per layer at 16 384
// tokens, 1.2-2.8 TB/s, as much as the chunked WKV kernel.)  A strip keeps the previous token's LN in registers, so each row is
// loaded and normalised once; only the strip's first token re-normalises its predecessor.  The strip that holds token 0 owns
// the sequence's shift-state row: it reads the old value (before griddepcontrol.wait: it is last written by this kernel one
// step earlier) and stores LN(x[b,T-1]) — computing that row itself when it lies in another strip, so no CTA depends on another.
// Decode (T = 1, strip = 1) may first finish the producing GEMM's deferred K split (glue.h): x <- fp16(x + fp16(p_0 + p_1 + ...)).
// Two shapes of the same kernel: prefill (T > 1) runs 128 threads x four 8-channel chunks per thread as described; a decode
// step (T = 1: 256 rows, at most two CTAs per SM whatever their size) keeps round 1's 512 threads x one chunk, where the row's
// latency, not the number of rows in flight, is what counts.
template <int LN_K>
struct RowF {   // one row's share of this thread in fp32: chunk j = channels (j * LN_T + tid) * 8 .. + 7
    float f[LN_K][8];
};

// fp32 two-pass statistics (as ATen) of a row spread over the CTA; `red` = 2 x 32 floats not used by the previous two calls.
// On return v holds x - mean.  The summation tree is the SAME for both kernel shapes (and the one round 1's kernel used):
// 8 elements of a chunk in order, then a shuffle tree over each group of 32 consecutive chunks, then a shuffle tree over the
// (at most 16) group sums — so a row normalises to the same bits whether it is reached through a T = 1 step or a longer one.
template <int LN_T, int LN_K, bool FULL>
DEVI void ln_stats(RowF<LN_K>& v, int nk, int C, float* red, float& rstd) {
    constexpr int NW = LN_T / 32, NV = NW * LN_K;   // chunk j of warp w is "virtual warp" j * NW + w: chunks 32v .. 32v + 31
    static_assert(NV <= 32, "one shuffle tree over the group sums");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int j = 0; j < LN_K; j++) {
        float s = 0.f;
        if (FULL || j < nk) {
#pragma unroll
            for (int e = 0; e < 8; e++) s += v.f[j][e];
        }
        s = warp_sum(s);
        if (lane == 0) red[j * NW + warp] = s;
    }
    __syncthreads();
    const float mean = warp_sum(lane < NV ? red[lane] : 0.f) / (float)C;
#pragma unroll
    for (int j = 0; j < LN_K; j++) {
        float q = 0.f;
        if (FULL || j < nk) {
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const float d = v.f[j][e] - mean;
                v.f[j][e] = d;
                q += d * d;
            }
        }
        q = warp_sum(q);
        if (lane == 0) red[32 + j * NW + warp] = q;
    }
    __syncthreads();
    rstd = 1.0f / sqrtf(warp_sum(lane < NV ? red[32 + lane] : 0.f) / (float)C + 1e-5f);
}

// FULL: C == LN_T * LN_K * 8, every thread owns LN_K whole chunks (no per-chunk predicate; C = 4096 in both shapes).
// The elementwise half ops after the LayerNorm (prev - cur, * mix, + cur: each an ATen fp16 op = fp32 op + one rounding) run
// as HSUB2 / HMUL2 / HADD2 with contraction disabled (__h*2_rn): for +, -, x of two fp16 values the fp32 intermediate is
// exact or rounds innocuously (24 >= 2 * 11 + 2 bits), so the packed op returns the same bits at a quarter of the
// instructions — the prefill shape of this kernel was issue-bound (57 % issue-active, 2.8 TB/s), not DRAM-bound.
template <int NMIX, int LN_T, int LN_K, bool FULL>
__global__ void __launch_bounds__(LN_T, LN_T == 128 ? 6 : 2) ln_shift_mix_kernel(int T, int C, int strip, __half* __restrict__ x,
                                                              const __half* __restrict__ ln_w, const __half* __restrict__ ln_b,
                                                              __half* __restrict__ shift, long shift_stride,
                                                              const int* __restrict__ slot_idx, MixPtrs<NMIX> mp,
                                                              const float* __restrict__ part, int n_part, long part_stride) {
    __shared__ float red[4][64];
    // prefill shape: this thread's LN affine parameters as fp32, [chunk j][w lo, w hi, b lo, b hi][tid] — thread-private
    // (written and read by the same thread, so no barrier), converted once per CTA instead of once per row
    __shared__ float4 lnc[LN_K > 1 ? 4 * LN_K * LN_T : 1];
    // decode shape only: this thread's 16 bytes of ln_w, ln_b and of every lerp coefficient, fetched by cp.async BEFORE
    // griddepcontrol.wait (they are weights) so that no L2 round trip for them is left on the row's critical path
    extern __shared__ __align__(16) int4 coef[];   // [(2 + NMIX)][LN_T], LN_K == 1 only
    pdl_launch_dependents();
    const int strips = (T + strip - 1) / strip;
    const int b = blockIdx.x / strips, t0 = (blockIdx.x - b * strips) * strip;
    const int t1 = min(T, t0 + strip);
    const int tid = threadIdx.x;
    const int nch = C >> 3;
    int nk = 0;   // chunks of this thread: channels (j * LN_T + tid) * 8, j < nk
#pragma unroll
    for (int j = 0; j < LN_K; j++)
        if (j * LN_T + tid < nch) nk = j + 1;
    __half* srow = shift + (size_t)(slot_idx ? slot_idx[b] : b) * shift_stride;
    int4 prevq[LN_K];   // previous token's LN, packed fp16
#pragma unroll
    for (int j = 0; j < LN_K; j++) prevq[j] = make_int4(0, 0, 0, 0);
    if (t0 == 0) {
#pragma unroll
        for (int j = 0; j < LN_K; j++)
            if (FULL || j < nk) prevq[j] = *reinterpret_cast<const int4*>(srow + (j * LN_T + tid) * 8);
    }
    if (LN_K == 1) {
        if (nk == 1) {
            cp_async16(&coef[0 * LN_T + tid], ln_w + tid * 8);
            cp_async16(&coef[1 * LN_T + tid], ln_b + tid * 8);
#pragma unroll
            for (int n = 0; n < NMIX; n++) cp_async16(&coef[(2 + n) * LN_T + tid], mp.mix[n] + tid * 8);
            cp_async_commit();
        }
    } else {
#pragma unroll
        for (int j = 0; j < LN_K; j++)
            if (FULL || j < nk) {
                const int c0 = (j * LN_T + tid) * 8;
                const H8 ww = cvt8(__ldg(reinterpret_cast<const int4*>(ln_w + c0))), bb = cvt8(__ldg(reinterpret_cast<const int4*>(ln_b + c0)));
                lnc[(j * 4 + 0) * LN_T + tid] = make_float4(ww.f[0], ww.f[1], ww.f[2], ww.f[3]);
                lnc[(j * 4 + 1) * LN_T + tid] = make_float4(ww.f[4], ww.f[5], ww.f[6], ww.f[7]);
                lnc[(j * 4 + 2) * LN_T + tid] = make_float4(bb.f[0], bb.f[1], bb.f[2], bb.f[3]);
                lnc[(j * 4 + 3) * LN_T + tid] = make_float4(bb.f[4], bb.f[5], bb.f[6], bb.f[7]);
            }
    }
    pdl_wait();
    int rb = 0;   // reduction buffer in use (a buffer comes round again after two more rows: every thread is long past it)

    // LN of the row in v (fp32) into packed fp16
    auto normalise = [&](RowF<LN_K>& v, int4 (&outq)[LN_K]) {
        float rstd;
        ln_stats<LN_T, LN_K, FULL>(v, nk, C, red[rb], rstd);
        rb = (rb + 1) & 3;
#pragma unroll
        for (int j = 0; j < LN_K; j++)
            if (FULL || j < nk) {
                float w[8], bb[8];
                if (LN_K == 1) {
                    cp_async_wait_all();
                    const H8 w8 = cvt8(coef[tid]), b8 = cvt8(coef[LN_T + tid]);
#pragma unroll
                    for (int e = 0; e < 8; e++) w[e] = w8.f[e], bb[e] = b8.f[e];
                } else {
                    const float4 w0 = lnc[(j * 4 + 0) * LN_T + tid], w1 = lnc[(j * 4 + 1) * LN_T + tid];
                    const float4 b0 = lnc[(j * 4 + 2) * LN_T + tid], b1 = lnc[(j * 4 + 3) * LN_T + tid];
                    w[0] = w0.x, w[1] = w0.y, w[2] = w0.z, w[3] = w0.w, w[4] = w1.x, w[5] = w1.y, w[6] = w1.z, w[7] = w1.w;
                    bb[0] = b0.x, bb[1] = b0.y, bb[2] = b0.z, bb[3] = b0.w, bb[4] = b1.x, bb[5] = b1.y, bb[6] = b1.z, bb[7] = b1.w;
                }
                __half2* h2 = reinterpret_cast<__half2*>(&outq[j]);
#pragma unroll
                for (int e = 0; e < 4; e++)
                    h2[e] = __floats2half2_rn(v.f[j][2 * e] * rstd * w[2 * e] + bb[2 * e],
                                              v.f[j][2 * e + 1] * rstd * w[2 * e + 1] + bb[2 * e + 1]);
            }
    };
    auto load_row = [&](const __half* row, RowF<LN_K>& v) {
        int4 q[LN_K];
#pragma unroll
        for (int j = 0; j < LN_K; j++)
            if (FULL || j < nk) q[j] = *reinterpret_cast<const int4*>(row + (j * LN_T + tid) * 8);
#pragma unroll
        for (int j = 0; j < LN_K; j++)
            if (FULL || j < nk) {
                const H8 h = cvt8(q[j]);
#pragma unroll
                for (int e = 0; e < 8; e++) v.f[j][e] = h.f[e];
            }
    };

    RowF<LN_K> v;
    if (t0 > 0) {   // the strip's first token needs LN of the token before it (another strip's row: normalised once more here)
        load_row(x + ((size_t)b * T + t0 - 1) * C, v);
        normalise(v, prevq);
    }
    for (int t = t0; t < t1; t++) {
        const size_t m = (size_t)b * T + t;
        __half* xrow = x + m * C;
        load_row(xrow, v);
        if (LN_K == 1 && part != nullptr && nk == 1) {
            // deferred K split of the producing GEMM (T == 1, decode shape): x <- fp16(x + fp16(p_0 + p_1 + ...)) in split
            // order, row written back
            const int c0 = tid * 8;
            const float* pr = part + m * C + c0;
            float acc[8];
#pragma unroll
            for (int e = 0; e < 8; e++) acc[e] = 0.f;
#pragma unroll 1
            for (int z0 = 0; z0 < n_part; z0 += 4) {   // up to four splits (all there are, in practice) in ONE round trip
                float4 lo[4], hi[4];
#pragma unroll
                for (int u = 0; u < 4; u++)
                    if (z0 + u < n_part) {
                        const float4* q = reinterpret_cast<const float4*>(pr + (size_t)(z0 + u) * part_stride);
                        lo[u] = __ldcg(q);
                        hi[u] = __ldcg(q + 1);
                    }
#pragma unroll
                for (int u = 0; u < 4; u++)
                    if (z0 + u < n_part) {
                        acc[0] += lo[u].x; acc[1] += lo[u].y; acc[2] += lo[u].z; acc[3] += lo[u].w;
                        acc[4] += hi[u].x; acc[5] += hi[u].y; acc[6] += hi[u].z; acc[7] += hi[u].w;
                    }
            }
            int4 xq;
            __half2* h2 = reinterpret_cast<__half2*>(&xq);
#pragma unroll
            for (int e = 0; e < 4; e++) {
                h2[e] = __floats2half2_rn(v.f[0][2 * e] + hr(acc[2 * e]), v.f[0][2 * e + 1] + hr(acc[2 * e + 1]));
                const float2 r = __half22float2(h2[e]);
                v.f[0][2 * e] = r.x, v.f[0][2 * e + 1] = r.y;
            }
            *reinterpret_cast<int4*>(xrow + c0) = xq;
        }
        int4 curq[LN_K];
        normalise(v, curq);
#pragma unroll
        for (int j = 0; j < LN_K; j++)
            if (FULL || j < nk) {
                const int c0 = (j * LN_T + tid) * 8;
                const __half2* c2 = reinterpret_cast<const __half2*>(&curq[j]);
                const __half2* p2 = reinterpret_cast<const __half2*>(&prevq[j]);
                __half2 dx[4];
#pragma unroll
                for (int e = 0; e < 4; e++) dx[e] = __hsub2_rn(p2[e], c2[e]);
#pragma unroll
                for (int n = 0; n < NMIX; n++) {
                    const int4 mq = LN_K == 1 ? coef[(2 + n) * LN_T + tid] : __ldg(reinterpret_cast<const int4*>(mp.mix[n] + c0));
                    const __half2* m2 = reinterpret_cast<const __half2*>(&mq);
                    int4 oq;
                    __half2* o2 = reinterpret_cast<__half2*>(&oq);
#pragma unroll
                    for (int e = 0; e < 4; e++) o2[e] = __hadd2_rn(c2[e], __hmul2_rn(dx[e], m2[e]));
                    *reinterpret_cast<int4*>(mp.out[n] + m * C + c0) = oq;
                }
                prevq[j] = curq[j];
            }
    }
    if (t0 == 0) {   // new shift state = LN(x[b, T-1])
        if (t1 < T) {
            load_row(x + ((size_t)b * T + T - 1) * C, v);
            normalise(v, prevq);
        }
#pragma unroll
        for (int j = 0; j < LN_K; j++)
            if (FULL || j < nk) *reinterpret_cast<int4*>(srow + (j * LN_T + tid) * 8) = prevq[j];
    }
}

// LayerNorm of selected rows: out[i] = LN(x[i*row_step + row_off]).
__global__ void __launch_bounds__(1024) ln_rows_kernel(int C, const __half* __restrict__ x, long row_step, long row_off,
                                                       const __half* __restrict__ ln_w, const __half* __restrict__ ln_b,
                                                       __half* __restrict__ out) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float red[33];
    const int c0 = threadIdx.x * 8;
    const bool active = c0 < C;
    const __half* xrow = x + ((size_t)blockIdx.x * row_step + row_off) * C;
    H8 xv;
    if (active) xv = ld8(xrow + c0);
    const H8 o = layer_norm8(xv, active, C, ln_w, ln_b, c0, 1e-5f, red);
    if (active) st8(out + (size_t)blockIdx.x * C + c0, o);
}

__global__ void add_elapsed_kernel(int B, int T, int* __restrict__ elapsed) {
    pdl_launch_dependents();
    pdl_wait();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < B) elapsed[i] += T;
}

DEVI float2 ld2(const __half* p) { return __half22float2(*reinterpret_cast<const __half2*>(p)); }
DEVI void st2(__half* p, float x, float y) { *reinterpret_cast<__half2*>(p) = __floats2half2_rn(x, y); }

// One warp per (token m, head h); lane holds channels h*64 + 2*lane, +1.                         rwkv7.py:632-637
__global__ void __launch_bounds__(128) tmix_prep_kernel(TmixPrepArgs p) {
    pdl_launch_dependents();
    pdl_wait();
    const int H = p.C >> 6;
    const int gw = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (gw >= p.M * H) return;
    const int lane = threadIdx.x & 31;
    const int m = gw / H, h = gw - m * H;
    const int c = h * 64 + lane * 2;
    const size_t i = (size_t)m * p.C + c;
    const float2 k = ld2(p.k + i), a = ld2(p.a + i), kk_w = ld2(p.k_k + c), ka_w = ld2(p.k_a + c);
    // kk = normalize(k * k_k) per head: norm accumulated in fp32 and rounded to fp16 (x.norm() on a half tensor), then divided
    const float q0 = hr(k.x * kk_w.x), q1 = hr(k.y * kk_w.y);
    const float nrm = hr(sqrtf(warp_sum(q0 * q0 + q1 * q1)));
    const float kk0 = hr(q0 / nrm), kk1 = hr(q1 / nrm);
    // k = k * (1 + (a - 1) * k_a)
    const float k0 = hr(k.x * hr(1.0f + hr(hr(a.x - 1.0f) * ka_w.x)));
    const float k1 = hr(k.y * hr(1.0f + hr(hr(a.y - 1.0f) * ka_w.y)));
    st2(p.k_out + i, k0, k1);
    st2(p.a_out + i, -kk0, -kk1);
    st2(p.b_out + i, hr(kk0 * a.x), hr(kk1 * a.y));
    if (p.v_first != nullptr) {  // v = v + (v_first - v) * sigmoid(...)
        const float2 v = ld2(p.v + i), vf = ld2(p.v_first + i), gt = ld2(p.vgate + i);
        st2(p.v_out + i, hr(v.x + hr(hr(vf.x - v.x) * gt.x)), hr(v.y + hr(hr(vf.y - v.y) * gt.y)));
    }
}

// One warp per (token, head): GroupNorm over the head's 64 outputs, bonus term, gate.               rwkv7.py:647-649
__global__ void __launch_bounds__(128) tmix_post_kernel(TmixPostArgs p) {
    pdl_launch_dependents();
    pdl_wait();
    const int H = p.C >> 6;
    const int gw = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (gw >= p.M * H) return;
    const int lane = threadIdx.x & 31;
    const int m = gw / H, h = gw - m * H;
    const int c = h * 64 + lane * 2;
    const size_t i = (size_t)m * p.C + c;
    const float2 y = ld2(p.y + i);
    const float mean = warp_sum(y.x + y.y) * (1.0f / 64.0f);
    const float d0 = y.x - mean, d1 = y.y - mean;
    const float var = warp_sum(d0 * d0 + d1 * d1) * (1.0f / 64.0f);
    const float rstd = 1.0f / sqrtf(var + 64e-5f);
    const float2 gw_ = ld2(p.ln_w + c), gb_ = ld2(p.ln_b + c);
    const float n0 = hr(d0 * rstd * gw_.x + gb_.x), n1 = hr(d1 * rstd * gw_.y + gb_.y);
    const float2 r = ld2(p.r + i), k = ld2(p.k + i), v = ld2(p.v + i), g = ld2(p.g + i), rk = ld2(p.r_k + c);
    const float s = hr(warp_sum(hr(hr(r.x * k.x) * rk.x) + hr(hr(r.y * k.y) * rk.y)));
    const float o0 = hr(n0 + hr(s * v.x)), o1 = hr(n1 + hr(s * v.y));
    st2(p.out + i, hr(o0 * g.x), hr(o1 * g.y));
}

// Greedy sampling: out[b] = first index of the maximum of logits[b, :] compared as fp32 (== torch.argmax(logits.float(), -1),
// the worker's top_k == 1 / temperature == 0 path, chirrup/utils/samplers.py:171-255).  One CTA per row.
__global__ void __launch_bounds__(256) argmax_rows_kernel(int V, const __half* __restrict__ logits, long ld,
                                                          int* __restrict__ out) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sv[8];
    __shared__ int si[8];
    const __half* row = logits + (size_t)blockIdx.x * ld;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int c = threadIdx.x * 8; c < V; c += 256 * 8) {
        int4 q = *reinterpret_cast<const int4*>(row + c);
        const __half* h = reinterpret_cast<const __half*>(&q);
#pragma unroll
        for (int e = 0; e < 8; e++) {
            const float v = __half2float(h[e]);
            if (v > best || (v == best && c + e < bi)) best = v, bi = c + e;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) best = ov, bi = oi;
    }
    if ((threadIdx.x & 31) == 0) sv[threadIdx.x >> 5] = best, si[threadIdx.x >> 5] = bi;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++)
            if (sv[w] > best || (sv[w] == best && si[w] < bi)) best = sv[w], bi = si[w];
        out[blockIdx.x] = bi;
    }
}

int ln_threads(int C) {
    int t = (C / 8 + 31) / 32 * 32;
    return t < 32 ? 32 : t;
}

}  // namespace

int launch_embed_gather(int M, int C, int V, const int* tokens, const __half* emb, __half* x, cudaStream_t s, bool pdl) {
    if (M <= 0) return 0;
    return launch_kernel(embed_gather_kernel, dim3(M), dim3(128), 0, s, pdl, 0, M, C, V, tokens, emb, x);
}
template <int NMIX, bool FULL>
int launch_ln_shift_mix_t(int B, int T, int C, __half* x, const __half* ln_w, const __half* ln_b, __half* shift, long shift_stride,
                          const int* slot_idx, const MixPtrs<NMIX>& mp, cudaStream_t s, bool pdl, const float* part, int n_part) {
    if (T == 1) {
        const size_t coef_bytes = (size_t)(2 + NMIX) * 512 * 16;
        static bool attr_done[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64 || !attr_done[dev]) {
            cudaError_t e = cudaFuncSetAttribute(ln_shift_mix_kernel<NMIX, 512, 1, FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)coef_bytes);
            if (e != cudaSuccess) return (int)e;
            if (dev >= 0 && dev < 64) attr_done[dev] = true;
        }
        return launch_kernel(ln_shift_mix_kernel<NMIX, 512, 1, FULL>, dim3(B), dim3(512), coef_bytes, s, pdl, 0, T, C, 1, x, ln_w, ln_b,
                             shift, shift_stride, slot_idx, mp, part, n_part, (long)B * T * C);
    }
    const int strip = T >= 8 ? 8 : T;
    return launch_kernel(ln_shift_mix_kernel<NMIX, 128, 4, FULL>, dim3(B * ((T + strip - 1) / strip)), dim3(128), 0, s, pdl, 0, T, C,
                         strip, x, ln_w, ln_b, shift, shift_stride, slot_idx, mp, part, n_part, (long)B * T * C);
}
template <int NMIX>
int launch_ln_shift_mix_n(int B, int T, int C, __half* x, const __half* ln_w, const __half* ln_b, __half* shift, long shift_stride,
                          const int* slot_idx, const MixPtrs<NMIX>& mp, cudaStream_t s, bool pdl, const float* part, int n_part) {
    if (B * T <= 0) return 0;
    if (C > 4096) return -4;  // B200_ERR_UNSUPPORTED: one thread per 8 channels, 512 threads (every RWKV-7 g1 size has C <= 4096)
    if (part && (T != 1 || n_part < 1)) return -1;
    return C == 4096 ? launch_ln_shift_mix_t<NMIX, true>(B, T, C, x, ln_w, ln_b, shift, shift_stride, slot_idx, mp, s, pdl, part, n_part)
                     : launch_ln_shift_mix_t<NMIX, false>(B, T, C, x, ln_w, ln_b, shift, shift_stride, slot_idx, mp, s, pdl, part, n_part);
}
int launch_ln_shift_mix6(int B, int T, int C, __half* x, const __half* ln_w, const __half* ln_b, __half* shift,
                         long shift_stride, const int* slot_idx, const MixPtrs<6>& mp, cudaStream_t s, bool pdl, const float* part,
                         int n_part) {
    return launch_ln_shift_mix_n<6>(B, T, C, x, ln_w, ln_b, shift, shift_stride, slot_idx, mp, s, pdl, part, n_part);
}
int launch_ln_shift_mix1(int B, int T, int C, __half* x, const __half* ln_w, const __half* ln_b, __half* shift,
                         long shift_stride, const int* slot_idx, const MixPtrs<1>& mp, cudaStream_t s, bool pdl, const float* part,
                         int n_part) {
    return launch_ln_shift_mix_n<1>(B, T, C, x, ln_w, ln_b, shift, shift_stride, slot_idx, mp, s, pdl, part, n_part);
}
int launch_ln_rows(int rows, int C, const __half* x, long row_step, long row_off, const __half* ln_w, const __half* ln_b,
                   __half* out, cudaStream_t s, bool pdl) {
    if (rows <= 0) return 0;
    return launch_kernel(ln_rows_kernel, dim3(rows), dim3(ln_threads(C)), 0, s, pdl, 0, C, x, row_step, row_off, ln_w, ln_b, out);
}
int launch_add_elapsed(int B, int T, int* elapsed, cudaStream_t s, bool pdl) {
    if (B <= 0) return 0;
    return launch_kernel(add_elapsed_kernel, dim3((B + 127) / 128), dim3(128), 0, s, pdl, 0, B, T, elapsed);
}
int launch_argmax_rows(int rows, int V, const __half* logits, long ld, int* out, cudaStream_t s) {
    if (rows <= 0) return 0;
    argmax_rows_kernel<<<rows, 256, 0, s>>>(V, logits, ld, out);
    return (int)cudaGetLastError();
}
int launch_tmix_prep(const TmixPrepArgs& a, cudaStream_t s, bool pdl) {
    const int warps = a.M * (a.C >> 6);
    if (warps <= 0) return 0;
    return launch_kernel(tmix_prep_kernel, dim3((warps + 3) / 4), dim3(128), 0, s, pdl, 0, a);
}
int launch_tmix_post(const TmixPostArgs& a, cudaStream_t s, bool pdl) {
    const int warps = a.M * (a.C >> 6);
    if (warps <= 0) return 0;
    return launch_kernel(tmix_post_kernel, dim3((warps + 3) / 4), dim3(128), 0, s, pdl, 0, a);
}

}  // namespace b200

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

// Sum of one float per thread over the block; every warp reduces the per-warp partials itself, so one barrier per call.
// `part` must be a buffer (32 floats) not used by the previous block_sum1 call.
DEVI float block_sum1(float v, float* part) {
    v = warp_sum(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    if (lane == 0) part[warp] = v;
    __syncthreads();
    return warp_sum(lane < nw ? part[lane] : 0.f);
}

// LayerNorm (eps 1e-5, two-pass fp32 statistics as ATen) of one row held as 8 values per active thread, affine parameters
// already in registers; result rounded to fp16.  Uses red[0..31] and red[32..63].
DEVI H8 layer_norm8_pre(const H8& x, bool active, int C, const int4& wq, const int4& bq, float* red) {
    float s = 0.f;
    if (active) {
#pragma unroll
        for (int e = 0; e < 8; e++) s += x.f[e];
    }
    const float mean = block_sum1(s, red) / (float)C;
    float q = 0.f;
    if (active) {
#pragma unroll
        for (int e = 0; e < 8; e++) {
            float d = x.f[e] - mean;
            q += d * d;
        }
    }
    const float var = block_sum1(q, red + 32) / (float)C;
    const float rstd = 1.0f / sqrtf(var + 1e-5f);
    const H8 ww = cvt8(wq), bb = cvt8(bq);
    H8 o;
#pragma unroll
    for (int e = 0; e < 8; e++) o.f[e] = hr((x.f[e] - mean) * rstd * ww.f[e] + bb.f[e]);
    return o;
}

// One CTA per (sequence b, token t) row.  cur = LN(x[b,t]); prev = t ? LN(x[b,t-1]) : shift_state[b];
// out_n = cur + (prev - cur) * mix_n.  The t == 0 CTA also owns the shift state of its sequence: it reads the old value and
// afterwards stores LN(x[b,T-1]) — no other CTA touches that row, so there is no ordering hazard for T > 1.
// Everything that the preceding kernels of the step do not produce — LN affine, the lerp coefficients and the shift-state row
// (last written by this same kernel one step earlier) — is fetched BEFORE griddepcontrol.wait, so under programmatic dependent
// launch only the x row load and two one-barrier reductions remain on the critical path.
template <int NMIX>
__global__ void __launch_bounds__(512, 2) ln_shift_mix_kernel(int T, int C, __half* __restrict__ x,
                                                            const __half* __restrict__ ln_w, const __half* __restrict__ ln_b,
                                                            __half* __restrict__ shift, long shift_stride,
                                                            const int* __restrict__ slot_idx, MixPtrs<NMIX> mp,
                                                            const float* __restrict__ part, int n_part, long part_stride) {
    __shared__ float red[3 * 64];
    pdl_launch_dependents();
    const int m = blockIdx.x;
    const int b = m / T, t = m - b * T;
    const int c0 = threadIdx.x * 8;
    const bool active = c0 < C;
    __half* xrow = x + (size_t)m * C;
    __half* srow = shift + (size_t)(slot_idx ? slot_idx[b] : b) * shift_stride;
    // kept packed (8 fp16 per int4) until used: 9 x 4 registers instead of 9 x 8
    const int4 z4 = make_int4(0, 0, 0, 0);
    // 64 registers per thread: two CTAs per SM, so the 256 rows of a decode step are ONE wave (and a CTA fits beside a GEMM CTA
    // of the previous kernel under programmatic dependent launch).  The lerp coefficients are therefore fetched where they are
    // used (L1/L2-resident: every CTA reads the same 6 x 8 KiB), not held across the LayerNorm.
    int4 ww = z4, bb = z4, prevq = z4;
    if (active) {
        ww = *reinterpret_cast<const int4*>(ln_w + c0);
        bb = *reinterpret_cast<const int4*>(ln_b + c0);
        if (t == 0) prevq = *reinterpret_cast<const int4*>(srow + c0);
    }
    pdl_wait();
    H8 prev = cvt8(prevq);
    H8 xv;
#pragma unroll
    for (int e = 0; e < 8; e++) xv.f[e] = 0.f;
    if (active) xv = ld8(xrow + c0);
    if (part != nullptr && active) {
        // The producing GEMM left its K split unreduced (glue.h, linear_f16_launch defer_splits): finish the residual add here,
        // x <- fp16(x + fp16(p_0 + p_1 + ...)) in split order — bit for bit what the GEMM's own reduction stores — and write the
        // row back for the residual stream.  T == 1 only (no CTA reads another CTA's row).
        const float* pr = part + (size_t)m * C + c0;
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; e++) acc[e] = 0.f;
#pragma unroll 1
        for (int z0 = 0; z0 < n_part; z0 += 2) {
            float4 lo[2], hi[2];
#pragma unroll
            for (int j = 0; j < 2; j++)
                if (z0 + j < n_part) {
                    const float4* q = reinterpret_cast<const float4*>(pr + (size_t)(z0 + j) * part_stride);
                    lo[j] = __ldcg(q);
                    hi[j] = __ldcg(q + 1);
                }
#pragma unroll
            for (int j = 0; j < 2; j++)
                if (z0 + j < n_part) {
                    acc[0] += lo[j].x; acc[1] += lo[j].y; acc[2] += lo[j].z; acc[3] += lo[j].w;
                    acc[4] += hi[j].x; acc[5] += hi[j].y; acc[6] += hi[j].z; acc[7] += hi[j].w;
                }
        }
#pragma unroll
        for (int e = 0; e < 8; e++) xv.f[e] = hr(xv.f[e] + hr(acc[e]));
        st8(xrow + c0, xv);
    }
    const H8 cur = layer_norm8_pre(xv, active, C, ww, bb, red);
    if (t > 0) {
        H8 pv = xv;
        if (active) pv = ld8(xrow - C + c0);
        prev = layer_norm8_pre(pv, active, C, ww, bb, red + 64);
    } else {
        H8 last = cur;
        if (T > 1) {
            H8 lv = xv;
            if (active) lv = ld8(xrow + (size_t)(T - 1) * C + c0);
            last = layer_norm8_pre(lv, active, C, ww, bb, red + 64);
        }
        if (active) st8(srow + c0, last);
    }
    if (!active) return;
    H8 dx;
#pragma unroll
    for (int e = 0; e < 8; e++) dx.f[e] = hr(prev.f[e] - cur.f[e]);
#pragma unroll
    for (int n = 0; n < NMIX; n++) {
        const H8 mv = cvt8(__ldg(reinterpret_cast<const int4*>(mp.mix[n] + c0)));
        H8 o;
#pragma unroll
        for (int e = 0; e < 8; e++) o.f[e] = hr(cur.f[e] + hr(dx.f[e] * mv.f[e]));
        st8(mp.out[n] + (size_t)m * C + c0, o);
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
int launch_ln_shift_mix6(int B, int T, int C, __half* x, const __half* ln_w, const __half* ln_b, __half* shift,
                         long shift_stride, const int* slot_idx, const MixPtrs<6>& mp, cudaStream_t s, bool pdl, const float* part,
                         int n_part) {
    if (B * T <= 0) return 0;
    if (C > 4096) return -4;  // B200_ERR_UNSUPPORTED: one thread per 8 channels, 512 threads (every RWKV-7 g1 size has C <= 4096)
    if (part && (T != 1 || n_part < 1)) return -1;
    return launch_kernel(ln_shift_mix_kernel<6>, dim3(B * T), dim3(ln_threads(C)), 0, s, pdl, 0, T, C, x, ln_w, ln_b, shift,
                         shift_stride, slot_idx, mp, part, n_part, (long)B * T * C);
}
int launch_ln_shift_mix1(int B, int T, int C, __half* x, const __half* ln_w, const __half* ln_b, __half* shift,
                         long shift_stride, const int* slot_idx, const MixPtrs<1>& mp, cudaStream_t s, bool pdl, const float* part,
                         int n_part) {
    if (B * T <= 0) return 0;
    if (C > 4096) return -4;
    if (part && (T != 1 || n_part < 1)) return -1;
    return launch_kernel(ln_shift_mix_kernel<1>, dim3(B * T), dim3(ln_threads(C)), 0, s, pdl, 0, T, C, x, ln_w, ln_b, shift,
                         shift_stride, slot_idx, mp, part, n_part, (long)B * T * C);
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

// WKV-7 state-forward kernels for sm_100a.
//
// Replaces the reference's kernel_forward_w0_fp16_dither_seq / _one
// (/root/reference/Albatross/cuda/rwkv7_state_fwd_fp16.cu:26-97, :99-167, launchers :313-322) behind the same
// argument meaning.  The arithmetic is the reference's, op for op (fp16 half2 chains, even/odd lane split, one rounding per
// op, the elapsed_t dither of .cu:20-23) so the result is bit-identical to the reference kernel; what is new is the data
// movement: the 64x64 fp16 state tile is staged global -> shared by one TMA tensor copy with the hardware 128-byte swizzle
// (conflict-free 128-bit row reads, no software transpose), updated in registers, written back through the same swizzled
// buffer with one TMA store.  A slot-index vector lets the caller address state rows indirectly (no physical slot swaps).
#include <cstdlib>

#include "common.cuh"
#include "tmap.h"
#include "glue.h"
#include "../../include/chirrup_b200.h"

namespace b200 {
namespace {

constexpr int HS = 64;  // head size (reference: _N_ = 64, Albatross/rwkv7.py:48)

// reference .cu:20-23, :59 — as compiled with --use_fast_math: FMUL, MUFU.EX2, FADD, MUFU.RCP, FMUL, MUFU.EX2, FADD, FFMA, F2F
DEVI __half decay_dither(__half w_pre, int t) {
    const float nlog2_e = -1.4426950408889634f;
    const float nexp_half_log2_e = -0.8750387749145276f;
    const float two_to_neg_41 = 4.547473508864641e-13f;
    float x = __fmul_rn(__half2float(w_pre), nlog2_e);
    float d = __fadd_rn(ex2_approx_ftz(x), 1.0f);
    float q = __fmul_rn(rcp_approx_ftz(d), nexp_half_log2_e);
    float e = __fadd_rn(ex2_approx_ftz(q), -1.0f);
    int rot = static_cast<int>(2654435769u * static_cast<uint32_t>(t));
    return __float2half_rn(__fmaf_rn(static_cast<float>(rot), two_to_neg_41, e));
}

struct TokenRegs {
    __half r, w, k, v, a, b;
};

// One CTA = one (sequence, head) tile, 64 threads = one thread per value row i (holds S[i][0..63] as 32 half2).
__global__ void __launch_bounds__(HS) wkv7_seq_kernel(const __grid_constant__ CUtensorMap tm_state, int T, int C, int H,
                                                      int rows_per_slot, const __half* __restrict__ r_,
                                                      const __half* __restrict__ w_, const __half* __restrict__ k_,
                                                      const __half* __restrict__ v_, const __half* __restrict__ a_,
                                                      const __half* __restrict__ b_, __half* __restrict__ y_,
                                                      const int* __restrict__ elapsed_t, const int* __restrict__ slot_idx) {
    __shared__ __align__(1024) uint8_t tile[HS * HS * 2];
    __shared__ __align__(16) __half sv[5][HS];  // w~, k, a, b, r of the current token
    __shared__ __align__(8) uint64_t bar;

    pdl_launch_dependents();
    const int bb = blockIdx.x / H;
    const int h = blockIdx.x - bb * H;
    const int i = threadIdx.x;
    const int slot = slot_idx ? slot_idx[bb] : bb;
    const int row0 = slot * rows_per_slot + h * HS;

    if (i == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    if (i == 0) {
        // the state tile (like slot_idx / elapsed_t) is not produced by the preceding kernels of the step: under PDL its load
        // starts while they are still draining
        mbar_arrive_expect_tx(&bar, HS * HS * 2);
        tma_load_2d(tile, &tm_state, 0, row0, &bar);
    }

    const int t0 = elapsed_t[bb];
    size_t idx = (size_t)bb * T * C + (size_t)h * HS + i;
    pdl_wait();
    TokenRegs cur;
    cur.r = r_[idx]; cur.w = w_[idx]; cur.k = k_[idx]; cur.v = v_[idx]; cur.a = a_[idx]; cur.b = b_[idx];

    mbar_wait(&bar, 0);
    __half2 S[HS / 2];
    {
        const int4* rowp = reinterpret_cast<const int4*>(tile + i * 128);
        const int sw = i & 7;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            int4 q = rowp[c ^ sw];
            S[4 * c + 0] = *reinterpret_cast<__half2*>(&q.x);
            S[4 * c + 1] = *reinterpret_cast<__half2*>(&q.y);
            S[4 * c + 2] = *reinterpret_cast<__half2*>(&q.z);
            S[4 * c + 3] = *reinterpret_cast<__half2*>(&q.w);
        }
    }

    for (int t = 0; t < T; t++) {
        if (t) __syncthreads();  // previous token's readers are done with sv
        sv[0][i] = decay_dither(cur.w, t0 + t);
        sv[1][i] = cur.k;
        sv[2][i] = cur.a;
        sv[3][i] = cur.b;
        sv[4][i] = cur.r;
        const __half vv = cur.v;
        __syncthreads();
        const size_t yidx = idx;
        if (t + 1 < T) {  // prefetch the next token under this token's math
            idx += C;
            cur.r = r_[idx]; cur.w = w_[idx]; cur.k = k_[idx]; cur.v = v_[idx]; cur.a = a_[idx]; cur.b = b_[idx];
        }
        const int4* wv = reinterpret_cast<const int4*>(sv[0]);
        const int4* kv = reinterpret_cast<const int4*>(sv[1]);
        const int4* av = reinterpret_cast<const int4*>(sv[2]);
        const int4* bv = reinterpret_cast<const int4*>(sv[3]);
        const int4* rv = reinterpret_cast<const int4*>(sv[4]);

        // sa_i = sum_j a_j * S_ij on the pre-update state: one half2 fma chain, lo = even j, hi = odd j (.cu:65-70)
        __half2 acc = __floats2half2_rn(0.f, 0.f);
#pragma unroll
        for (int c = 0; c < 8; c++) {
            int4 q = av[c];
            const __half2* p = reinterpret_cast<const __half2*>(&q);
#pragma unroll
            for (int u = 0; u < 4; u++) acc = __hfma2(p[u], S[4 * c + u], acc);
        }
        const __half sa = __hadd_rn(__low2half(acc), __high2half(acc));
        const __half2 sa2 = __half2half2(sa);
        const __half2 vv2 = __half2half2(vv);

        // S_ij <- S_ij + fma(sa, b_j, fma(S_ij, w~_j, k_j * v_i)) ; y_i = sum_j S_ij(new) * r_j   (.cu:72-81)
        acc = __floats2half2_rn(0.f, 0.f);
#pragma unroll
        for (int c = 0; c < 8; c++) {
            int4 qw = wv[c], qk = kv[c], qb = bv[c], qr = rv[c];
            const __half2* pw = reinterpret_cast<const __half2*>(&qw);
            const __half2* pk = reinterpret_cast<const __half2*>(&qk);
            const __half2* pb = reinterpret_cast<const __half2*>(&qb);
            const __half2* pr = reinterpret_cast<const __half2*>(&qr);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                __half2 s = S[4 * c + u];
                __half2 d = __hmul2_rn(pk[u], vv2);
                d = __hfma2(s, pw[u], d);
                d = __hfma2(sa2, pb[u], d);
                s = __hadd2_rn(s, d);
                S[4 * c + u] = s;
                acc = __hfma2(s, pr[u], acc);
            }
        }
        y_[yidx] = __hadd_rn(__low2half(acc), __high2half(acc));
    }

    {
        int4* rowp = reinterpret_cast<int4*>(tile + i * 128);
        const int sw = i & 7;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            int4 q;
            q.x = *reinterpret_cast<int*>(&S[4 * c + 0]);
            q.y = *reinterpret_cast<int*>(&S[4 * c + 1]);
            q.z = *reinterpret_cast<int*>(&S[4 * c + 2]);
            q.w = *reinterpret_cast<int*>(&S[4 * c + 3]);
            rowp[c ^ sw] = q;
        }
    }
    fence_proxy_async();  // generic-proxy smem writes -> visible to the TMA (async proxy) store
    __syncthreads();
    if (i == 0) {
        tma_store_2d(&tm_state, 0, row0, tile);
        tma_store_commit();
        tma_store_wait_read<0>();  // smem must stay valid until the bulk store has read it
    }
}


// ---------------------------------------------------------------------------------------------------------------------
// Fused TMix core (decode and prefill): everything between the projection GEMMs and the output GEMM in one kernel —
// kk = normalize(k*k_k), k-mix, v-mix, a = -kk, b = kk*a (rwkv7.py:632-637); the WKV recurrence above (:645); GroupNorm over the
// head's 64 outputs (eps 64e-5), bonus (sum r*k*r_k) v and the gate (:647-649).  Same arithmetic, op for op, as the unfused
// tmix_prep / wkv7_seq / tmix_post chain (glue.cu): every half op computed in fp32 and rounded (`hr`), WKV in fp16 half2.
// While the 8 KiB state tile is in flight (TMA) the 64 threads of the CTA (thread i = channel i of the head = state row i)
// compute the vector prep; nothing but the tile, 8 input rows of 128 B and one output row of 128 B touches HBM.
DEVI float hr(float x) { return __half2float(__float2half_rn(x)); }

struct FusedIn {
    float r, w, k, v, a, g, vg, vf;
};

DEVI void fused_load(FusedIn& f, const WkvFusedArgs& p, size_t idx) {
    f.r = __half2float(p.r[idx]); f.w = __half2float(p.w[idx]); f.k = __half2float(p.k[idx]); f.v = __half2float(p.v[idx]);
    f.a = __half2float(p.a[idx]); f.g = __half2float(p.g[idx]);
    if (p.v_first) {
        f.vg = __half2float(p.vgate[idx]);
        f.vf = __half2float(p.v_first[idx]);
    }
}

__global__ void __launch_bounds__(HS, 14) wkv7_fused_kernel(const __grid_constant__ CUtensorMap tm_state, const WkvFusedArgs p) {
    __shared__ __align__(1024) uint8_t tile[HS * HS * 2];
    __shared__ __align__(16) __half sv[5][HS];  // w~, k, a, b, r of the current token
    __shared__ float red[4][2];
    __shared__ __align__(8) uint64_t bar;

    pdl_launch_dependents();
    const int T = p.T, C = p.C, H = p.H;
    const int bb = blockIdx.x / H;
    const int h = blockIdx.x - bb * H;
    const int i = threadIdx.x;
    const int wi = i >> 5;
    const int slot = p.slot_idx ? p.slot_idx[bb] : bb;
    const int row0 = slot * p.rows_per_slot + h * HS;

    if (i == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    if (i == 0) {
        mbar_arrive_expect_tx(&bar, HS * HS * 2);
        tma_load_2d(tile, &tm_state, 0, row0, &bar);
    }
    const int t0 = p.elapsed_t[bb];
    const int c = h * HS + i;
    const float kk_w = __half2float(p.k_k[c]), ka_w = __half2float(p.k_a[c]), rk_w = __half2float(p.r_k[c]);
    const float ln_w = __half2float(p.ln_w[c]), ln_b = __half2float(p.ln_b[c]);
    size_t idx = (size_t)bb * T * C + c;
    // PDL: the state-tile load and the per-channel weights above do not depend on the preceding kernels; r..g below do
    pdl_wait();
    FusedIn cur;
    fused_load(cur, p, idx);

    __half2 S[HS / 2];
    bool have_state = false;

    for (int t = 0; t < T; t++) {
        if (t) __syncthreads();  // previous token's readers are done with sv / red
        // ---- vector prep (rwkv7.py:632-637)
        const float q = hr(cur.k * kk_w);
        const float k2 = hr(cur.k * hr(1.0f + hr(hr(cur.a - 1.0f) * ka_w)));
        const float v2 = p.v_first ? hr(cur.v + hr(hr(cur.vf - cur.v) * cur.vg)) : cur.v;
        const float bonus_i = hr(hr(cur.r * k2) * rk_w);
        const float s0 = warp_sum(q * q), s1 = warp_sum(bonus_i);
        if ((i & 31) == 0) red[0][wi] = s0, red[1][wi] = s1;
        __syncthreads();
        const float nrm = hr(sqrtf(red[0][0] + red[0][1]));
        const float bonus = hr(red[1][0] + red[1][1]);
        const float kk = hr(q / nrm);
        sv[0][i] = decay_dither(__float2half_rn(cur.w), t0 + t);
        sv[1][i] = __float2half_rn(k2);
        sv[2][i] = __float2half_rn(-kk);
        sv[3][i] = __float2half_rn(hr(kk * cur.a));
        sv[4][i] = __float2half_rn(cur.r);
        const __half vv = __float2half_rn(v2);
        const float gate = cur.g;
        __syncthreads();
        const size_t oidx = idx;
        if (t + 1 < T) {
            idx += C;
            fused_load(cur, p, idx);
        }
        if (!have_state) {  // the tile has had the whole prep to arrive
            mbar_wait(&bar, 0);
            const int4* rowp = reinterpret_cast<const int4*>(tile + i * 128);
            const int sw = i & 7;
#pragma unroll
            for (int cc = 0; cc < 8; cc++) {
                int4 qq = rowp[cc ^ sw];
                S[4 * cc + 0] = *reinterpret_cast<__half2*>(&qq.x);
                S[4 * cc + 1] = *reinterpret_cast<__half2*>(&qq.y);
                S[4 * cc + 2] = *reinterpret_cast<__half2*>(&qq.z);
                S[4 * cc + 3] = *reinterpret_cast<__half2*>(&qq.w);
            }
            have_state = true;
        }
        // ---- WKV-7 (identical to wkv7_seq_kernel)
        const int4* wv = reinterpret_cast<const int4*>(sv[0]);
        const int4* kv = reinterpret_cast<const int4*>(sv[1]);
        const int4* av = reinterpret_cast<const int4*>(sv[2]);
        const int4* bv = reinterpret_cast<const int4*>(sv[3]);
        const int4* rv = reinterpret_cast<const int4*>(sv[4]);
        __half2 acc = __floats2half2_rn(0.f, 0.f);
#pragma unroll
        for (int cc = 0; cc < 8; cc++) {
            int4 qq = av[cc];
            const __half2* pp = reinterpret_cast<const __half2*>(&qq);
#pragma unroll
            for (int u = 0; u < 4; u++) acc = __hfma2(pp[u], S[4 * cc + u], acc);
        }
        const __half2 sa2 = __half2half2(__hadd_rn(__low2half(acc), __high2half(acc)));
        const __half2 vv2 = __half2half2(vv);
        acc = __floats2half2_rn(0.f, 0.f);
#pragma unroll
        for (int cc = 0; cc < 8; cc++) {
            int4 qw = wv[cc], qk = kv[cc], qb = bv[cc], qr = rv[cc];
            const __half2* pw = reinterpret_cast<const __half2*>(&qw);
            const __half2* pk = reinterpret_cast<const __half2*>(&qk);
            const __half2* pb = reinterpret_cast<const __half2*>(&qb);
            const __half2* pr = reinterpret_cast<const __half2*>(&qr);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                __half2 s = S[4 * cc + u];
                __half2 d = __hmul2_rn(pk[u], vv2);
                d = __hfma2(s, pw[u], d);
                d = __hfma2(sa2, pb[u], d);
                s = __hadd2_rn(s, d);
                S[4 * cc + u] = s;
                acc = __hfma2(s, pr[u], acc);
            }
        }
        const float y = __half2float(__hadd_rn(__low2half(acc), __high2half(acc)));
        if (t == T - 1) {
            // last token: send the new state on its way (registers -> swizzled tile -> one TMA store) BEFORE the output math, so
            // that the write-back drains while the GroupNorm / bonus / gate reductions run
            int4* rowp = reinterpret_cast<int4*>(tile + i * 128);
            const int sw = i & 7;
#pragma unroll
            for (int cc = 0; cc < 8; cc++) {
                int4 qq;
                qq.x = *reinterpret_cast<int*>(&S[4 * cc + 0]);
                qq.y = *reinterpret_cast<int*>(&S[4 * cc + 1]);
                qq.z = *reinterpret_cast<int*>(&S[4 * cc + 2]);
                qq.w = *reinterpret_cast<int*>(&S[4 * cc + 3]);
                rowp[cc ^ sw] = qq;
            }
            fence_proxy_async();
        }
        // ---- GroupNorm over the head (fp32 statistics from one pass: sum and sum of squares; eps 64e-5) + bonus + gate
        //      (rwkv7.py:647-649)
        const float m0 = warp_sum(y), m1 = warp_sum(y * y);
        if ((i & 31) == 0) red[2][wi] = m0, red[3][wi] = m1;
        __syncthreads();
        if (t == T - 1 && i == 0) {
            tma_store_2d(&tm_state, 0, row0, tile);
            tma_store_commit();
        }
        const float mean = (red[2][0] + red[2][1]) * (1.0f / 64.0f);
        const float var = fmaxf((red[3][0] + red[3][1]) * (1.0f / 64.0f) - mean * mean, 0.0f);
        const float rstd = 1.0f / sqrtf(var + 64e-5f);
        const float gn = hr((y - mean) * rstd * ln_w + ln_b);
        const float o = hr(gn + hr(bonus * v2));
        p.out[oidx] = __float2half_rn(hr(o * gate));
    }
    if (i == 0) tma_store_wait_read<0>();  // the tile must stay valid until the bulk store has read it
}

// Decode form (T = 1) of the fused TMix core: same arithmetic, op for op, as wkv7_fused_kernel, with the token loop, the
// next-token prefetch and every value that would have to stay live across the WKV chain removed from the registers (the
// per-thread outputs of the vector prep wait in shared memory), so that 16 CTAs fit on an SM instead of 12: 128 KiB of state
// loads in flight per SM.  This is the launch the decode step makes 32 times.
__global__ void __launch_bounds__(HS, 16) wkv7_fused_one_kernel(const __grid_constant__ CUtensorMap tm_state, const WkvFusedArgs p) {
    __shared__ __align__(1024) uint8_t tile[HS * HS * 2];
    __shared__ __align__(16) __half sv[5][HS];  // w~, k, a, b, r
    __shared__ float keep[3][HS];               // v', bonus share, gate of thread i (parked across the WKV chain)
    __shared__ float red[4][2];
    __shared__ __align__(8) uint64_t bar;

    pdl_launch_dependents();
    const int C = p.C, H = p.H;
    const int bb = blockIdx.x / H;
    const int h = blockIdx.x - bb * H;
    const int i = threadIdx.x;
    const int wi = i >> 5;
    const int slot = p.slot_idx ? p.slot_idx[bb] : bb;
    const int row0 = slot * p.rows_per_slot + h * HS;

    if (i == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    if (i == 0) {
        mbar_arrive_expect_tx(&bar, HS * HS * 2);
        tma_load_2d(tile, &tm_state, 0, row0, &bar);
    }
    const int t0 = p.elapsed_t[bb];
    const int c = h * HS + i;
    const size_t idx = (size_t)bb * C + c;
    {
        const float kk_w = __half2float(p.k_k[c]), ka_w = __half2float(p.k_a[c]), rk_w = __half2float(p.r_k[c]);
        pdl_wait();  // the state tile and the per-channel weights do not depend on the preceding kernels; r..g below do
        FusedIn cur;
        fused_load(cur, p, idx);
        // ---- vector prep (rwkv7.py:632-637)
        const float q = hr(cur.k * kk_w);
        const float k2 = hr(cur.k * hr(1.0f + hr(hr(cur.a - 1.0f) * ka_w)));
        const float v2 = p.v_first ? hr(cur.v + hr(hr(cur.vf - cur.v) * cur.vg)) : cur.v;
        const float bonus_i = hr(hr(cur.r * k2) * rk_w);
        const float s0 = warp_sum(q * q), s1 = warp_sum(bonus_i);
        if ((i & 31) == 0) red[0][wi] = s0, red[1][wi] = s1;
        __syncthreads();
        const float nrm = hr(sqrtf(red[0][0] + red[0][1]));
        const float kk = hr(q / nrm);
        sv[0][i] = decay_dither(__float2half_rn(cur.w), t0);
        sv[1][i] = __float2half_rn(k2);
        sv[2][i] = __float2half_rn(-kk);
        sv[3][i] = __float2half_rn(hr(kk * cur.a));
        sv[4][i] = __float2half_rn(cur.r);
        keep[0][i] = v2;
        keep[1][i] = cur.g;
        __syncthreads();
    }
    // ---- WKV-7 (identical to wkv7_seq_kernel)
    mbar_wait(&bar, 0);
    __half2 S[HS / 2];
    {
        const int4* rowp = reinterpret_cast<const int4*>(tile + i * 128);
        const int sw = i & 7;
#pragma unroll
        for (int cc = 0; cc < 8; cc++) {
            int4 qq = rowp[cc ^ sw];
            S[4 * cc + 0] = *reinterpret_cast<__half2*>(&qq.x);
            S[4 * cc + 1] = *reinterpret_cast<__half2*>(&qq.y);
            S[4 * cc + 2] = *reinterpret_cast<__half2*>(&qq.z);
            S[4 * cc + 3] = *reinterpret_cast<__half2*>(&qq.w);
        }
    }
    const int4* wv = reinterpret_cast<const int4*>(sv[0]);
    const int4* kv = reinterpret_cast<const int4*>(sv[1]);
    const int4* av = reinterpret_cast<const int4*>(sv[2]);
    const int4* bv = reinterpret_cast<const int4*>(sv[3]);
    const int4* rv = reinterpret_cast<const int4*>(sv[4]);
    __half2 acc = __floats2half2_rn(0.f, 0.f);
#pragma unroll
    for (int cc = 0; cc < 8; cc++) {
        int4 qq = av[cc];
        const __half2* pp = reinterpret_cast<const __half2*>(&qq);
#pragma unroll
        for (int u = 0; u < 4; u++) acc = __hfma2(pp[u], S[4 * cc + u], acc);
    }
    const __half2 sa2 = __half2half2(__hadd_rn(__low2half(acc), __high2half(acc)));
    const __half2 vv2 = __half2half2(__float2half_rn(keep[0][i]));
    acc = __floats2half2_rn(0.f, 0.f);
    {
        int4* rowp = reinterpret_cast<int4*>(tile + i * 128);
        const int sw = i & 7;
#pragma unroll
        for (int cc = 0; cc < 8; cc++) {
            int4 qw = wv[cc], qk = kv[cc], qb = bv[cc], qr = rv[cc];
            const __half2* pw = reinterpret_cast<const __half2*>(&qw);
            const __half2* pk = reinterpret_cast<const __half2*>(&qk);
            const __half2* pb = reinterpret_cast<const __half2*>(&qb);
            const __half2* pr = reinterpret_cast<const __half2*>(&qr);
            int4 out;
            __half2* po = reinterpret_cast<__half2*>(&out);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                __half2 s = S[4 * cc + u];
                __half2 d = __hmul2_rn(pk[u], vv2);
                d = __hfma2(s, pw[u], d);
                d = __hfma2(sa2, pb[u], d);
                s = __hadd2_rn(s, d);
                po[u] = s;
                acc = __hfma2(s, pr[u], acc);
            }
            rowp[cc ^ sw] = out;  // new state straight back into the swizzled tile
        }
    }
    fence_proxy_async();
    const float y = __half2float(__hadd_rn(__low2half(acc), __high2half(acc)));
    // ---- GroupNorm over the head (eps 64e-5) + bonus + gate (rwkv7.py:647-649); the state write-back drains underneath
    const float m0 = warp_sum(y), m1 = warp_sum(y * y);
    if ((i & 31) == 0) red[2][wi] = m0, red[3][wi] = m1;
    __syncthreads();
    if (i == 0) {
        tma_store_2d(&tm_state, 0, row0, tile);
        tma_store_commit();
    }
    const float mean = (red[2][0] + red[2][1]) * (1.0f / 64.0f);
    const float var = fmaxf((red[3][0] + red[3][1]) * (1.0f / 64.0f) - mean * mean, 0.0f);
    const float rstd = 1.0f / sqrtf(var + 64e-5f);
    const float gn = hr((y - mean) * rstd * __half2float(p.ln_w[c]) + __half2float(p.ln_b[c]));
    const float bonus = hr(red[1][0] + red[1][1]);
    const float o = hr(gn + hr(bonus * keep[0][i]));
    p.out[idx] = __float2half_rn(hr(o * keep[1][i]));
    if (i == 0) tma_store_wait_read<0>();  // the tile must stay valid until the bulk store has read it
}

}  // namespace

int wkv7_make_state_tmap(CUtensorMap* tm, const void* state, int n_slots, int H, long state_slot_stride) {
    if (state_slot_stride % HS != 0 || n_slots < 1) return B200_ERR_BAD_ARG;
    const uint64_t rows_per_slot = (uint64_t)state_slot_stride / HS;
    const uint64_t rows = (uint64_t)(n_slots - 1) * rows_per_slot + (uint64_t)H * HS;
    return make_tmap_2d_f16(tm, state, rows, HS, HS * 2, HS, HS, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
}

int launch_wkv7_fused(const WkvFusedArgs& a, void* state, long state_slot_stride, int n_slots, cudaStream_t stream, bool pdl) {
    if (a.B <= 0 || a.T <= 0) return 0;
    if (a.H * HS != a.C) return B200_ERR_BAD_ARG;
    CUtensorMap tm;
    int rc = wkv7_make_state_tmap(&tm, state, a.slot_idx ? n_slots : a.B, a.H, state_slot_stride);
    if (rc) return rc;
    WkvFusedArgs p = a;
    p.rows_per_slot = (int)(state_slot_stride / HS);
    static const bool generic_only = getenv("B200_WKV_GENERIC") != nullptr;  // A/B switch
    if (a.T == 1 && !generic_only) return launch_kernel(wkv7_fused_one_kernel, dim3(a.B * a.H), dim3(HS), 0, stream, pdl, 0, tm, p);
    return launch_kernel(wkv7_fused_kernel, dim3(a.B * a.H), dim3(HS), 0, stream, pdl, 0, tm, p);
}

}  // namespace b200

using namespace b200;

namespace b200 {
int launch_wkv7_seq(int B, int T, int C, int H, void* state, long state_slot_stride, int n_slots, const void* r, const void* w,
                    const void* k, const void* v, const void* a, const void* b, void* y, const int* elapsed_t,
                    const int* slot_idx, cudaStream_t stream, bool pdl) {
    if (B < 0 || T < 0 || H <= 0 || H * HS != C) return B200_ERR_BAD_ARG;  // reference: assert(H*_N_ == C), .cu:314
    if (B == 0 || T == 0) return 0;
    if (!slot_idx) n_slots = B;
    if ((reinterpret_cast<uintptr_t>(state) & 15) != 0) return B200_ERR_BAD_ARG;
    CUtensorMap tm;
    int rc = wkv7_make_state_tmap(&tm, state, n_slots, H, state_slot_stride);
    if (rc) return rc;
    return launch_kernel(wkv7_seq_kernel, dim3(B * H), dim3(HS), 0, stream, pdl, 0, tm, T, C, H, (int)(state_slot_stride / HS),
                         static_cast<const __half*>(r), static_cast<const __half*>(w), static_cast<const __half*>(k),
                         static_cast<const __half*>(v), static_cast<const __half*>(a), static_cast<const __half*>(b),
                         static_cast<__half*>(y), elapsed_t, slot_idx);
}
}  // namespace b200

extern "C" int b200_wkv7_forward_seq(int B, int T, int C, int H, void* state, long state_slot_stride, int n_slots,
                                     const void* r, const void* w, const void* k, const void* v, const void* a,
                                     const void* b, void* y, const int* elapsed_t, const int* slot_idx, void* stream) {
    return b200::launch_wkv7_seq(B, T, C, H, state, state_slot_stride, n_slots, r, w, k, v, a, b, y, elapsed_t, slot_idx,
                                 static_cast<cudaStream_t>(stream), false);
}

extern "C" int b200_wkv7_forward_one(int B, int C, int H, void* state, long state_slot_stride, int n_slots, const void* r,
                                     const void* w, const void* k, const void* v, const void* a, const void* b, void* y,
                                     const int* elapsed_t, const int* slot_idx, void* stream) {
    return b200_wkv7_forward_seq(B, 1, C, H, state, state_slot_stride, n_slots, r, w, k, v, a, b, y, elapsed_t, slot_idx,
                                 stream);
}

extern "C" int b200_abi_version(void) { return 1; }

extern "C" int b200_tmix_core_fused(int B, int T, int C, int H, void* state, long state_slot_stride, int n_slots, const void* r,
                                    const void* w, const void* k, const void* v, const void* a, const void* g,
                                    const void* vgate, const void* v_first, const void* k_k, const void* k_a, const void* r_k,
                                    const void* ln_w, const void* ln_b, void* out, const int* elapsed_t, const int* slot_idx,
                                    void* stream) {
    if (B < 0 || T < 0 || H <= 0 || H * HS != C) return B200_ERR_BAD_ARG;
    if ((v_first == nullptr) != (vgate == nullptr)) return B200_ERR_BAD_ARG;
    auto h = [](const void* p) { return static_cast<const __half*>(p); };
    WkvFusedArgs fa;
    fa.B = B; fa.T = T; fa.C = C; fa.H = H; fa.rows_per_slot = 0;
    fa.r = h(r); fa.w = h(w); fa.k = h(k); fa.v = h(v); fa.a = h(a); fa.g = h(g); fa.vgate = h(vgate); fa.v_first = h(v_first);
    fa.k_k = h(k_k); fa.k_a = h(k_a); fa.r_k = h(r_k); fa.ln_w = h(ln_w); fa.ln_b = h(ln_b);
    fa.out = static_cast<__half*>(out); fa.elapsed_t = elapsed_t; fa.slot_idx = slot_idx;
    return launch_wkv7_fused(fa, state, state_slot_stride, n_slots, static_cast<cudaStream_t>(stream));
}

// Chunked WKV-7 prefill kernel for sm_100a: the fused TMix core (vector prep + WKV-7 + GroupNorm / bonus / gate, the same
// inputs and outputs as wkv7_fused_kernel in wkv7.cu) with the recurrence evaluated 16 tokens at a time in its parallel
// ("chunked", WY / UT-transform) form on the tensor cores instead of token by token.
//
// Reference semantics: kernel_forward_w0_fp16_dither_seq (/root/reference/Albatross/cuda/rwkv7_state_fwd_fp16.cu:26-97) is a
// strictly sequential T loop with two barriers per token (.cu:56-82); the reference has no chunked form.
// Per head (N = 64), S[i][j] (i value row, j key column), per token t:
//     d_t = 1 + w~_t            w~ = fp16(decay - 1 + dither(elapsed_t + t))                       (.cu:20-23, :59)
//     u_t = S_{t-1} a_t                                                                            (.cu:65-70, "sa")
//     S_t = S_{t-1} diag(d_t) + u_t b_t^T + v_t k_t^T                                              (.cu:78)
//     y_t = S_t r_t                                                                                (.cu:79-81)
// Chunk of 16 tokens from S_0 with c_t = sum_{s<=t} log2 d_s per channel and a per-channel reference point m = c_7 (keeps
// every rescaled operand within 0.545^(+-8) of its value: fp16-safe):
//     a~_t = a_t 2^(c_{t-1}-m)   r~_t = r_t 2^(c_t-m)   b~_s = b_s 2^(m-c_s)   k~_s = k_s 2^(m-c_s)   S0' = S_0 diag(2^m)
//     U = (S0' A~ + V Aak)(I - Aab)^-1        Aab[s,t] = b~_s.a~_t (s<t)    Aak[s,t] = k~_s.a~_t (s<t)
//     Y = S0' R~ + U Abr + V Akr              Abr[s,t] = b~_s.r~_t (s<=t)   Akr[s,t] = k~_s.r~_t (s<=t)
//     S_16 = (S0' + U B~^T + V K~^T) diag(2^(c_16 - m))
// (tools/wkv_chunk_emulate.py checks this algebra and the operand roundings against the sequential fp16 oracle: the chunked
// form is CLOSER to the exact recurrence than the reference kernel is — it rounds the state to fp16 once per launch, not
// once per token — so what separates the two is the reference's own fp16 noise, ~6e-4 relative on the state.)
//
// One CTA (4 warps) per (sequence, head).  Every product runs on warp-level tensor-core MMAs (mma.sync m16n8k16, fp16
// operands, fp32 accumulate): the operands are 16..64-wide per head, far below a tcgen05 tile (M >= 64 per CTA, operands in
// shared memory, accumulators in TMEM), and the chain S0'A~ -> U -> Y -> state feeds each accumulator straight back as the
// next operand IN REGISTERS (C-fragment -> A-fragment), which TMEM accumulators cannot do without a round trip.
// The state lives in fp32 accumulator fragments for the whole launch; it is staged in / out through the same TMA tile as the
// sequential kernels.  Decay, dither and every elementwise op before / after the recurrence are the reference's, op for op.
// Tolerance-only parity (the sequential kernels stay the bit-exact mode: B200_WKV_SEQUENTIAL=1).
#include <cstdlib>

#include "common.cuh"
#include "tmap.h"
#include "glue.h"
#include "../../include/chirrup_b200.h"

namespace b200 {
int wkv7_make_state_tmap(CUtensorMap* tm, const void* state, int n_slots, int H, long state_slot_stride);

namespace {

constexpr int HS = 64;
constexpr int TC = 16;        // tokens per chunk
constexpr int NT = 128;       // threads per CTA
constexpr int LDO = 72;       // halves per row of the 64-wide operand tiles (+8: conflict-free fragment loads)
constexpr int LDS16 = 24;     // halves per row of the 16-wide operand tiles
constexpr int LDY = 65;       // floats per row of the Y staging tile

DEVI float hr(float x) { return __half2float(__float2half_rn(x)); }

// reference .cu:20-23, :59 as compiled (see wkv7.cu)
DEVI __half decay_dither_c(__half w_pre, int t) {
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

DEVI uint32_t pack2(float lo, float hi) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}
DEVI uint32_t lds32(const __half* p) { return *reinterpret_cast<const uint32_t*>(p); }
DEVI uint32_t pack_u16(const __half* lo, const __half* hi) {
    return (uint32_t)(*reinterpret_cast<const unsigned short*>(lo)) | ((uint32_t)(*reinterpret_cast<const unsigned short*>(hi)) << 16);
}

// D = A (16x16, row) * B (16x8, col) + D, fp16 operands, fp32 accumulate
DEVI void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

struct ChunkSmem {
    alignas(1024) uint8_t tile[HS * HS * 2];   // state tile, TMA SWIZZLE_128B layout
    float L[TC][HS];                           // log2 d, then its inclusive cumulative sum over the chunk's tokens
    float em[HS], ecTm[HS];                    // 2^m, 2^(c_16 - m) per channel
    alignas(16) __half opT[2 * TC][LDO];       // rows 0..15 a~_t, rows 16..31 r~_t           [t][j]
    alignas(16) __half opS[2 * TC][LDO];       // rows 0..15 b~_s, rows 16..31 k~_s           [s][j]
    alignas(16) __half vT[HS][LDS16];          // v' transposed                               [i][s]
    alignas(16) __half aakT[TC][LDS16];        // Aak, Abr, Akr, (I - Aab)^-1 as B operands   [t][s]
    alignas(16) __half abrT[TC][LDS16];
    alignas(16) __half akrT[TC][LDS16];
    alignas(16) __half tinvT[TC][LDS16];
    float aab[TC][TC + 1];                     // Aab in fp32 for the triangular inverse      [s][t]
    float ys[TC][LDY];                         // y of the chunk                               [t][i]
    alignas(16) __half v2s[TC][HS];            // v' and gate per token for the output stage   [t][i]
    alignas(16) __half gs[TC][HS];
    float bonus[TC];
    alignas(8) uint64_t bar;
};

__global__ void __launch_bounds__(NT, 4) wkv7_chunk_kernel(const __grid_constant__ CUtensorMap tm_state, const WkvFusedArgs p) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    ChunkSmem& sm = *reinterpret_cast<ChunkSmem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));   // TMA swizzle atom alignment

    pdl_launch_dependents();
    const int T = p.T, C = p.C, H = p.H;
    const int bb = blockIdx.x / H;
    const int h = blockIdx.x - bb * H;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, tq = lane & 3;          // mma fragment coordinates
    const int slot = p.slot_idx ? p.slot_idx[bb] : bb;
    const int row0 = slot * p.rows_per_slot + h * HS;

    if (tid == 0) {
        mbar_init(&sm.bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    if (tid == 0) {
        mbar_arrive_expect_tx(&sm.bar, HS * HS * 2);
        tma_load_2d(sm.tile, &tm_state, 0, row0, &sm.bar);
    }
    const int t_base = p.elapsed_t[bb];
    // prep mapping: thread -> (token tk of the chunk, 8 consecutive channels)
    const int tk = tid >> 3, cg = tid & 7;
    const int ch0 = h * HS + cg * 8;
    // per-channel weights are re-read (L1-resident) where they are used instead of living in 40 registers next to the state
    auto ldw8 = [&](const __half* base, float (&out)[8]) {
        const int4 q = *reinterpret_cast<const int4*>(base + ch0);
        const __half* hh = reinterpret_cast<const __half*>(&q);
#pragma unroll
        for (int e = 0; e < 8; e++) out[e] = __half2float(hh[e]);
    };
    pdl_wait();  // the state tile and the per-channel weights do not depend on the preceding kernels; r..g below do

    // ---- the state: fp32 accumulator fragments for the whole launch.  Warp w owns rows 16w..16w+15; S[n][0..3] = rows
    //      (16w+g, 16w+g+8) x columns (8n+2tq, 8n+2tq+1)
    float S[8][4];
    mbar_wait(&sm.bar, 0);
    {
        const int r_lo = 16 * warp + g, r_hi = r_lo + 8;
#pragma unroll
        for (int n = 0; n < 8; n++) {
            const __half2 lo = *reinterpret_cast<const __half2*>(sm.tile + r_lo * 128 + ((n ^ (r_lo & 7)) << 4) + 4 * tq);
            const __half2 hi = *reinterpret_cast<const __half2*>(sm.tile + r_hi * 128 + ((n ^ (r_hi & 7)) << 4) + 4 * tq);
            S[n][0] = __low2float(lo); S[n][1] = __high2float(lo);
            S[n][2] = __low2float(hi); S[n][3] = __high2float(hi);
        }
    }

    for (int t0 = 0; t0 < T; t0 += TC) {
        // ================================================================ 1. vector prep of 16 tokens (rwkv7.py:632-637)
        const int t = t0 + tk;
        const bool live = t < T;
        float a_vec[8], b_vec[8], k2[8], rr[8], ld[8];
        {
            float kin[8], vin[8], ain[8], win[8], gin[8], vg[8], vf[8], kk_w[8], ka_w[8], rk_w[8];
            ldw8(p.k_k, kk_w); ldw8(p.k_a, ka_w); ldw8(p.r_k, rk_w);
            const size_t idx = ((size_t)bb * T + (live ? t : 0)) * C + ch0;
            auto ld8 = [&](const __half* base, float (&out)[8]) {
                const int4 q = *reinterpret_cast<const int4*>(base + idx);
                const __half* hh = reinterpret_cast<const __half*>(&q);
#pragma unroll
                for (int e = 0; e < 8; e++) out[e] = __half2float(hh[e]);
            };
            ld8(p.r, rr); ld8(p.w, win); ld8(p.k, kin); ld8(p.v, vin); ld8(p.a, ain); ld8(p.g, gin);
            if (p.v_first) {
                ld8(p.vgate, vg);
                ld8(p.v_first, vf);
            }
            float q[8], v2[8], ssq = 0.f, sbon = 0.f;
#pragma unroll
            for (int e = 0; e < 8; e++) {
                q[e] = hr(kin[e] * kk_w[e]);
                k2[e] = hr(kin[e] * hr(1.0f + hr(hr(ain[e] - 1.0f) * ka_w[e])));
                v2[e] = p.v_first ? hr(vin[e] + hr(hr(vf[e] - vin[e]) * vg[e])) : vin[e];
                ssq += q[e] * q[e];
                sbon += hr(hr(rr[e] * k2[e]) * rk_w[e]);
            }
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {   // the 8 threads of a token are 8 consecutive lanes
                ssq += __shfl_xor_sync(0xffffffffu, ssq, o);
                sbon += __shfl_xor_sync(0xffffffffu, sbon, o);
            }
            const float nrm = hr(sqrtf(ssq));
            __half v2h[8], gh[8];
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const float kk = hr(q[e] / nrm);
                a_vec[e] = -kk;
                b_vec[e] = hr(kk * ain[e]);
                const float wt = __half2float(decay_dither_c(__float2half_rn(win[e]), t_base + t));
                ld[e] = __log2f(1.0f + wt);
                if (!live) {   // past the end of the sequence: an identity step (d = 1, no update, no read-out)
                    a_vec[e] = 0.f; b_vec[e] = 0.f; k2[e] = 0.f; v2[e] = 0.f; rr[e] = 0.f; ld[e] = 0.f;
                }
                v2h[e] = __float2half_rn(v2[e]);
                gh[e] = __float2half_rn(gin[e]);
                sm.L[tk][cg * 8 + e] = ld[e];
                sm.vT[cg * 8 + e][tk] = v2h[e];
            }
            *reinterpret_cast<int4*>(&sm.v2s[tk][cg * 8]) = *reinterpret_cast<const int4*>(v2h);
            *reinterpret_cast<int4*>(&sm.gs[tk][cg * 8]) = *reinterpret_cast<const int4*>(gh);
            if (cg == 0) sm.bonus[tk] = hr(sbon);
        }
        __syncthreads();
        // ---- per channel: inclusive cumulative log2-decay over the chunk, reference point m = c_7
        if (tid < HS) {
            float c = 0.f, m = 0.f;
#pragma unroll
            for (int s = 0; s < TC; s++) {
                c += sm.L[s][tid];
                sm.L[s][tid] = c;
                if (s == TC / 2 - 1) m = c;
            }
            sm.em[tid] = exp2f(m);
            sm.ecTm[tid] = exp2f(c - m);
        }
        __syncthreads();
        // ---- rescaled operands
        {
            __half at[8], rt[8], bt[8], kt[8];
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const int j = cg * 8 + e;
                const float c = sm.L[tk][j];
                const float m = sm.L[TC / 2 - 1][j];
                const float up = exp2f(c - m), dn = exp2f(m - c);
                at[e] = __float2half_rn(a_vec[e] * exp2f(c - ld[e] - m));
                rt[e] = __float2half_rn(rr[e] * up);
                bt[e] = __float2half_rn(b_vec[e] * dn);
                kt[e] = __float2half_rn(k2[e] * dn);
            }
            *reinterpret_cast<int4*>(&sm.opT[tk][cg * 8]) = *reinterpret_cast<const int4*>(at);
            *reinterpret_cast<int4*>(&sm.opT[TC + tk][cg * 8]) = *reinterpret_cast<const int4*>(rt);
            *reinterpret_cast<int4*>(&sm.opS[tk][cg * 8]) = *reinterpret_cast<const int4*>(bt);
            *reinterpret_cast<int4*>(&sm.opS[TC + tk][cg * 8]) = *reinterpret_cast<const int4*>(kt);
        }
        __syncthreads();

        // ================================================================ 2. the four 16x16 token-token matrices (K = 64)
        //      warp 0: Aab = B~ A~^T   warp 1: Aak = K~ A~^T   warp 2: Abr = B~ R~^T   warp 3: Akr = K~ R~^T
        {
            const int mt = warp & 1, np = warp >> 1;
            float acc[2][4] = {};
#pragma unroll
            for (int kt = 0; kt < 4; kt++) {
                uint32_t a[4];
                a[0] = lds32(&sm.opS[16 * mt + g][16 * kt + 2 * tq]);
                a[1] = lds32(&sm.opS[16 * mt + g + 8][16 * kt + 2 * tq]);
                a[2] = lds32(&sm.opS[16 * mt + g][16 * kt + 2 * tq + 8]);
                a[3] = lds32(&sm.opS[16 * mt + g + 8][16 * kt + 2 * tq + 8]);
#pragma unroll
                for (int nt = 0; nt < 2; nt++) {
                    const __half* brow = &sm.opT[16 * np + 8 * nt + g][16 * kt + 2 * tq];
                    mma16816(acc[nt], a, lds32(brow), lds32(brow + 8));
                }
            }
            // fragment (row s = g / g+8, col t = 8nt + 2tq / +1); keep s < t (A~ products) or s <= t (R~ products)
#pragma unroll
            for (int nt = 0; nt < 2; nt++) {
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const int s = g + 8 * (e >> 1), tt = 8 * nt + 2 * tq + (e & 1);
                    const bool keep = np == 0 ? (s < tt) : (s <= tt);
                    const float v = keep ? acc[nt][e] : 0.f;
                    if (warp == 0) sm.aab[s][tt] = v;
                    else if (warp == 1) sm.aakT[tt][s] = __float2half_rn(v);
                    else if (warp == 2) sm.abrT[tt][s] = __float2half_rn(v);
                    else sm.akrT[tt][s] = __float2half_rn(v);
                }
            }
        }
        __syncthreads();

        // ================================================================ 3. (I - Aab)^-1 (warp 0, fp32) under the S0' products
        if (warp == 0 && lane < TC) {
            // lane = column t of X = (I - A)^-1, A strictly upper: X[t][t] = 1, X[s][t] = sum_{q > s} A[s][q] X[q][t] for s < t
            float x[TC];
#pragma unroll
            for (int q = 0; q < TC; q++) x[q] = q == lane ? 1.f : 0.f;
#pragma unroll
            for (int s = TC - 2; s >= 0; s--) {
                float acc = 0.f;
#pragma unroll
                for (int q = s + 1; q < TC; q++) acc = fmaf(sm.aab[s][q], x[q], acc);
                if (s < lane) x[s] = acc;
            }
            __half xh[TC];
#pragma unroll
            for (int q = 0; q < TC; q++) xh[q] = __float2half_rn(x[q]);
            *reinterpret_cast<int4*>(&sm.tinvT[lane][0]) = *reinterpret_cast<const int4*>(&xh[0]);
            *reinterpret_cast<int4*>(&sm.tinvT[lane][8]) = *reinterpret_cast<const int4*>(&xh[8]);
        }
        __syncwarp();
        // S <- S0' = S diag(2^m) (fp32); its fp16 rounding is the A operand of  SA|SR = S0' [A~ ; R~]^T  (16 x 32 per warp, K = 64)
        float acc[4][4] = {};   // n-tiles 0,1: S0' A~ (t = 0..15)   n-tiles 2,3: S0' R~
        {
#pragma unroll
            for (int n = 0; n < 8; n++) {
                const float2 e = *reinterpret_cast<const float2*>(&sm.em[8 * n + 2 * tq]);
                S[n][0] *= e.x; S[n][1] *= e.y; S[n][2] *= e.x; S[n][3] *= e.y;
            }
#pragma unroll
            for (int kt = 0; kt < 4; kt++) {
                uint32_t a[4];
                a[0] = pack2(S[2 * kt][0], S[2 * kt][1]);
                a[1] = pack2(S[2 * kt][2], S[2 * kt][3]);
                a[2] = pack2(S[2 * kt + 1][0], S[2 * kt + 1][1]);
                a[3] = pack2(S[2 * kt + 1][2], S[2 * kt + 1][3]);
#pragma unroll
                for (int nt = 0; nt < 4; nt++) {
                    const __half* brow = &sm.opT[8 * nt + g][16 * kt + 2 * tq];
                    mma16816(acc[nt], a, lds32(brow), lds32(brow + 8));
                }
            }
        }
        __syncthreads();

        // ================================================================ 4. rhs -> U -> Y -> state (per warp: its 16 value rows)
        {
            uint32_t va[4];   // V as A operand: rows i, k = s
            const int i_lo = 16 * warp + g;
            va[0] = lds32(&sm.vT[i_lo][2 * tq]);
            va[1] = lds32(&sm.vT[i_lo + 8][2 * tq]);
            va[2] = lds32(&sm.vT[i_lo][2 * tq + 8]);
            va[3] = lds32(&sm.vT[i_lo + 8][2 * tq + 8]);
            // rhs = S0' A~ + V Aak
#pragma unroll
            for (int nt = 0; nt < 2; nt++) mma16816(acc[nt], va, lds32(&sm.aakT[8 * nt + g][2 * tq]), lds32(&sm.aakT[8 * nt + g][2 * tq + 8]));
            // U = rhs (I - Aab)^-1
            uint32_t ra[4] = {pack2(acc[0][0], acc[0][1]), pack2(acc[0][2], acc[0][3]), pack2(acc[1][0], acc[1][1]),
                              pack2(acc[1][2], acc[1][3])};
            float u[2][4] = {};
#pragma unroll
            for (int nt = 0; nt < 2; nt++) mma16816(u[nt], ra, lds32(&sm.tinvT[8 * nt + g][2 * tq]), lds32(&sm.tinvT[8 * nt + g][2 * tq + 8]));
            uint32_t ua[4] = {pack2(u[0][0], u[0][1]), pack2(u[0][2], u[0][3]), pack2(u[1][0], u[1][1]), pack2(u[1][2], u[1][3])};
            // Y = S0' R~ + U Abr + V Akr
#pragma unroll
            for (int nt = 0; nt < 2; nt++) {
                mma16816(acc[2 + nt], ua, lds32(&sm.abrT[8 * nt + g][2 * tq]), lds32(&sm.abrT[8 * nt + g][2 * tq + 8]));
                mma16816(acc[2 + nt], va, lds32(&sm.akrT[8 * nt + g][2 * tq]), lds32(&sm.akrT[8 * nt + g][2 * tq + 8]));
            }
#pragma unroll
            for (int nt = 0; nt < 2; nt++) {
#pragma unroll
                for (int e = 0; e < 4; e++) sm.ys[8 * nt + 2 * tq + (e & 1)][i_lo + 8 * (e >> 1)] = acc[2 + nt][e];
            }
            // S <- (S0' + U B~^T + V K~^T) diag(2^(c_16 - m))   (16 x 64 per warp, K = 32: s' = [u-tokens ; v-tokens])
#pragma unroll
            for (int n = 0; n < 8; n++) {
                const int j = 8 * n + g;
                const uint32_t b0 = pack_u16(&sm.opS[2 * tq][j], &sm.opS[2 * tq + 1][j]);
                const uint32_t b1 = pack_u16(&sm.opS[2 * tq + 8][j], &sm.opS[2 * tq + 9][j]);
                mma16816(S[n], ua, b0, b1);
                const uint32_t b2 = pack_u16(&sm.opS[TC + 2 * tq][j], &sm.opS[TC + 2 * tq + 1][j]);
                const uint32_t b3 = pack_u16(&sm.opS[TC + 2 * tq + 8][j], &sm.opS[TC + 2 * tq + 9][j]);
                mma16816(S[n], va, b2, b3);
                const float2 e = *reinterpret_cast<const float2*>(&sm.ecTm[8 * n + 2 * tq]);
                S[n][0] *= e.x; S[n][1] *= e.y; S[n][2] *= e.x; S[n][3] *= e.y;
            }
        }
        __syncthreads();

        // ================================================================ 5. GroupNorm (eps 64e-5) + bonus + gate (rwkv7.py:647-649)
        {
            float y[8], s1 = 0.f, s2 = 0.f, ln_w[8], ln_b[8];
            ldw8(p.ln_w, ln_w); ldw8(p.ln_b, ln_b);
#pragma unroll
            for (int e = 0; e < 8; e++) {
                y[e] = hr(sm.ys[tk][cg * 8 + e]);   // the WKV output is an fp16 tensor in the reference
                s1 += y[e];
                s2 += y[e] * y[e];
            }
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {
                s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            }
            const float mean = s1 * (1.0f / 64.0f);
            const float var = fmaxf(s2 * (1.0f / 64.0f) - mean * mean, 0.0f);
            const float rstd = 1.0f / sqrtf(var + 64e-5f);
            const float bonus = sm.bonus[tk];
            const int4 qv = *reinterpret_cast<const int4*>(&sm.v2s[tk][cg * 8]);
            const int4 qg = *reinterpret_cast<const int4*>(&sm.gs[tk][cg * 8]);
            const __half *vh = reinterpret_cast<const __half*>(&qv), *gh = reinterpret_cast<const __half*>(&qg);
            __half oh[8];
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const float gn = hr((y[e] - mean) * rstd * ln_w[e] + ln_b[e]);
                const float o = hr(gn + hr(bonus * __half2float(vh[e])));
                oh[e] = __float2half_rn(hr(o * __half2float(gh[e])));
            }
            if (live) *reinterpret_cast<int4*>(p.out + ((size_t)bb * T + t) * C + ch0) = *reinterpret_cast<const int4*>(oh);
        }
        // the next chunk's prep writes L / vT / v2s / gs / bonus: every reader of those is past the barrier above except
        // this output stage, which reads ys / v2s / gs / bonus
        __syncthreads();
    }

    // ---- new state: one fp16 rounding, back through the swizzled tile, one TMA store
    {
        const int r_lo = 16 * warp + g, r_hi = r_lo + 8;
#pragma unroll
        for (int n = 0; n < 8; n++) {
            *reinterpret_cast<__half2*>(sm.tile + r_lo * 128 + ((n ^ (r_lo & 7)) << 4) + 4 * tq) = __floats2half2_rn(S[n][0], S[n][1]);
            *reinterpret_cast<__half2*>(sm.tile + r_hi * 128 + ((n ^ (r_hi & 7)) << 4) + 4 * tq) = __floats2half2_rn(S[n][2], S[n][3]);
        }
    }
    fence_proxy_async();
    __syncthreads();
    if (tid == 0) {
        tma_store_2d(&tm_state, 0, row0, sm.tile);
        tma_store_commit();
        tma_store_wait_read<0>();
    }
}

}  // namespace

int launch_wkv7_chunk(const WkvFusedArgs& a, void* state, long state_slot_stride, int n_slots, cudaStream_t stream, bool pdl) {
    if (a.B <= 0 || a.T <= 0) return 0;
    if (a.H * HS != a.C) return B200_ERR_BAD_ARG;
    CUtensorMap tm;
    int rc = wkv7_make_state_tmap(&tm, state, a.slot_idx ? n_slots : a.B, a.H, state_slot_stride);
    if (rc) return rc;
    WkvFusedArgs p = a;
    p.rows_per_slot = (int)(state_slot_stride / HS);
    const size_t smem = sizeof(ChunkSmem) + 1024;
    static bool attr_done[64] = {};   // per device: worker threads of one process drive several GPUs
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !attr_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(wkv7_chunk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        if (dev >= 0 && dev < 64) attr_done[dev] = true;
    }
    return launch_kernel(wkv7_chunk_kernel, dim3(a.B * a.H), dim3(NT), smem, stream, pdl, 0, tm, p);
}

}  // namespace b200

extern "C" int b200_tmix_core_chunked(int B, int T, int C, int H, void* state, long state_slot_stride, int n_slots, const void* r,
                                      const void* w, const void* k, const void* v, const void* a, const void* g,
                                      const void* vgate, const void* v_first, const void* k_k, const void* k_a, const void* r_k,
                                      const void* ln_w, const void* ln_b, void* out, const int* elapsed_t, const int* slot_idx,
                                      void* stream) {
    using namespace b200;
    if (B < 0 || T < 0 || H <= 0 || H * 64 != C) return B200_ERR_BAD_ARG;
    if ((v_first == nullptr) != (vgate == nullptr)) return B200_ERR_BAD_ARG;
    auto h = [](const void* p) { return static_cast<const __half*>(p); };
    WkvFusedArgs fa;
    fa.B = B; fa.T = T; fa.C = C; fa.H = H; fa.rows_per_slot = 0;
    fa.r = h(r); fa.w = h(w); fa.k = h(k); fa.v = h(v); fa.a = h(a); fa.g = h(g); fa.vgate = h(vgate); fa.v_first = h(v_first);
    fa.k_k = h(k_k); fa.k_a = h(k_a); fa.r_k = h(r_k); fa.ln_w = h(ln_w); fa.ln_b = h(ln_b);
    fa.out = static_cast<__half*>(out); fa.elapsed_t = elapsed_t; fa.slot_idx = slot_idx;
    return launch_wkv7_chunk(fa, state, state_slot_stride, n_slots, static_cast<cudaStream_t>(stream));
}

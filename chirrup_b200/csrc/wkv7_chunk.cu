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
constexpr int LDY = 68;       // floats per row of the Y staging tile: 4 t + i -> the fragment store (t = 2tq, i = g) hits 32 banks

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
// four 8x8 b16 matrices, TRANSPOSED on the way in: lane l supplies the address of row (l & 7) of matrix (l >> 3); thread
// (g, tq) receives elements (row 2tq, 2tq+1; column g) of each — the mma fragment of an operand stored k-major
DEVI void ldmatrix_x4_trans(uint32_t (&r)[4], const __half* row_ptr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(smem_u32(row_ptr)));
}

// D = A (16x16, row) * B (16x8, col) + D, fp16 operands, fp32 accumulate
DEVI void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// 39 KB per CTA: five CTAs (20 warps) per SM, which also turns the 2 048 (sequence, head) tiles of a bsz-32 launch into three
// rounds of 740 instead of four rounds of 592.  Two regions are time-shared: the TMA state tile is only live before the first
// and after the last chunk, the input staging slots only in between; the cumulative decays (L) are dead once the operands
// are rescaled, the chunk's y tile is born after that.
struct ChunkSmem {
    union alignas(1024) {
        uint8_t tile[HS * HS * 2];             // state tile, TMA SWIZZLE_128B layout (launch prologue / epilogue)
        int4 stage[8][NT];                     // next chunk's r,w,k,v,a,g,vgate,v_first rows: one private 16-byte slot per thread
    };
    union {
        float L[TC][HS];                       // log2 d, then its inclusive cumulative sum over the chunk's tokens (stages 1-3)
        float ys[TC][LDY];                     // y of the chunk [t][i] (stages 4-5)
    };
    float em[HS], ecTm[HS];                    // 2^m, 2^(c_16 - m) per channel
    alignas(16) __half opT[2 * TC][LDO];       // rows 0..15 a~_t, rows 16..31 r~_t           [t][j]
    alignas(16) __half opS[2 * TC][LDO];       // rows 0..15 b~_s, rows 16..31 k~_s           [s][j]
    alignas(16) __half aakT[TC][LDS16];        // Aak, Abr, Akr, (I - Aab)^-1 as B operands   [t][s]
    alignas(16) __half abrT[TC][LDS16];
    alignas(16) __half akrT[TC][LDS16];
    alignas(16) __half tinvT[TC][LDS16];
    float aab[TC][TC + 1];                     // Aab in fp32 for the triangular inverse      [s][t]
    alignas(16) __half v2s[TC][LDO];           // v' per token: A operand (via ldmatrix.trans) and output stage   [s][i]
    alignas(16) __half gs[TC][HS];
    float bonus[TC];
    alignas(8) uint64_t bar;
};

__global__ void __launch_bounds__(NT, 5) wkv7_chunk_kernel(const __grid_constant__ CUtensorMap tm_state, const WkvFusedArgs p) {
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

    // Each thread stages exactly the 8 x 16 bytes it will consume itself (token tk, channels 8cg..8cg+7 of r,w,k,v,a,g,vgate,
    // v_first), so the prefetch needs no barrier: program order + cp.async.wait_group.  Tokens past T re-read token 0 (valid
    // memory) and are masked below.
    auto prefetch = [&](int t_first) {
        const int tt = t_first + tk;
        const size_t idx = ((size_t)bb * T + (tt < T ? tt : 0)) * C + ch0;
        cp_async16(&sm.stage[0][tid], p.r + idx);
        cp_async16(&sm.stage[1][tid], p.w + idx);
        cp_async16(&sm.stage[2][tid], p.k + idx);
        cp_async16(&sm.stage[3][tid], p.v + idx);
        cp_async16(&sm.stage[4][tid], p.a + idx);
        cp_async16(&sm.stage[5][tid], p.g + idx);
        if (p.v_first) {
            cp_async16(&sm.stage[6][tid], p.vgate + idx);
            cp_async16(&sm.stage[7][tid], p.v_first + idx);
        }
        cp_async_commit();
    };
    __syncthreads();   // every thread has taken its part of the state tile: the region becomes the staging slots
    prefetch(0);

    for (int t0 = 0; t0 < T; t0 += TC) {
        // ================================================================ 1. vector prep of 16 tokens (rwkv7.py:632-637)
        // Every reference op here acts on fp16 tensors (fp32 inside the op, one rounding): exactly one correctly rounded
        // half op (sums and products of two binary16 values are exact in fp32), so the prep runs on half2 pairs.
        const int t = t0 + tk;
        const bool live = t < T;
        __half2 a_vec[4], b_vec[4], k2[4], rr[4];
        float ld[8];
        {
            auto ldc = [&](const __half* base) { return *reinterpret_cast<const int4*>(base + ch0); };
            cp_async_wait_all();
            const int4 q_r = sm.stage[0][tid], q_w = sm.stage[1][tid], q_k = sm.stage[2][tid], q_v = sm.stage[3][tid];
            const int4 q_a = sm.stage[4][tid], q_g = sm.stage[5][tid];
            int4 q_vg = q_v, q_vf = q_v;
            if (p.v_first) {
                q_vg = sm.stage[6][tid];
                q_vf = sm.stage[7][tid];
            }
            if (t0 + TC < T) prefetch(t0 + TC);   // this thread's slots are free again: the next chunk's rows start moving now
            const int4 w_kk = ldc(p.k_k), w_ka = ldc(p.k_a), w_rk = ldc(p.r_k);
            const __half2 *hr_ = reinterpret_cast<const __half2*>(&q_r), *hk = reinterpret_cast<const __half2*>(&q_k);
            const __half2 *hv = reinterpret_cast<const __half2*>(&q_v), *ha = reinterpret_cast<const __half2*>(&q_a);
            const __half2 *hvg = reinterpret_cast<const __half2*>(&q_vg), *hvf = reinterpret_cast<const __half2*>(&q_vf);
            const __half2 *kkw = reinterpret_cast<const __half2*>(&w_kk), *kaw = reinterpret_cast<const __half2*>(&w_ka);
            const __half2* rkw = reinterpret_cast<const __half2*>(&w_rk);
            const __half* hw = reinterpret_cast<const __half*>(&q_w);
            const __half2 one = __floats2half2_rn(1.f, 1.f);
            __half2 q[4], v2[4];
            float ssq = 0.f, sbon = 0.f;
#pragma unroll
            for (int u = 0; u < 4; u++) {
                q[u] = __hmul2_rn(hk[u], kkw[u]);                                                           // k * k_k
                k2[u] = __hmul2_rn(hk[u], __hadd2_rn(one, __hmul2_rn(__hsub2_rn(ha[u], one), kaw[u])));     // k * (1 + (a-1) * k_a)
                v2[u] = p.v_first ? __hadd2_rn(hv[u], __hmul2_rn(__hsub2_rn(hvf[u], hv[u]), hvg[u])) : hv[u];
                const float2 qf = __half22float2(q[u]);
                ssq = fmaf(qf.x, qf.x, fmaf(qf.y, qf.y, ssq));
                const float2 bf = __half22float2(__hmul2_rn(__hmul2_rn(hr_[u], k2[u]), rkw[u]));            // r * k * r_k
                sbon += bf.x + bf.y;
                rr[u] = hr_[u];
            }
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {   // the 8 threads of a token are 8 consecutive lanes
                ssq += __shfl_xor_sync(0xffffffffu, ssq, o);
                sbon += __shfl_xor_sync(0xffffffffu, sbon, o);
            }
            const float rnrm = 1.0f / hr(sqrtf(ssq));
            const __half2 zero2 = __floats2half2_rn(0.f, 0.f);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float2 qf = __half22float2(q[u]);
                const __half2 kk = __floats2half2_rn(qf.x * rnrm, qf.y * rnrm);                            // normalize(k * k_k)
                a_vec[u] = __hneg2(kk);
                b_vec[u] = __hmul2_rn(kk, ha[u]);
                ld[2 * u] = __log2f(1.0f + __half2float(decay_dither_c(hw[2 * u], t_base + t)));
                ld[2 * u + 1] = __log2f(1.0f + __half2float(decay_dither_c(hw[2 * u + 1], t_base + t)));
                if (!live) {   // past the end of the sequence: an identity step (d = 1, no update, no read-out)
                    a_vec[u] = b_vec[u] = k2[u] = v2[u] = rr[u] = zero2;
                    ld[2 * u] = ld[2 * u + 1] = 0.f;
                }
            }
            // L row layout: position 32 j + 4 cg + e holds channel 8 cg + 4 j + e (a quarter-warp's 16-byte accesses are contiguous)
            *reinterpret_cast<float4*>(&sm.L[tk][cg * 4]) = make_float4(ld[0], ld[1], ld[2], ld[3]);
            *reinterpret_cast<float4*>(&sm.L[tk][32 + cg * 4]) = make_float4(ld[4], ld[5], ld[6], ld[7]);
            *reinterpret_cast<int4*>(&sm.v2s[tk][cg * 8]) = *reinterpret_cast<const int4*>(v2);
            *reinterpret_cast<int4*>(&sm.gs[tk][cg * 8]) = q_g;
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
            const int chn = ((tid & 31) >> 2) * 8 + (tid >> 5) * 4 + (tid & 3);   // the channel at position tid of an L row
            sm.em[chn] = ex2_approx_ftz(m);
            sm.ecTm[chn] = ex2_approx_ftz(c - m);
        }
        __syncthreads();
        // ---- rescaled operands (|c - m| <= 8 * 0.875: plain ex2.approx, no range handling needed)
        {
            const float4 c0 = *reinterpret_cast<const float4*>(&sm.L[tk][cg * 4]), c1 = *reinterpret_cast<const float4*>(&sm.L[tk][32 + cg * 4]);
            const float4 m0 = *reinterpret_cast<const float4*>(&sm.L[TC / 2 - 1][cg * 4]);
            const float4 m1 = *reinterpret_cast<const float4*>(&sm.L[TC / 2 - 1][32 + cg * 4]);
            const float cc[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
            const float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
            __half2 at[4], rt[4], bt[4], kt[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                float up[2], dn[2], pv[2];
#pragma unroll
                for (int h2 = 0; h2 < 2; h2++) {
                    const float d = cc[2 * u + h2] - mm[2 * u + h2];
                    up[h2] = ex2_approx_ftz(d);
                    dn[h2] = ex2_approx_ftz(-d);
                    pv[h2] = ex2_approx_ftz(d - ld[2 * u + h2]);
                }
                const float2 af = __half22float2(a_vec[u]), rf = __half22float2(rr[u]);
                const float2 bf = __half22float2(b_vec[u]), kf = __half22float2(k2[u]);
                at[u] = __floats2half2_rn(af.x * pv[0], af.y * pv[1]);
                rt[u] = __floats2half2_rn(rf.x * up[0], rf.y * up[1]);
                bt[u] = __floats2half2_rn(bf.x * dn[0], bf.y * dn[1]);
                kt[u] = __floats2half2_rn(kf.x * dn[0], kf.y * dn[1]);
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
            uint32_t va[4];   // V as A operand (rows i, k = s) out of the token-major v' tile: transposed 8x8 loads
            const int i_lo = 16 * warp + g;
            // matrices: (i 0-7, s 0-7), (i 8-15, s 0-7), (i 0-7, s 8-15), (i 8-15, s 8-15) of this warp's 16 rows
            ldmatrix_x4_trans(va, &sm.v2s[(lane & 7) + 8 * (lane >> 4)][16 * warp + 8 * ((lane >> 3) & 1)]);
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
                uint32_t bq[4];   // B operand (k = s', n = j) out of the [s'][j] tile: rows s' = lane, columns 8n..8n+7
                ldmatrix_x4_trans(bq, &sm.opS[lane][8 * n]);
                mma16816(S[n], ua, bq[0], bq[1]);   // u-tokens x b~
                mma16816(S[n], va, bq[2], bq[3]);   // v-tokens x k~
                const float2 e = *reinterpret_cast<const float2*>(&sm.ecTm[8 * n + 2 * tq]);
                S[n][0] *= e.x; S[n][1] *= e.y; S[n][2] *= e.x; S[n][3] *= e.y;
            }
        }
        __syncthreads();

        // ================================================================ 5. GroupNorm (eps 64e-5) + bonus + gate (rwkv7.py:647-649)
        {
            float y[8], s1 = 0.f, s2 = 0.f;
#pragma unroll
            const float4 y0 = *reinterpret_cast<const float4*>(&sm.ys[tk][cg * 8]), y1 = *reinterpret_cast<const float4*>(&sm.ys[tk][cg * 8 + 4]);
            const float yr[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
#pragma unroll
            for (int e = 0; e < 8; e++) {
                y[e] = hr(yr[e]);   // the WKV output is an fp16 tensor in the reference
                s1 += y[e];
                s2 = fmaf(y[e], y[e], s2);
            }
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {
                s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            }
            const float mean = s1 * (1.0f / 64.0f);
            const float var = fmaxf(s2 * (1.0f / 64.0f) - mean * mean, 0.0f);
            const float rstd = 1.0f / sqrtf(var + 64e-5f);
            const __half2 bonus = __float2half2_rn(sm.bonus[tk]);
            const int4 qv = *reinterpret_cast<const int4*>(&sm.v2s[tk][cg * 8]);
            const int4 qg = *reinterpret_cast<const int4*>(&sm.gs[tk][cg * 8]);
            const int4 qlw = *reinterpret_cast<const int4*>(p.ln_w + ch0), qlb = *reinterpret_cast<const int4*>(p.ln_b + ch0);
            const __half2 *vh = reinterpret_cast<const __half2*>(&qv), *gh = reinterpret_cast<const __half2*>(&qg);
            const __half2 *lw = reinterpret_cast<const __half2*>(&qlw), *lb = reinterpret_cast<const __half2*>(&qlb);
            __half2 oh[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float2 w2 = __half22float2(lw[u]), b2 = __half22float2(lb[u]);
                const __half2 gn = __floats2half2_rn((y[2 * u] - mean) * rstd * w2.x + b2.x, (y[2 * u + 1] - mean) * rstd * w2.y + b2.y);
                oh[u] = __hmul2_rn(__hadd2_rn(gn, __hmul2_rn(bonus, vh[u])), gh[u]);
            }
            if (live) *reinterpret_cast<int4*>(p.out + ((size_t)bb * T + t) * C + ch0) = *reinterpret_cast<const int4*>(oh);
        }
        // the next chunk's prep writes L / v2s / gs / bonus: every reader of those is past the barrier above except
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

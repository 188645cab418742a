// Pieces shared by the two tcgen05 GEMM kernels (gemm.cu: one CTA per tile; gemm_pair.cu: cta_group::2 CTA pairs):
// the fused epilogues (each reproduces the reference's per-op fp16 rounding, Albatross/rwkv7.py:626-630, :637, :649, :678-679)
// and the per-problem argument block.
#pragma once
#include "common.cuh"
#include "../../include/chirrup_b200.h"

namespace b200 {

constexpr int BK = 64;      // K elements per pipeline stage = one 128-byte swizzle atom row
constexpr int STG_LD = 36;  // row pitch (floats) of staged 32-column units: 16-byte aligned rows, conflict-free 128-bit accesses

DEVI float round_h(float x) { return __half2float(__float2half_rn(x)); }

template <int EPI>
DEVI float epilogue_op(float acc, float bias, float res) {
    if (EPI == B200_EPI_NONE) return acc + bias;
    if (EPI == B200_EPI_TANH) return tanhf(round_h(acc + bias));
    if (EPI == B200_EPI_SIGMOID)  // 1 / (1 + 2^(-x log2 e)) on the MUFU units: error ~1e-6 relative, far below the fp16 rounding
        return rcp_approx_ftz(1.0f + ex2_approx_ftz(-1.4426950408889634f * round_h(acc + bias)));
    if (EPI == B200_EPI_RELU_SQ) {
        float t = fmaxf(round_h(acc), 0.0f);
        return t * t;
    }
    return res + round_h(acc);  // B200_EPI_RESADD
}

// Four consecutive output columns of one row: fp32 -> epilogue -> fp16, one 8-byte store.  In the epilogue's coalesced
// layout a warp instruction covers 4 rows x 64 contiguous bytes.
template <int EPI>
DEVI void finish4(const float4 acc, __half* __restrict__ out_row, const __half* __restrict__ bias,
                  const __half* __restrict__ res_row, int c) {
    float a[4] = {acc.x, acc.y, acc.z, acc.w};
    float bv[4] = {0.f, 0.f, 0.f, 0.f}, rv[4] = {0.f, 0.f, 0.f, 0.f};
    if ((EPI == B200_EPI_NONE || EPI == B200_EPI_TANH || EPI == B200_EPI_SIGMOID) && bias != nullptr) {
        uint2 q = *reinterpret_cast<const uint2*>(bias + c);
        const __half* p = reinterpret_cast<const __half*>(&q);
#pragma unroll
        for (int e = 0; e < 4; e++) bv[e] = __half2float(p[e]);
    }
    if (EPI == B200_EPI_RESADD) {
        uint2 q = *reinterpret_cast<const uint2*>(res_row + c);
        const __half* p = reinterpret_cast<const __half*>(&q);
#pragma unroll
        for (int e = 0; e < 4; e++) rv[e] = __half2float(p[e]);
    }
    uint2 o;
    __half* po = reinterpret_cast<__half*>(&o);
#pragma unroll
    for (int e = 0; e < 4; e++) po[e] = __float2half_rn(epilogue_op<EPI>(a[e], bv[e], rv[e]));
    *reinterpret_cast<uint2*>(out_row + c) = o;
}

DEVI void finish4_rt(int epi, const float4 acc, __half* out_row, const __half* bias, const __half* res_row, int c) {
    switch (epi) {
        case B200_EPI_NONE: finish4<B200_EPI_NONE>(acc, out_row, bias, res_row, c); break;
        case B200_EPI_TANH: finish4<B200_EPI_TANH>(acc, out_row, bias, res_row, c); break;
        case B200_EPI_SIGMOID: finish4<B200_EPI_SIGMOID>(acc, out_row, bias, res_row, c); break;
        case B200_EPI_RELU_SQ: finish4<B200_EPI_RELU_SQ>(acc, out_row, bias, res_row, c); break;
        default: finish4<B200_EPI_RESADD>(acc, out_row, bias, res_row, c); break;
    }
}

struct GemmArgs {
    int M, N, K;
    __half* out;
    long ldo;
    const __half* bias;
    const __half* residual;
    int epi;
    float* partials;  // [tile][split][unit][128][32] fp32
    int kb_per_split;
    int n_split;      // == cluster size along z (gridDim.z with cluster z = 1 when `defer` is set)
    float* defer;     // non-null: K split whose reduction is DEFERRED to the consumer kernel — split z stores its fp32 partial
                      // tile row-major at defer[(z * M + row) * N + col]; no bias / residual / rounding here (gemm_pair.cu only)
    unsigned long long* dbg;  // optional timeline probe (globaltimer ns), 8 slots per CTA; nullptr in production
};

// ---- epilogue helpers.  Everything the inner loops need is hoisted into registers (EpiCtx) and the epilogue kind is a template
// parameter chosen once per CTA: the first version re-read the problem block with indexed constant loads and ran a 5-way switch
// per 4 outputs (~70 instructions per 4 outputs, ncu source page profiles/r01_pair_ffnkey_ncu_v5.txt), which made the
// epilogue as long as 40 % of the main loop.
struct EpiCtx {
    __half* out;
    const __half* bias;
    const __half* residual;
    long ldo;
    int M, N;
};
DEVI void sts128(uint32_t addr, float x, float y, float z, float w) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
DEVI float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
// Eight consecutive outputs of one row: fp32 -> epilogue -> 8 x fp16 (one 16-byte store for the caller).
template <int EPI>
DEVI uint4 finish8(const float4 lo, const float4 hi, const uint4 b8, const uint4 r8) {
    const float a[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    const __half2* bp = reinterpret_cast<const __half2*>(&b8);
    const __half2* rp = reinterpret_cast<const __half2*>(&r8);
    uint4 o;
    __half2* op = reinterpret_cast<__half2*>(&o);
#pragma unroll
    for (int e = 0; e < 4; e++) {
        const float2 bv = __half22float2(bp[e]), rv = __half22float2(rp[e]);
        op[e] = __floats2half2_rn(epilogue_op<EPI>(a[2 * e], bv.x, rv.x), epilogue_op<EPI>(a[2 * e + 1], bv.y, rv.y));
    }
    return o;
}

// One 32-row x 32-column unit of this warp.  In: thread = row (lane), v = its 32 consecutive columns.  The unit is transposed
// through a padded shared-memory buffer so that a warp store covers 8 rows x 64 contiguous bytes (lane = row i*8 + lane/4,
// 8 columns), bias / residual are fetched as 16-byte vectors BEFORE the staging so their latency hides under it.
template <int EPI>
DEVI void emit_unit(const float (&v)[32], uint32_t stg, const EpiCtx& c, int row0, int col0, int lane) {
    const int sub_r = lane >> 2, c8 = (lane & 3) * 8;
    const int col = col0 + c8;
    const bool col_ok = col < c.N;
    uint4 b8 = make_uint4(0u, 0u, 0u, 0u);
    uint4 r8[4];
    if ((EPI == B200_EPI_NONE || EPI == B200_EPI_TANH || EPI == B200_EPI_SIGMOID) && c.bias != nullptr && col_ok)
        b8 = *reinterpret_cast<const uint4*>(c.bias + col);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        r8[i] = make_uint4(0u, 0u, 0u, 0u);
        if (EPI == B200_EPI_RESADD) {
            const int row = row0 + i * 8 + sub_r;
            if (row < c.M && col_ok) r8[i] = *reinterpret_cast<const uint4*>(c.residual + (size_t)row * c.ldo + col);
        }
    }
#pragma unroll
    for (int e = 0; e < 8; e++) sts128(stg + (uint32_t)(lane * STG_LD + 4 * e) * 4u, v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]);
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int r = i * 8 + sub_r;
        const uint32_t a = stg + (uint32_t)(r * STG_LD + c8) * 4u;
        const float4 lo = lds128(a), hi = lds128(a + 16u);
        const int row = row0 + r;
        if (row < c.M && col_ok) *reinterpret_cast<uint4*>(c.out + (size_t)row * c.ldo + col) = finish8<EPI>(lo, hi, b8, r8[i]);
    }
}

// The same unit stored as raw fp32 (deferred K-split partial): a warp store covers 8 rows x 128 contiguous bytes.
DEVI void emit_unit_f32(const float (&v)[32], uint32_t stg, float* __restrict__ dst, long ld, int M, int N, int row0, int col0,
                        int lane) {
    const int sub_r = lane >> 2, c8 = (lane & 3) * 8;
    const int col = col0 + c8;
#pragma unroll
    for (int e = 0; e < 8; e++) sts128(stg + (uint32_t)(lane * STG_LD + 4 * e) * 4u, v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]);
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int r = i * 8 + sub_r;
        const uint32_t a = stg + (uint32_t)(r * STG_LD + c8) * 4u;
        const float4 lo = lds128(a), hi = lds128(a + 16u);
        const int row = row0 + r;
        if (row < M && col < N) {
            float4* q = reinterpret_cast<float4*>(dst + (size_t)row * ld + col);
            __stcg(q, lo);
            __stcg(q + 1, hi);
        }
    }
}

#define EPI_DISPATCH(epi, CALL)                                       \
    switch (epi) {                                                    \
        case B200_EPI_NONE: { constexpr int E = B200_EPI_NONE; CALL; } break;       \
        case B200_EPI_TANH: { constexpr int E = B200_EPI_TANH; CALL; } break;       \
        case B200_EPI_SIGMOID: { constexpr int E = B200_EPI_SIGMOID; CALL; } break; \
        case B200_EPI_RELU_SQ: { constexpr int E = B200_EPI_RELU_SQ; CALL; } break; \
        default: { constexpr int E = B200_EPI_RESADD; CALL; } break;                \
    }

}  // namespace b200

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
    int n_split;      // == cluster size along z
    unsigned long long* dbg;  // optional timeline probe (globaltimer ns), 8 slots per CTA; nullptr in production
};

}  // namespace b200

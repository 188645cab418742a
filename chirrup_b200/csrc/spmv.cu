// Sparse vector x dense matrix for the bsz-1 FFN value product:  out[c] = sum_{d : vec[d] != 0} vec[d] * mat[d, c].
//
// Replaces spvecmatmul_noindices (/root/reference/Albatross/cuda/rwkv7_state_fwd_fp16.cu:222-310, launcher :324-330) and the
// Triton rwkv_mm_sparsity (/root/reference/Albatross/rwkv_mm_op_triton.py:6-61).  Unlike the reference (fp16 atomics into a
// pre-zeroed output, arbitrary order) this accumulates in fp32 in a fixed order and overwrites `out`: deterministic.
//
// CTA = 64 output columns x 32 row lanes (256 threads, 8 columns = one 16-byte load per thread per row); rows whose vec entry
// is zero are never fetched; the 32 row-lane partials are combined through shared memory in lane order.
#include "common.cuh"
#include "../../include/chirrup_b200.h"

namespace b200 {
namespace {

constexpr int SP_COLS = 64;
constexpr int SP_LANES = 32;

__global__ void __launch_bounds__(256) spmv_kernel(int D, int C, const __half* __restrict__ vec,
                                                   const __half* __restrict__ mat, __half* __restrict__ out) {
    __shared__ float part[SP_LANES][SP_COLS + 4];
    const int cg = threadIdx.x & 7;    // column group of 8
    const int rl = threadIdx.x >> 3;   // row lane 0..31
    const int c0 = blockIdx.x * SP_COLS + cg * 8;
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; e++) acc[e] = 0.f;
    if (c0 < C) {
        for (int d = rl; d < D; d += SP_LANES) {
            const __half x = vec[d];
            if ((__half_as_ushort(x) & 0x7fff) == 0) continue;
            const float xf = __half2float(x);
            int4 q = __ldg(reinterpret_cast<const int4*>(mat + (size_t)d * C + c0));
            const __half* p = reinterpret_cast<const __half*>(&q);
#pragma unroll
            for (int e = 0; e < 8; e++) acc[e] = fmaf(xf, __half2float(p[e]), acc[e]);
        }
    }
#pragma unroll
    for (int e = 0; e < 8; e++) part[rl][cg * 8 + e] = acc[e];
    __syncthreads();
    if (threadIdx.x < SP_COLS) {
        const int c = blockIdx.x * SP_COLS + threadIdx.x;
        if (c < C) {
            float s = 0.f;
#pragma unroll
            for (int l = 0; l < SP_LANES; l++) s += part[l][threadIdx.x];
            out[c] = __float2half_rn(s);
        }
    }
}

}  // namespace
}  // namespace b200

extern "C" int b200_spmv_forward(int D, int C, const void* vec, const void* mat, void* out, void* stream) {
    if (D <= 0 || C <= 0 || (C & 7)) return B200_ERR_BAD_ARG;
    const int grid = (C + b200::SP_COLS - 1) / b200::SP_COLS;
    b200::spmv_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        D, C, static_cast<const __half*>(vec), static_cast<const __half*>(mat), static_cast<__half*>(out));
    return (int)cudaGetLastError();
}

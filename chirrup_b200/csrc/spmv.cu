// Sparse vector x dense matrix for the bsz-1 FFN value product:  out[c] = sum_{d : vec[d] != 0} vec[d] * mat[d, c].
//
// Replaces spvecmatmul_noindices (/root/reference/Albatross/cuda/rwkv7_state_fwd_fp16.cu:222-310, launcher :324-330) and the
// Triton rwkv_mm_sparsity (/root/reference/Albatross/rwkv_mm_op_triton.py:6-61).  The reference builds each 64-row slab's
// non-zero list on ONE thread, streams rows 4 bytes per thread, accumulates in fp16 and merges slabs with fp16 atomicAdd into a
// pre-zeroed output (run-to-run different bits).  Here:
//   * grid = (C / 64 column blocks) x (8 row ranges) = 512 CTAs at C = 4096, the 8 CTAs of a column block forming one
//     thread-block cluster;
//   * every CTA compacts the non-zero rows of its range with warp ballots (ordered, all 256 threads), then 32 row lanes x 8
//     column groups stream ONLY those rows with 16-byte loads, four in flight per thread — the weight-bandwidth-bound part;
//   * fp32 accumulation; the 32 lane partials are summed in lane order, the 8 range partials in rank order by the cluster's
//     rank-0 CTA through distributed shared memory: one rounding, a fixed order, no atomics, `out` overwritten (no pre-zeroing);
//   * optional fused residual: out[c] = fp16(res[c] + fp16(sum)) — the `x + rwkv_mm_sparsity(k, V_)` of rwkv7.py:659, :536.
#include <cooperative_groups.h>

#include "common.cuh"
#include "glue.h"
#include "../../include/chirrup_b200.h"

namespace cg = cooperative_groups;

namespace b200 {
namespace {

constexpr int SP_COLS = 64;     // output columns per CTA
constexpr int SP_LANES = 32;    // row lanes per CTA
constexpr int SP_SPLIT = 8;     // row ranges = cluster size
constexpr int SP_MAXR = 4096;   // rows per CTA range handled per pass (D / SP_SPLIT <= SP_MAXR per pass; larger D loops)

__global__ void __cluster_dims__(1, 1, SP_SPLIT) __launch_bounds__(256)
spmv_kernel(int D, int C, const __half* __restrict__ vec, const __half* __restrict__ mat, const __half* __restrict__ res,
            __half* __restrict__ out) {
    __shared__ float part[SP_LANES][SP_COLS + 4];
    __shared__ float psum[SP_COLS];
    __shared__ unsigned short nz_idx[SP_MAXR];
    __shared__ float nz_val[SP_MAXR];
    __shared__ int warp_cnt[8];
    cg::cluster_group cluster = cg::this_cluster();
    pdl_launch_dependents();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int cgp = tid & 7;    // column group of 8
    const int rl = tid >> 3;    // row lane 0..31
    const int c0 = blockIdx.x * SP_COLS + cgp * 8;
    const int z = blockIdx.z;
    const int rows_per = (D + SP_SPLIT - 1) / SP_SPLIT;
    const int r_begin = z * rows_per, r_end = min(D, r_begin + rows_per);
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; e++) acc[e] = 0.f;
    pdl_wait();
    for (int base = r_begin; base < r_end; base += SP_MAXR) {
        const int lim = min(r_end, base + SP_MAXR);
        // ---- ordered compaction of the non-zero rows of [base, lim): 256 rows per round, ballot + prefix
        int count = 0;
        for (int r0 = base; r0 < lim; r0 += 256) {
            const int d = r0 + tid;
            __half x = __float2half(0.f);
            if (d < lim) x = vec[d];
            const bool nzf = (__half_as_ushort(x) & 0x7fff) != 0;
            const unsigned m = __ballot_sync(0xffffffffu, nzf);
            if (lane == 0) warp_cnt[warp] = __popc(m);
            __syncthreads();
            int off = count;
#pragma unroll
            for (int w = 0; w < 8; w++) {
                if (w < warp) off += warp_cnt[w];
                count += warp_cnt[w];
            }
            if (nzf) {
                const int pos = off + __popc(m & ((1u << lane) - 1u));
                nz_idx[pos] = (unsigned short)(d - base);
                nz_val[pos] = __half2float(x);
            }
            __syncthreads();
        }
        // ---- stream the non-zero rows: lane rl takes entries rl, rl + 32, ...; four 16-byte loads in flight per thread
        if (c0 < C) {
            const __half* mp = mat + (size_t)base * C + c0;
            int i = rl;
            for (; i + 3 * SP_LANES < count; i += 4 * SP_LANES) {
                int4 q[4];
                float xv[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    q[u] = __ldcs(reinterpret_cast<const int4*>(mp + (size_t)nz_idx[i + u * SP_LANES] * C));
                    xv[u] = nz_val[i + u * SP_LANES];
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const __half* ph = reinterpret_cast<const __half*>(&q[u]);
#pragma unroll
                    for (int e = 0; e < 8; e++) acc[e] = fmaf(xv[u], __half2float(ph[e]), acc[e]);
                }
            }
            for (; i < count; i += SP_LANES) {
                const int4 q = __ldcs(reinterpret_cast<const int4*>(mp + (size_t)nz_idx[i] * C));
                const float xv = nz_val[i];
                const __half* ph = reinterpret_cast<const __half*>(&q);
#pragma unroll
                for (int e = 0; e < 8; e++) acc[e] = fmaf(xv, __half2float(ph[e]), acc[e]);
            }
        }
        __syncthreads();   // the index list is rebuilt by the next pass
    }
    // ---- lane partials -> this CTA's partial (lane order), range partials -> the result (rank order, rank 0 of the cluster)
#pragma unroll
    for (int e = 0; e < 8; e++) part[rl][cgp * 8 + e] = acc[e];
    __syncthreads();
    if (tid < SP_COLS) {
        float s = 0.f;
#pragma unroll
        for (int l = 0; l < SP_LANES; l++) s += part[l][tid];
        psum[tid] = s;
    }
    cluster.sync();
    if (cluster.block_rank() == 0 && tid < SP_COLS) {
        const int c = blockIdx.x * SP_COLS + tid;
        if (c < C) {
            float s = 0.f;
#pragma unroll
            for (int r = 0; r < SP_SPLIT; r++) s += cluster.map_shared_rank(psum, r)[tid];
            __half o = __float2half_rn(s);
            if (res) o = __float2half_rn(__half2float(res[c]) + __half2float(o));
            out[c] = o;
        }
    }
    cluster.sync();   // peers' shared memory must outlive rank 0's reads
}

}  // namespace

int launch_spmv(int D, int C, const __half* vec, const __half* mat, const __half* res, __half* out, cudaStream_t stream, bool pdl) {
    if (D <= 0 || C <= 0 || (C & 7)) return B200_ERR_BAD_ARG;
    const int grid_x = (C + SP_COLS - 1) / SP_COLS;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid_x, 1, SP_SPLIT);
    cfg.blockDim = dim3(256);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    int na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    return (int)cudaLaunchKernelEx(&cfg, spmv_kernel, D, C, vec, mat, res, out);
}

}  // namespace b200

extern "C" int b200_spmv_forward(int D, int C, const void* vec, const void* mat, void* out, void* stream) {
    return b200::launch_spmv(D, C, static_cast<const __half*>(vec), static_cast<const __half*>(mat), nullptr,
                             static_cast<__half*>(out), static_cast<cudaStream_t>(stream), false);
}

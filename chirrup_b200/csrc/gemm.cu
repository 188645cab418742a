// Dense fp16 linear layer on tcgen05 tensor cores:  out (M,N) = epilogue( A (M,K) . W (N,K)^T ),  fp32 accumulation in TMEM.
//
// Replaces the cuBLAS calls behind F.linear / `@` in the reference model
// (/root/reference/Albatross/rwkv7.py:625-630, :637, :649, :678-679, :551).
//
// One CTA computes a (128*MT) x BN output tile over a K range:
//   warp 0 (one lane)  TMA producer: A and W tiles, 64 K-columns (128 B, SWIZZLE_128B) per stage, mbarrier ring
//   warp 1 (one lane)  tcgen05.mma issuer: MT accumulators of 128 x BN fp32 in TMEM share each W tile
//   warps 2-5          epilogue: tcgen05.ld -> registers -> bias / activation / residual -> fp16 -> global
// BN = 256 for the big weight-streaming shapes (fewest re-reads of the activations through L2), BN = 32 for the skinny LoRA
// shapes (many small tiles so that every SM takes part).
// Decode-sized problems (M <= 256: a handful of tiles) are split along K over the CTAs of one thread-block CLUSTER: every CTA
// parks its fp32 partial tile in an L2-resident workspace, the cluster barrier orders the writes, and then each CTA reduces and
// finishes a 1/S share of the tile, summing the partials in split order (deterministic, no atomics, no spinning).
#include "common.cuh"
#include "glue.h"
#include "tmap.h"
#include "../../include/chirrup_b200.h"

namespace b200 {
namespace {

constexpr int BK = 64;
constexpr int GEMM_THREADS = 192;
constexpr int MAX_SPLIT = 8;                      // portable cluster size
constexpr long WS_BYTES = 160L * 256 * 256 * 4;  // partial-tile capacity: > 148 CTAs x the largest (256 x 256 fp32) tile

template <int MT, int BN>
struct Cfg {
    static constexpr int kABytes = MT * 128 * BK * 2;
    static constexpr int kWBytes = BN * BK * 2;
    static constexpr int kStageBytes = kABytes + kWBytes;
    static constexpr int kStages = (200 * 1024) / kStageBytes > 8 ? 8 : (200 * 1024) / kStageBytes;
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
    static constexpr uint32_t kTmemCols = (MT * BN) < 32 ? 32 : MT * BN;
    static constexpr int kUnits = MT * (BN / 32);  // epilogue work units of 128 rows x 32 columns
};

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

// 32 consecutive output columns of one row: acc[] fp32 -> fp16 global, 16-byte stores.
template <int EPI>
DEVI void store_row32(const float (&acc)[32], __half* __restrict__ out_row, const __half* __restrict__ bias,
                      const __half* __restrict__ res_row, int col0, int N) {
#pragma unroll
    for (int g = 0; g < 4; g++) {
        const int c = col0 + g * 8;
        if (c >= N) break;
        float bv[8], rv[8];
#pragma unroll
        for (int e = 0; e < 8; e++) bv[e] = 0.f, rv[e] = 0.f;
        if ((EPI == B200_EPI_NONE || EPI == B200_EPI_TANH || EPI == B200_EPI_SIGMOID) && bias != nullptr) {
            int4 q = *reinterpret_cast<const int4*>(bias + c);
            const __half* p = reinterpret_cast<const __half*>(&q);
#pragma unroll
            for (int e = 0; e < 8; e++) bv[e] = __half2float(p[e]);
        }
        if (EPI == B200_EPI_RESADD) {
            int4 q = *reinterpret_cast<const int4*>(res_row + c);
            const __half* p = reinterpret_cast<const __half*>(&q);
#pragma unroll
            for (int e = 0; e < 8; e++) rv[e] = __half2float(p[e]);
        }
        int4 o;
        __half* po = reinterpret_cast<__half*>(&o);
#pragma unroll
        for (int e = 0; e < 8; e++) po[e] = __float2half_rn(epilogue_op<EPI>(acc[g * 8 + e], bv[e], rv[e]));
        *reinterpret_cast<int4*>(out_row + c) = o;
    }
}

DEVI void store_row32_rt(int epi, const float (&acc)[32], __half* out_row, const __half* bias, const __half* res_row,
                         int col0, int N) {
    switch (epi) {
        case B200_EPI_NONE: store_row32<B200_EPI_NONE>(acc, out_row, bias, res_row, col0, N); break;
        case B200_EPI_TANH: store_row32<B200_EPI_TANH>(acc, out_row, bias, res_row, col0, N); break;
        case B200_EPI_SIGMOID: store_row32<B200_EPI_SIGMOID>(acc, out_row, bias, res_row, col0, N); break;
        case B200_EPI_RELU_SQ: store_row32<B200_EPI_RELU_SQ>(acc, out_row, bias, res_row, col0, N); break;
        default: store_row32<B200_EPI_RESADD>(acc, out_row, bias, res_row, col0, N); break;
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
};

DEVI void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <int MT, int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
    gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const GemmArgs g) {
    using C = Cfg<MT, BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
    uint64_t* empty_bar = full_bar + C::kStages;
    uint64_t* accum_bar = empty_bar + C::kStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * BN;
    const int m0 = blockIdx.y * (128 * MT);
    const int split = blockIdx.z;
    const int kb_total = (g.K + BK - 1) / BK;
    const int kb0 = split * g.kb_per_split;
    const int kb1 = min(kb_total, kb0 + g.kb_per_split);
    const int n_kb = kb1 - kb0;  // >= 1 by construction of the launch
    const bool splitk = g.n_split > 1;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmW);
        for (int s = 0; s < C::kStages; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(accum_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<C::kTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int q = warp & 3;  // TMEM lane quarter of an epilogue warp
    const int tile_id = blockIdx.y * gridDim.x + blockIdx.x;
    const int lrow = q * 32 + lane;

    if (warp == 0) {
        if (lane == 0) {
            const uint64_t pol_w = policy_evict_first();
            const uint64_t pol_a = policy_evict_last();
            for (int it = 0; it < n_kb; it++) {
                const int s = it % C::kStages;
                const uint32_t ph = (it / C::kStages) & 1;
                mbar_wait(&empty_bar[s], ph ^ 1);
                uint8_t* st = smem + s * C::kStageBytes;
                mbar_arrive_expect_tx(&full_bar[s], C::kStageBytes);
                const int kc = (kb0 + it) * BK;
#pragma unroll
                for (int mt = 0; mt < MT; mt++)
                    tma_load_2d_hint(st + mt * (128 * BK * 2), &tmA, kc, m0 + mt * 128, &full_bar[s], pol_a);
                tma_load_2d_hint(st + C::kABytes, &tmW, kc, n0, &full_bar[s], pol_w);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(128, BN);
            for (int it = 0; it < n_kb; it++) {
                const int s = it % C::kStages;
                const uint32_t ph = (it / C::kStages) & 1;
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                const uint32_t st = smem_u32(smem + s * C::kStageBytes);
                const uint64_t bdesc = umma_desc_sw128(st + C::kABytes);
#pragma unroll
                for (int k = 0; k < BK / 16; k++) {
#pragma unroll
                    for (int mt = 0; mt < MT; mt++) {
                        const uint64_t adesc = umma_desc_sw128(st + mt * (128 * BK * 2));
                        // advancing 16 K-elements = 32 B inside the 128-B swizzle atom = +2 in the 16-byte address field
                        umma_f16_ss(tmem_base + mt * BN, adesc + 2 * k, bdesc + 2 * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
                    }
                }
                umma_commit(&empty_bar[s]);  // frees the smem stage once these MMAs have read it
            }
            umma_commit(accum_bar);  // accumulators complete
        }
        __syncwarp();
    } else {
        // ---- epilogue warps, phase 1: drain TMEM
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        float* my_part = splitk ? g.partials + ((size_t)tile_id * g.n_split + split) * (size_t)C::kUnits * 128 * 32 : nullptr;
#pragma unroll 1
        for (int u = 0; u < C::kUnits; u++) {
            const int mt = u / (BN / 32), c = u - mt * (BN / 32);
            if (n0 + c * 32 >= g.N) continue;  // columns past N (N tail of the last tile)
            uint32_t v[32];
            tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + mt * BN + c * 32, v);
            tmem_ld_wait();
            if (splitk) {
                float4* dst = reinterpret_cast<float4*>(my_part + ((size_t)u * 128 + lrow) * 32);
#pragma unroll
                for (int e = 0; e < 8; e++)
                    __stcg(dst + e, make_float4(__uint_as_float(v[4 * e]), __uint_as_float(v[4 * e + 1]),
                                                __uint_as_float(v[4 * e + 2]), __uint_as_float(v[4 * e + 3])));
            } else {
                const int row = m0 + mt * 128 + lrow;
                if (row < g.M) {
                    float acc[32];
#pragma unroll
                    for (int e = 0; e < 32; e++) acc[e] = __uint_as_float(v[e]);
                    store_row32_rt(g.epi, acc, g.out + (size_t)row * g.ldo, g.bias,
                                   g.residual ? g.residual + (size_t)row * g.ldo : nullptr, n0 + c * 32, g.N);
                }
            }
        }
    }

    if (splitk) {
        // every CTA of the cluster (= every K split of this tile) has parked its partial: release/acquire at cluster scope
        __threadfence();
        cluster_sync_all();
        if (warp >= 2) {
            const float* base = g.partials + (size_t)tile_id * g.n_split * (size_t)C::kUnits * 128 * 32;
#pragma unroll 1
            for (int u = split; u < C::kUnits; u += g.n_split) {  // this CTA's share of the tile
                const int mt = u / (BN / 32), c = u - mt * (BN / 32);
                const int row = m0 + mt * 128 + lrow;
                if (n0 + c * 32 >= g.N || row >= g.M) continue;
                float acc[32];
#pragma unroll
                for (int e = 0; e < 32; e++) acc[e] = 0.f;
#pragma unroll 2
                for (int z = 0; z < g.n_split; z++) {  // fixed order: deterministic sum
                    const float4* src =
                        reinterpret_cast<const float4*>(base + (((size_t)z * C::kUnits + u) * 128 + lrow) * 32);
                    float4 t[8];
#pragma unroll
                    for (int e = 0; e < 8; e++) t[e] = __ldcg(src + e);
#pragma unroll
                    for (int e = 0; e < 8; e++) {
                        acc[4 * e] += t[e].x; acc[4 * e + 1] += t[e].y; acc[4 * e + 2] += t[e].z; acc[4 * e + 3] += t[e].w;
                    }
                }
                store_row32_rt(g.epi, acc, g.out + (size_t)row * g.ldo, g.bias,
                               g.residual ? g.residual + (size_t)row * g.ldo : nullptr, n0 + c * 32, g.N);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<C::kTmemCols>(tmem_base);
    }
}

struct GemmPlan {
    int mt, bn, grid_x, grid_y, n_split, kb_per_split;
};

// Tile shape and K split for a problem.  Decode-sized problems get one wave of <= 148 CTAs.
static GemmPlan plan_gemm(int M, int N, int K) {
    GemmPlan p;
    const int kb = (K + BK - 1) / BK;
    p.mt = (M > 128) ? 2 : 1;
    p.bn = (N >= 1024 && kb >= 16) ? 256 : 32;
    p.grid_x = (N + p.bn - 1) / p.bn;
    p.grid_y = (M + 128 * p.mt - 1) / (128 * p.mt);
    const long tiles = (long)p.grid_x * p.grid_y;
    int s = (int)(148 / tiles);
    if (s > kb / 2) s = kb / 2;  // >= 2 K-blocks per split
    if (s > MAX_SPLIT) s = MAX_SPLIT;
    if (s < 1) s = 1;
    p.kb_per_split = (kb + s - 1) / s;
    p.n_split = (kb + p.kb_per_split - 1) / p.kb_per_split;  // no empty splits
    return p;
}

template <int MT, int BN>
static int launch_cfg(const GemmPlan& p, const CUtensorMap& tmA, const CUtensorMap& tmW, const GemmArgs& g, cudaStream_t stream) {
    using C = Cfg<MT, BN>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(gemm_kernel<MT, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.grid_x, p.grid_y, p.n_split);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = C::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = p.n_split;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return (int)cudaLaunchKernelEx(&cfg, gemm_kernel<MT, BN>, tmA, tmW, g);
}

}  // namespace

int linear_f16_launch(int M, int N, int K, const void* A, long lda, const void* W, long ldw, void* out, long ldo,
                      const void* bias, const void* residual, int epilogue, void* workspace, long workspace_bytes,
                      cudaStream_t stream) {
    if (M < 0 || N <= 0 || K <= 0 || (K & 7) || (N & 7) || (lda & 7) || (ldw & 7) || (ldo & 7)) return B200_ERR_BAD_ARG;
    if (epilogue < B200_EPI_NONE || epilogue > B200_EPI_RESADD) return B200_ERR_BAD_ARG;
    if (epilogue == B200_EPI_RESADD && residual == nullptr) return B200_ERR_BAD_ARG;
    if (M == 0) return 0;
    GemmPlan p = plan_gemm(M, N, K);
    if (p.n_split > 1) {
        const long need = (long)p.grid_x * p.grid_y * p.n_split * (128L * p.mt) * p.bn * 4;
        if (workspace == nullptr || workspace_bytes < need) return B200_ERR_WORKSPACE;
    }
    CUtensorMap tmA, tmW;
    int rc = make_tmap_2d_f16(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 2, 128, BK, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
    if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
    rc = make_tmap_2d_f16(&tmW, W, (uint64_t)N, (uint64_t)K, (uint64_t)ldw * 2, p.bn, BK, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
    if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
    GemmArgs g;
    g.M = M; g.N = N; g.K = K;
    g.out = static_cast<__half*>(out);
    g.ldo = ldo;
    g.bias = static_cast<const __half*>(bias);
    g.residual = static_cast<const __half*>(residual);
    g.epi = epilogue;
    g.partials = static_cast<float*>(workspace);
    g.kb_per_split = p.kb_per_split;
    g.n_split = p.n_split;
    if (p.mt == 2 && p.bn == 256) return launch_cfg<2, 256>(p, tmA, tmW, g, stream);
    if (p.mt == 1 && p.bn == 256) return launch_cfg<1, 256>(p, tmA, tmW, g, stream);
    if (p.mt == 2) return launch_cfg<2, 32>(p, tmA, tmW, g, stream);
    return launch_cfg<1, 32>(p, tmA, tmW, g, stream);
}

}  // namespace b200

extern "C" long b200_linear_workspace_bytes(int /*max_M*/, int /*max_N*/) { return b200::WS_BYTES; }

extern "C" int b200_linear_f16(int M, int N, int K, const void* A, long lda, const void* W, long ldw, void* out, long ldo,
                               const void* bias, const void* residual, int epilogue, void* workspace, long workspace_bytes,
                               void* stream) {
    return b200::linear_f16_launch(M, N, K, A, lda, W, ldw, out, ldo, bias, residual, epilogue, workspace, workspace_bytes,
                                   static_cast<cudaStream_t>(stream));
}

// Dense fp16 linear layer on tcgen05 tensor cores:  out (M,N) = epilogue( A (M,K) . W (N,K)^T ),  fp32 accumulation in TMEM.
//
// Replaces the cuBLAS calls behind F.linear / `@` in the reference model
// (/root/reference/Albatross/rwkv7.py:625-630, :637, :649, :678-679, :551).
//
// One CTA computes a (128*MT) x 128 output tile over a K range:
//   warp 0 (one lane)  TMA producer: A and W tiles, 64 K-columns (128 B, SWIZZLE_128B) per stage, mbarrier ring
//   warp 1 (one lane)  tcgen05.mma issuer: MT accumulators of 128x128 fp32 in TMEM share each W tile
//   warps 2-5          epilogue: tcgen05.ld -> registers -> bias / activation / residual -> fp16 -> global
// Skinny problems (decode: M <= 256, few N tiles) are split along K across CTAs; partial tiles go to an fp32 workspace
// and the last CTA to arrive on the tile's counter sums them in split order (deterministic) and runs the epilogue.
#include "common.cuh"
#include "tmap.h"
#include "../../include/chirrup_b200.h"

namespace b200 {
namespace {

constexpr int BN = 128;
constexpr int BK = 64;
constexpr int TILE_BYTES = 128 * BK * 2;  // one 128-row x 64-col fp16 operand tile = 16 KiB
constexpr int GEMM_THREADS = 192;
constexpr int MAX_SPLIT = 16;
constexpr int WS_TILES = 160;  // partial-tile capacity of the workspace (>= SMs)
constexpr int N_COUNTERS = 4096;

template <int MT>
struct Cfg {
    static constexpr int kStages = (MT == 2) ? 4 : 6;
    static constexpr int kStageBytes = (MT + 1) * TILE_BYTES;
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
    static constexpr uint32_t kTmemCols = MT * BN;
};

DEVI float round_h(float x) { return __half2float(__float2half_rn(x)); }

template <int EPI>
DEVI float epilogue_op(float acc, float bias, float res) {
    if (EPI == B200_EPI_NONE) return acc + bias;
    if (EPI == B200_EPI_TANH) return tanhf(round_h(acc + bias));
    if (EPI == B200_EPI_SIGMOID) return 1.0f / (1.0f + expf(-round_h(acc + bias)));
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

template <int EPI>
DEVI void store_row32_dispatch(int, const float (&acc)[32], __half* out_row, const __half* bias, const __half* res_row,
                               int col0, int N) {
    store_row32<EPI>(acc, out_row, bias, res_row, col0, N);
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
    float* partials;  // [tile][split][128*MT][128] fp32
    int* counters;    // [tile]
    int kb_per_split;
    int n_split;
};

template <int MT>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
    gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const GemmArgs g) {
    using C = Cfg<MT>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
    uint64_t* empty_bar = full_bar + C::kStages;
    uint64_t* accum_bar = empty_bar + C::kStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
    int* last_flag = reinterpret_cast<int*>(tmem_slot + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * BN;
    const int m0 = blockIdx.y * (128 * MT);
    const int split = blockIdx.z;
    const int kb_total = (g.K + BK - 1) / BK;
    const int kb0 = split * g.kb_per_split;
    const int kb1 = min(kb_total, kb0 + g.kb_per_split);
    const int n_kb = kb1 - kb0;  // >= 1 by construction of the launch

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
                for (int mt = 0; mt < MT; mt++) tma_load_2d_hint(st + mt * TILE_BYTES, &tmA, kc, m0 + mt * 128, &full_bar[s], pol_a);
                tma_load_2d_hint(st + MT * TILE_BYTES, &tmW, kc, n0, &full_bar[s], pol_w);
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
                const uint64_t bdesc = umma_desc_sw128(st + MT * TILE_BYTES);
#pragma unroll
                for (int k = 0; k < BK / 16; k++) {
#pragma unroll
                    for (int mt = 0; mt < MT; mt++) {
                        const uint64_t adesc = umma_desc_sw128(st + mt * TILE_BYTES);
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
        // ---- epilogue warps: TMEM lane quarter = warp % 4
        const int q = warp & 3;
        const int et = threadIdx.x - 64;  // 0..127
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        const int tile_id = blockIdx.y * gridDim.x + blockIdx.x;
        const bool splitk = g.n_split > 1;
        float* my_part = splitk ? g.partials + ((size_t)tile_id * g.n_split + split) * (size_t)(128 * MT) * BN : nullptr;
#pragma unroll 1
        for (int mt = 0; mt < MT; mt++) {
            const int lrow = mt * 128 + q * 32 + lane;
            const int row = m0 + lrow;
#pragma unroll 1
            for (int c = 0; c < BN / 32; c++) {
                uint32_t v[32];
                tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + mt * BN + c * 32, v);
                tmem_ld_wait();
                if (splitk) {
                    float4* dst = reinterpret_cast<float4*>(my_part + (size_t)lrow * BN + c * 32);
#pragma unroll
                    for (int e = 0; e < 8; e++)
                        dst[e] = make_float4(__uint_as_float(v[4 * e]), __uint_as_float(v[4 * e + 1]),
                                             __uint_as_float(v[4 * e + 2]), __uint_as_float(v[4 * e + 3]));
                } else if (row < g.M) {
                    float acc[32];
#pragma unroll
                    for (int e = 0; e < 32; e++) acc[e] = __uint_as_float(v[e]);
                    store_row32_rt(g.epi, acc, g.out + (size_t)row * g.ldo, g.bias,
                                   g.residual ? g.residual + (size_t)row * g.ldo : nullptr, n0 + c * 32, g.N);
                }
            }
        }
        if (splitk) {
            __threadfence();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (et == 0) {
                const int old = atomicAdd(&g.counters[tile_id], 1);
                *last_flag = (old == g.n_split - 1);
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (*last_flag) {
                __threadfence();
                const float* base = g.partials + (size_t)tile_id * g.n_split * (size_t)(128 * MT) * BN;
#pragma unroll 1
                for (int mt = 0; mt < MT; mt++) {
                    const int lrow = mt * 128 + q * 32 + lane;
                    const int row = m0 + lrow;
                    if (row >= g.M) continue;
#pragma unroll 1
                    for (int c = 0; c < BN / 32; c++) {
                        if (n0 + c * 32 >= g.N) break;
                        float acc[32];
#pragma unroll
                        for (int e = 0; e < 32; e++) acc[e] = 0.f;
                        for (int z = 0; z < g.n_split; z++) {
                            const float4* src = reinterpret_cast<const float4*>(
                                base + (size_t)z * (128 * MT) * BN + (size_t)lrow * BN + c * 32);
#pragma unroll
                            for (int e = 0; e < 8; e++) {
                                float4 t = __ldcg(src + e);
                                acc[4 * e] += t.x; acc[4 * e + 1] += t.y; acc[4 * e + 2] += t.z; acc[4 * e + 3] += t.w;
                            }
                        }
                        store_row32_rt(g.epi, acc, g.out + (size_t)row * g.ldo, g.bias,
                                       g.residual ? g.residual + (size_t)row * g.ldo : nullptr, n0 + c * 32, g.N);
                    }
                }
                if (et == 0) g.counters[tile_id] = 0;  // leave the counters zeroed for the next launch
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

}  // namespace

// Launch plan shared by the C entry point and the model step.
struct GemmPlan {
    int mt, grid_x, grid_y, n_split, kb_per_split;
};

static GemmPlan plan_gemm(int M, int N, int K) {
    GemmPlan p;
    p.mt = (M > 128) ? 2 : 1;
    p.grid_x = (N + BN - 1) / BN;
    p.grid_y = (M + 128 * p.mt - 1) / (128 * p.mt);
    const int tiles = p.grid_x * p.grid_y;
    const int kb = (K + BK - 1) / BK;
    int s = 148 / tiles;                       // fill the SMs once
    if (s > kb / 2) s = kb / 2;                // keep >= 2 K-blocks per split
    if (s > MAX_SPLIT) s = MAX_SPLIT;
    if (s < 1) s = 1;
    if (tiles > N_COUNTERS || tiles * s > WS_TILES) s = 1;
    p.kb_per_split = (kb + s - 1) / s;
    p.n_split = (kb + p.kb_per_split - 1) / p.kb_per_split;  // no empty splits
    return p;
}

int linear_f16_launch(int M, int N, int K, const void* A, long lda, const void* W, long ldw, void* out, long ldo,
                      const void* bias, const void* residual, int epilogue, void* workspace, long workspace_bytes,
                      cudaStream_t stream) {
    if (M < 0 || N <= 0 || K <= 0 || (K & 7) || (N & 7) || (lda & 7) || (ldw & 7) || (ldo & 7)) return B200_ERR_BAD_ARG;
    if (epilogue < B200_EPI_NONE || epilogue > B200_EPI_RESADD) return B200_ERR_BAD_ARG;
    if (epilogue == B200_EPI_RESADD && residual == nullptr) return B200_ERR_BAD_ARG;
    if (M == 0) return 0;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(gemm_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<1>::kSmemBytes);
        if (e != cudaSuccess) return (int)e;
        e = cudaFuncSetAttribute(gemm_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<2>::kSmemBytes);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }
    GemmPlan p = plan_gemm(M, N, K);
    if (p.n_split > 1) {
        const long need = (long)N_COUNTERS * 4 + (long)p.grid_x * p.grid_y * p.n_split * (128L * p.mt) * BN * 4;
        if (workspace == nullptr || workspace_bytes < need) return B200_ERR_WORKSPACE;
    }
    CUtensorMap tmA, tmW;
    int rc = make_tmap_2d_f16(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 2, 128, BK, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
    if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
    rc = make_tmap_2d_f16(&tmW, W, (uint64_t)N, (uint64_t)K, (uint64_t)ldw * 2, 128, BK, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
    if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
    GemmArgs g;
    g.M = M; g.N = N; g.K = K;
    g.out = static_cast<__half*>(out);
    g.ldo = ldo;
    g.bias = static_cast<const __half*>(bias);
    g.residual = static_cast<const __half*>(residual);
    g.epi = epilogue;
    g.counters = static_cast<int*>(workspace);
    g.partials = workspace ? reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + (size_t)N_COUNTERS * 4) : nullptr;
    g.kb_per_split = p.kb_per_split;
    g.n_split = p.n_split;
    dim3 grid(p.grid_x, p.grid_y, p.n_split);
    if (p.mt == 2)
        gemm_kernel<2><<<grid, GEMM_THREADS, Cfg<2>::kSmemBytes, stream>>>(tmA, tmW, g);
    else
        gemm_kernel<1><<<grid, GEMM_THREADS, Cfg<1>::kSmemBytes, stream>>>(tmA, tmW, g);
    return (int)cudaGetLastError();
}

}  // namespace b200

extern "C" long b200_linear_workspace_bytes(int /*max_M*/, int /*max_N*/) {
    return (long)b200::N_COUNTERS * 4 + (long)b200::WS_TILES * 256L * b200::BN * 4;
}

extern "C" int b200_linear_f16(int M, int N, int K, const void* A, long lda, const void* W, long ldw, void* out, long ldo,
                               const void* bias, const void* residual, int epilogue, void* workspace, long workspace_bytes,
                               void* stream) {
    return b200::linear_f16_launch(M, N, K, A, lda, W, ldw, out, ldo, bias, residual, epilogue, workspace, workspace_bytes,
                                   static_cast<cudaStream_t>(stream));
}

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
#include <cstdlib>

#include "common.cuh"
#include "gemm_common.cuh"
#include "glue.h"
#include "tmap.h"
#include "../../include/chirrup_b200.h"

namespace b200 {
namespace {

constexpr int GEMM_THREADS = 192;
constexpr int MAX_SPLIT = 8;                      // portable cluster size
constexpr int UNIT_FLOATS = 128 * STG_LD;         // one parked unit: 128 rows x 32 columns (+4 pad) fp32 = 18 KiB
constexpr long WS_BYTES = 160L * 64 * UNIT_FLOATS * 4;  // > 148 CTAs x the largest tile (256 x 256 = 64 units)

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

DEVI unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define PROBE(slot)                                                                                         \
    do {                                                                                                    \
        if (g.dbg && (threadIdx.x == 64 || ((slot) == 2 && threadIdx.x == 32)))                             \
            g.dbg[((size_t)(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 8 + (slot)] = gtime(); \
    } while (0)

// shared -> global linear bulk copy (TMA engine), tracked by the issuing thread's bulk async-group
DEVI void bulk_store(void* gdst, const void* ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(reinterpret_cast<uint64_t>(gdst)),
                 "r"(smem_u32(ssrc)), "r"(bytes)
                 : "memory");
}

DEVI void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// The body of one CTA: output tile (m0, n0) of problem `g`, K split blockIdx.z.  `tmAp` / `tmWp` point into kernel-parameter space.
template <int MT, int BN>
DEVI void gemm_body(const CUtensorMap* tmAp, const CUtensorMap* tmWp, const GemmArgs& g, const int n0, const int m0,
                    const int tile_id) {
    using C = Cfg<MT, BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
    uint64_t* empty_bar = full_bar + C::kStages;
    uint64_t* accum_bar = empty_bar + C::kStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int split = blockIdx.z;
    const int kb_total = (g.K + BK - 1) / BK;
    const int kb0 = split * g.kb_per_split;
    const int kb1 = min(kb_total, kb0 + g.kb_per_split);
    const int n_kb = kb1 - kb0;  // >= 1 by construction of the launch
    const bool splitk = g.n_split > 1;
    const int epi = g.epi;
    EpiCtx ec;
    ec.out = g.out; ec.bias = g.bias; ec.residual = g.residual; ec.ldo = g.ldo; ec.M = g.M; ec.N = g.N;
    pdl_launch_dependents();
    PROBE(0);

    if (threadIdx.x == 0) {
        tma_prefetch_desc(tmAp);
        tma_prefetch_desc(tmWp);
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
    pdl_wait();  // everything above overlapped the previous kernel's tail; A / residual / workspace are touched only below
    PROBE(1);

    const int q = warp & 3;  // TMEM lane quarter of an epilogue warp

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
                    tma_load_2d_hint(st + mt * (128 * BK * 2), tmAp, kc, m0 + mt * 128, &full_bar[s], pol_a);
                tma_load_2d_hint(st + C::kABytes, tmWp, kc, n0, &full_bar[s], pol_w);
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
                if (it == 0) PROBE(2);
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
        // ---- epilogue warps: drain TMEM.  tcgen05.ld gives thread = row.  Without a K split each 32 x 32 unit goes through
        // emit_unit (gemm_common.cuh: per-warp shared-memory transpose -> coalesced 16-byte stores with the fused epilogue);
        // with a K split the unit is parked in the L2-resident workspace in thread-major order (512 contiguous bytes per warp
        // store, straight from the registers).
        mbar_wait(accum_bar, 0);
        PROBE(3);
        tc_fence_after();
        const uint32_t stg0 = smem_u32(smem) + (uint32_t)q * (2 * 32 * STG_LD * 4);  // two staging buffers per warp (idle stages)
        const int row = q * 32 + lane;
        float* my_part = splitk ? g.partials + ((size_t)tile_id * g.n_split + split) * (size_t)C::kUnits * UNIT_FLOATS : nullptr;
        int cnt = 0;
#pragma unroll 1
        for (int u = 0; u < C::kUnits; u++) {
            const int mt = u / (BN / 32), c = u - mt * (BN / 32);
            if (n0 + c * 32 >= g.N) continue;  // columns past N (N tail of the last tile)
            float v[32];
            tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + mt * BN + c * 32,
                               reinterpret_cast<uint32_t (&)[32]>(v));
            tmem_ld_wait();
            if (splitk) {
                float4* dst = reinterpret_cast<float4*>(my_part + (size_t)u * UNIT_FLOATS) + row;
#pragma unroll
                for (int e = 0; e < 8; e++) __stcg(dst + e * 128, make_float4(v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]));
            } else {
                EPI_DISPATCH(epi, emit_unit<E>(v, stg0 + (uint32_t)(cnt & 1) * (32 * STG_LD * 4), ec, m0 + mt * 128 + q * 32,
                                               n0 + c * 32, lane));
                cnt++;
            }
        }
        if (splitk) __threadfence();
    }

    if (splitk) {
        // every CTA of the cluster (= every K split of this tile) has parked its partial: release/acquire at cluster scope
        PROBE(4);
        cluster_sync_all();
        PROBE(5);
        if (warp >= 2) {
            const float* base = g.partials + (size_t)tile_id * g.n_split * (size_t)C::kUnits * UNIT_FLOATS;
            const uint32_t stg0 = smem_u32(smem) + (uint32_t)q * (2 * 32 * STG_LD * 4);
            const int row = q * 32 + lane;
            int cnt = 0;
#pragma unroll 1
            for (int u = split; u < C::kUnits; u += g.n_split) {  // this CTA's share of the tile
                const int mt = u / (BN / 32), c = u - mt * (BN / 32);
                if (n0 + c * 32 >= g.N) continue;
                float acc[32];
#pragma unroll
                for (int e = 0; e < 32; e++) acc[e] = 0.f;
#pragma unroll 2
                for (int z = 0; z < g.n_split; z++) {  // fixed order: deterministic sum
                    const float4* src = reinterpret_cast<const float4*>(base + ((size_t)z * C::kUnits + u) * UNIT_FLOATS) + row;
                    float4 t[8];
#pragma unroll
                    for (int e = 0; e < 8; e++) t[e] = __ldcg(src + e * 128);
#pragma unroll
                    for (int e = 0; e < 8; e++) {
                        acc[4 * e] += t[e].x; acc[4 * e + 1] += t[e].y; acc[4 * e + 2] += t[e].z; acc[4 * e + 3] += t[e].w;
                    }
                }
                EPI_DISPATCH(epi, emit_unit<E>(acc, stg0 + (uint32_t)(cnt & 1) * (32 * STG_LD * 4), ec, m0 + mt * 128 + q * 32,
                                               n0 + c * 32, lane));
                cnt++;
            }
        }
    }

    PROBE(6);
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<C::kTmemCols>(tmem_base);
    }
}

template <int MT, int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
    gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const GemmArgs g) {
    gemm_body<MT, BN>(&tmA, &tmW, g, blockIdx.x * BN, blockIdx.y * (128 * MT), blockIdx.y * gridDim.x + blockIdx.x);
}

// Several independent small problems sharing M, the tile shape and the K-split factor in ONE launch (the four LoRA
// down-projections, then the four up-projections of a layer): blockIdx.x enumerates the N-tiles of all problems back to back.
constexpr int MAX_GROUP = 4;
struct alignas(64) GemmGroup {
    CUtensorMap tmA[MAX_GROUP];
    CUtensorMap tmW[MAX_GROUP];
    GemmArgs g[MAX_GROUP];
    int tile_end[MAX_GROUP];
    int n;
};

template <int MT, int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1) gemm_group_kernel(const __grid_constant__ GemmGroup grp) {
    int p = 0;
    while (p + 1 < grp.n && (int)blockIdx.x >= grp.tile_end[p]) p++;
    const int local = blockIdx.x - (p ? grp.tile_end[p - 1] : 0);
    gemm_body<MT, BN>(&grp.tmA[p], &grp.tmW[p], grp.g[p], local * BN, blockIdx.y * (128 * MT),
                      blockIdx.y * gridDim.x + blockIdx.x);
}

struct GemmPlan {
    int mt, bn, grid_x, grid_y, n_split, kb_per_split;
};

template <int MT, int BN>
static int launch_cfg(const GemmPlan& p, const CUtensorMap& tmA, const CUtensorMap& tmW, const GemmArgs& g, cudaStream_t stream,
                      bool pdl) {
    using C = Cfg<MT, BN>;
    cudaError_t e = cudaFuncSetAttribute(gemm_kernel<MT, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes);
    if (e != cudaSuccess) return (int)e;  // per device, cheap: set on every launch (one process may drive several GPUs)
    return launch_kernel(gemm_kernel<MT, BN>, dim3(p.grid_x, p.grid_y, p.n_split), dim3(GEMM_THREADS), C::kSmemBytes, stream, pdl,
                         p.n_split, tmA, tmW, g);
}

// How many clusters of `s` CTAs of this configuration the device can hold at once (GPC boundaries strand SMs for big clusters).
template <int MT, int BN>
static int max_clusters_cfg(int s) {
    using C = Cfg<MT, BN>;
    static int cache[MAX_SPLIT + 1] = {0};
    if (cache[s]) return cache[s];
    cudaFuncSetAttribute(gemm_kernel<MT, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1, 1, s);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = C::kSmemBytes;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = s;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, gemm_kernel<MT, BN>, &cfg) != cudaSuccess || n <= 0) {
        cudaGetLastError();
        n = 148 / s - (s > 2 ? 2 : 0);  // conservative guess
    }
    cache[s] = n;
    return n;
}

static int max_clusters(int mt, int bn, int s) {
    if (mt == 2) return bn == 256 ? max_clusters_cfg<2, 256>(s) : bn == 128 ? max_clusters_cfg<2, 128>(s) : max_clusters_cfg<2, 32>(s);
    return bn == 256 ? max_clusters_cfg<1, 256>(s) : bn == 128 ? max_clusters_cfg<1, 128>(s) : max_clusters_cfg<1, 32>(s);
}

// Tile shape and K split.  Candidates (BN, S) are ranked by a small cost model of one CTA's life (measured rates):
// operand bytes through its shared memory, fp32 partial bytes parked + re-read when S > 1, and the number of waves.
static GemmPlan plan_gemm(int M, int N, int K) {
    const int kb = (K + BK - 1) / BK;
    const int mt = (M > 128) ? 2 : 1;
    const int gy = (M + 128 * mt - 1) / (128 * mt);
    GemmPlan best = {};
    double best_cost = 1e30;
    const int bns[3] = {256, 128, 32};
    for (int bi = 0; bi < 3; bi++) {
        const int bn = bns[bi];
        if (bn > 32 && N < 2 * bn) continue;  // skinny outputs: small tiles
        const int gx = (N + bn - 1) / bn;
        const long tiles = (long)gx * gy;
        for (int s = 1; s <= MAX_SPLIT; s++) {
            if (s > 1 && (s > kb / 2 || tiles * s > 148)) break;
            const int kps = (kb + s - 1) / s;
            const int ns = (kb + kps - 1) / kps;
            if (ns != s) continue;
            const long slots = (long)max_clusters(mt, bn, s);
            const long waves = (tiles + slots - 1) / slots;
            const double ingest = (double)kps * (mt * 128 + bn) * BK * 2;            // bytes through one CTA's smem
            const double chip = ingest * (double)(tiles < slots ? tiles : slots) * s;  // concurrent L2 -> SM traffic of a wave
            double t_main = ingest / 110e3;                                          // us at ~110 GB/s per SM
            if (chip / 10.0e6 > t_main) t_main = chip / 10.0e6;                      // or the ~10 TB/s L2 ceiling
            const double part = (s > 1) ? (double)mt * 128 * bn * 4 * (STG_LD / 32.0) : 0.0;
            const double t_epi = (s > 1) ? part / 120e3 + 1.0 + part / 70e3 : (double)mt * 128 * bn * 2 / 40e3;
            const double cost = waves * (2.0 + t_main + t_epi);
            if (cost < best_cost) {
                best_cost = cost;
                best.mt = mt; best.bn = bn; best.grid_x = gx; best.grid_y = gy; best.n_split = s; best.kb_per_split = kps;
            }
        }
    }
    return best;
}

}  // namespace

static unsigned long long* g_gemm_dbg = nullptr;
static bool no_pair() {
    static const bool v = getenv("B200_NO_PAIR") != nullptr;  // A/B switch: one-CTA tiles (this file) for every M
    return v;
}
void gemm_set_debug_buffer(unsigned long long* p) { g_gemm_dbg = p; }

int linear_f16_launch(int M, int N, int K, const void* A, long lda, const void* W, long ldw, void* out, long ldo,
                      const void* bias, const void* residual, int epilogue, void* workspace, long workspace_bytes,
                      cudaStream_t stream, bool pdl, int* defer_splits) {
    if (defer_splits) *defer_splits = 0;
    if (M < 0 || N <= 0 || K <= 0 || (K & 7) || (N & 7) || (lda & 7) || (ldw & 7) || (ldo & 7)) return B200_ERR_BAD_ARG;
    if (epilogue < B200_EPI_NONE || epilogue > B200_EPI_RESADD) return B200_ERR_BAD_ARG;
    if (epilogue == B200_EPI_RESADD && residual == nullptr) return B200_ERR_BAD_ARG;
    // TMA needs 16-byte aligned operands; the epilogue moves 8 fp16 per access
    if ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(W) | reinterpret_cast<uintptr_t>(out) |
         reinterpret_cast<uintptr_t>(bias) | reinterpret_cast<uintptr_t>(residual)) & 15)
        return B200_ERR_BAD_ARG;
    if (M == 0) return 0;
    if (M > 128 && !no_pair()) {  // decode / prefill sized token counts: CTA pairs (gemm_pair.cu)
        LinearDesc d1 = {N, K, A, lda, W, ldw, out, ldo, bias, residual, epilogue};
        const int rc = pair_group_launch(1, &d1, M, workspace, workspace_bytes, stream, pdl, g_gemm_dbg, defer_splits);
        if (rc != B200_ERR_UNSUPPORTED) return rc;
    }
    GemmPlan p = plan_gemm(M, N, K);
    if (p.n_split > 1) {
        const long need = (long)p.grid_x * p.grid_y * p.n_split * (long)p.mt * (p.bn / 32) * UNIT_FLOATS * 4;
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
    g.dbg = g_gemm_dbg;
    g.defer = nullptr;
    if (p.mt == 2)
        return p.bn == 256 ? launch_cfg<2, 256>(p, tmA, tmW, g, stream, pdl)
               : p.bn == 128 ? launch_cfg<2, 128>(p, tmA, tmW, g, stream, pdl) : launch_cfg<2, 32>(p, tmA, tmW, g, stream, pdl);
    return p.bn == 256 ? launch_cfg<1, 256>(p, tmA, tmW, g, stream, pdl)
           : p.bn == 128 ? launch_cfg<1, 128>(p, tmA, tmW, g, stream, pdl) : launch_cfg<1, 32>(p, tmA, tmW, g, stream, pdl);
}

template <int MT, int BN>
static int launch_group_cfg(const GemmGroup& grp, int grid_x, int grid_y, int s, cudaStream_t stream, bool pdl) {
    using C = Cfg<MT, BN>;
    cudaError_t e = cudaFuncSetAttribute(gemm_group_kernel<MT, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes);
    if (e != cudaSuccess) return (int)e;
    return launch_kernel(gemm_group_kernel<MT, BN>, dim3(grid_x, grid_y, s), dim3(GEMM_THREADS), C::kSmemBytes, stream, pdl, s, grp);
}

// n <= 4 problems with the same M in one launch.  Falls back to one launch per problem when the group does not fit the
// single-wave shapes it is meant for (large M).
int linear_group_launch(int n, const LinearDesc* d, int M, void* workspace, long workspace_bytes, cudaStream_t stream, bool pdl) {
    if (n <= 0 || M <= 0) return 0;
    if (M > 128 && !no_pair()) {
        const int rc = pair_group_launch(n, d, M, workspace, workspace_bytes, stream, pdl, nullptr);
        if (rc != B200_ERR_UNSUPPORTED) return rc;
    }
    bool ok = n <= MAX_GROUP && M <= 256;
    int max_n = 0, min_kb = 1 << 30;
    for (int i = 0; i < n; i++) {
        if ((d[i].K & 7) || (d[i].N & 7) || (d[i].lda & 7) || (d[i].ldw & 7) || (d[i].ldo & 7)) return B200_ERR_BAD_ARG;
        if ((reinterpret_cast<uintptr_t>(d[i].A) | reinterpret_cast<uintptr_t>(d[i].W) | reinterpret_cast<uintptr_t>(d[i].out) |
             reinterpret_cast<uintptr_t>(d[i].bias) | reinterpret_cast<uintptr_t>(d[i].residual)) & 15)
            return B200_ERR_BAD_ARG;
        if (d[i].N > max_n) max_n = d[i].N;
        const int kb = (d[i].K + BK - 1) / BK;
        if (kb < min_kb) min_kb = kb;
    }
    const int mt = (M > 128) ? 2 : 1;
    const int bn = (max_n <= 512) ? 32 : 128;
    int tiles = 0;
    for (int i = 0; i < n; i++) tiles += (d[i].N + bn - 1) / bn;
    if (tiles > 148) ok = false;
    if (!ok) {
        for (int i = 0; i < n; i++) {
            int rc = linear_f16_launch(M, d[i].N, d[i].K, d[i].A, d[i].lda, d[i].W, d[i].ldw, d[i].out, d[i].ldo, d[i].bias,
                                       d[i].residual, d[i].epi, workspace, workspace_bytes, stream, pdl);
            if (rc) return rc;
        }
        return 0;
    }
    // common split factor: one wave, >= 2 K-blocks per split in every problem, no empty split in any problem
    int s = 148 / tiles;
    if (s > min_kb / 2) s = min_kb / 2;
    if (s > MAX_SPLIT) s = MAX_SPLIT;
    if (s < 1) s = 1;
    for (; s > 1; s--) {
        bool fits = tiles <= max_clusters(mt, bn, s);
        for (int i = 0; i < n && fits; i++) {
            const int kb = (d[i].K + BK - 1) / BK, kps = (kb + s - 1) / s;
            if ((kb + kps - 1) / kps != s) fits = false;
        }
        if (fits) break;
    }
    if (s > 1) {
        const long need = (long)tiles * s * mt * (bn / 32) * UNIT_FLOATS * 4;
        if (workspace == nullptr || workspace_bytes < need) return B200_ERR_WORKSPACE;
    }
    GemmGroup grp;
    grp.n = n;
    int end = 0;
    for (int i = 0; i < n; i++) {
        int rc = make_tmap_2d_f16(&grp.tmA[i], d[i].A, (uint64_t)M, (uint64_t)d[i].K, (uint64_t)d[i].lda * 2, 128, BK,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
        if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
        rc = make_tmap_2d_f16(&grp.tmW[i], d[i].W, (uint64_t)d[i].N, (uint64_t)d[i].K, (uint64_t)d[i].ldw * 2, bn, BK,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
        if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
        GemmArgs& g = grp.g[i];
        g.M = M; g.N = d[i].N; g.K = d[i].K;
        g.out = static_cast<__half*>(d[i].out);
        g.ldo = d[i].ldo;
        g.bias = static_cast<const __half*>(d[i].bias);
        g.residual = static_cast<const __half*>(d[i].residual);
        g.epi = d[i].epi;
        g.partials = static_cast<float*>(workspace);
        const int kb = (d[i].K + BK - 1) / BK;
        g.kb_per_split = (kb + s - 1) / s;
        g.n_split = s;
        g.dbg = nullptr;
        g.defer = nullptr;
        end += (d[i].N + bn - 1) / bn;
        grp.tile_end[i] = end;
    }
    for (int i = n; i < MAX_GROUP; i++) grp.tile_end[i] = end;
    if (mt == 2) return bn == 32 ? launch_group_cfg<2, 32>(grp, tiles, 1, s, stream, pdl) : launch_group_cfg<2, 128>(grp, tiles, 1, s, stream, pdl);
    return bn == 32 ? launch_group_cfg<1, 32>(grp, tiles, 1, s, stream, pdl) : launch_group_cfg<1, 128>(grp, tiles, 1, s, stream, pdl);
}

}  // namespace b200

extern "C" long b200_linear_workspace_bytes(int /*max_M*/, int /*max_N*/) { return b200::WS_BYTES; }

extern "C" int b200_linear_f16(int M, int N, int K, const void* A, long lda, const void* W, long ldw, void* out, long ldo,
                               const void* bias, const void* residual, int epilogue, void* workspace, long workspace_bytes,
                               void* stream) {
    return b200::linear_f16_launch(M, N, K, A, lda, W, ldw, out, ldo, bias, residual, epilogue, workspace, workspace_bytes,
                                   static_cast<cudaStream_t>(stream), /*pdl=*/false);
}

extern "C" int b200_linear_group_f16(int n, const b200_linear_desc* problems, int M, void* workspace, long workspace_bytes,
                                     void* stream) {
    if (n < 0 || M < 0 || (n > 0 && !problems)) return B200_ERR_BAD_ARG;
    static_assert(sizeof(b200_linear_desc) == sizeof(b200::LinearDesc), "descriptor layouts must match");
    for (int i = 0; i < n; i++) {
        const int e = problems[i].epilogue;
        if (e < B200_EPI_NONE || e > B200_EPI_RESADD || (e == B200_EPI_RESADD && !problems[i].residual)) return B200_ERR_BAD_ARG;
        if (problems[i].N <= 0 || problems[i].K <= 0) return B200_ERR_BAD_ARG;
    }
    return b200::linear_group_launch(n, reinterpret_cast<const b200::LinearDesc*>(problems), M, workspace, workspace_bytes,
                                     static_cast<cudaStream_t>(stream), /*pdl=*/false);
}

namespace b200 { int pair_slots_query(int s); }
// Debug aid (not part of the public header): CTA pairs x K-splits the device can co-schedule (cudaOccupancyMaxActiveClusters).
extern "C" int b200_debug_pair_slots(int s) { return b200::pair_slots_query(s); }

// Debug aid (not part of the public header): point the GEMM timeline probes at a device buffer of >= 8 * n_ctas uint64.
extern "C" void b200_debug_gemm_probe(void* buf) { b200::gemm_set_debug_buffer(static_cast<unsigned long long*>(buf)); }

// Decode-sized dense linear layers on CTA PAIRS:  out (M,N) = epilogue( A (M,K) . W (N,K)^T )  for 128 < M, fp16 in, fp32 in TMEM.
//
// Same reference call sites as gemm.cu (/root/reference/Albatross/rwkv7.py:625-630, :637, :649, :678-679, :551).  What differs
// is the mapping to the machine.  At M = 256 one output tile needs every token row, so the per-SM cost of a K-step is the
// bytes that enter the SM's shared memory (measured ceiling ~100 GB/s per SM).  Two SMs of a TPC therefore share each tile with
// tcgen05.mma.cta_group::2 (UMMA 256 x BN x 16): CTA r of the pair loads token rows [128r, 128r+128) of A and rows
// [r BN/2, (r+1) BN/2) of the W tile, the leader CTA issues the MMAs for both, and each CTA ends up with the 128 x BN fp32
// accumulator of its own token rows in its own TMEM — half the shared-memory fill per SM of a one-CTA tile of the same shape,
// twice the CTAs per tile, so that far less K-splitting is needed to occupy the chip.
//
//   warp 0 (one lane)   TMA producer: its CTA's A rows and W-half rows per 64-wide K block, completion on the LEADER's mbarrier
//   warp 1 (one lane)   leader only: tcgen05.mma issuer; tcgen05.commit multicast frees the stage in both CTAs
//   warps 2-9           epilogue: tcgen05.ld -> fp32 staging in shared memory -> coalesced bias / activation / residual -> fp16;
//                       two warps per TMEM lane quarter (warp & 3), one takes the even 32-column units, the other the odd ones:
//                       a one-tile launch's epilogue is on the step's critical path and a single warp per scheduler issues
//                       one instruction per ~5 cycles (ncu: stall_wait / short_scoreboard), so a second warp halves it
//
// One launch carries up to PAIR_MAX_GROUP independent problems that share M (r,k,v and the four LoRA down-projections of a
// layer are one launch), each with its own tile width BN (multiple of 32, <= 256).  K may additionally be split over the z
// dimension of the cluster (cluster = 2 x 1 x S): every CTA parks its fp32 partial in an L2-resident workspace in a
// thread-major layout (coalesced 512-byte warp stores straight from the TMEM registers), the cluster barrier orders the
// writes, and each CTA then pulls the S partials of its 1/S share of the tile's columns back with bulk copies (all in flight
// at once) and finishes them in split order: deterministic, no atomics.
//
// Programmatic dependent launch: the weight tiles of the first pipeline stages are requested BEFORE griddepcontrol.wait —
// they do not depend on the preceding kernel — so the weight stream starts under the previous kernel's tail.
#include <cstdlib>

#include "common.cuh"
#include "gemm_common.cuh"
#include "glue.h"
#include "tmap.h"

namespace b200 {
namespace {

constexpr int PAIR_THREADS = 320;      // one-tile kernel: TMA warp, MMA warp, EIGHT epilogue warps (two per TMEM lane quarter)
constexpr int PERSIST_THREADS = 320;   // persistent kernel: the same roles
constexpr int PAIR_MAX_GROUP = 8;
constexpr int PAIR_STAGE_BYTES = 32 * 1024;       // A half 128 x 64 fp16 (16 KiB) + W half up to 128 x 64 fp16 (16 KiB)
constexpr int PAIR_MAX_STAGES = 6;
constexpr int PAIR_MAX_SPLIT = 4;                 // cluster = 2 x 1 x S <= 8 CTAs (portable size)
constexpr int PAIR_UNIT_FLOATS = 128 * 32;        // one parked unit: 128 rows x 32 columns fp32 = 16 KiB, thread-major
constexpr int PAIR_STAGING_BYTES = 4 * 2 * 32 * STG_LD * 4;  // two 32 x 32(+4) fp32 buffers per epilogue warp = 36 KiB (four warps: K-split reduce)
constexpr int PAIR_STAGING_BYTES_8 = 2 * PAIR_STAGING_BYTES;  // the same for all eight epilogue warps (no K split / deferred split)
constexpr uint32_t PAIR_TMEM_COLS = 256;

struct PairProblem {
    CUtensorMap tmA;  // box 128 rows x 64
    CUtensorMap tmW;  // box bn/2 rows x 64
    GemmArgs g;
    int tile_end;     // one past this problem's last tile index along blockIdx.x / 2
    int bn;           // tile width: multiple of 32, <= 256
    // persistent kernel only — chained problems inside one launch (FFN key -> FFN value):
    int kb0;          // first K block of this problem's K range (its length is g.kb_per_split)
    int dep_flag;     // >= 0: tiles of this problem may load A only once flags[dep_flag] >= dep_need
    int dep_need;
    int sig_flag;     // >= 0: every epilogue warp adds 1 to flags[sig_flag + local_tile / sig_div] when its rows are stored
    int sig_div;
};

struct alignas(64) PairGroup {
    PairProblem p[PAIR_MAX_GROUP];
    int n;
    int n_stages;
    int* flags;       // chained launches: [0, 8) arrival counters, [8, 16) consumer counters (both back to 0 when the launch ends)
};

DEVI uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
DEVI uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
DEVI void cluster_arrive_wait() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// 2-D tiled load whose completion bytes are signalled on an mbarrier given as a shared::cluster address (the pair leader's).
DEVI void tma_load_2d_pair(void* dst, const CUtensorMap* m, int c0, int c1, uint32_t bar_cluster_addr, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, "
        "%3}], [%4], %5;" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(c0), "r"(c1), "r"(bar_cluster_addr), "l"(policy)
        : "memory");
}
DEVI void umma_f16_ss_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Arrive (count 1) on the mbarrier at this shared-memory offset in every CTA of `cta_mask` once all MMAs issued so far are done.
DEVI void umma_commit_pair(uint64_t* bar, uint16_t cta_mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)),
                 "h"(cta_mask)
                 : "memory");
}
DEVI void tmem_alloc_pair(uint32_t* smem_dst, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
DEVI void tmem_dealloc_pair(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// global -> shared linear bulk copy, completion on a local mbarrier
DEVI void bulk_load(void* sdst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(sdst)),
                 "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
DEVI unsigned long long gtime_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define PPROBE(slot)                                                                                              \
    do {                                                                                                          \
        if (g.dbg && (threadIdx.x == 64 || ((slot) == 2 && threadIdx.x == 32)))                                   \
            g.dbg[((size_t)(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 8 + (slot)] = gtime_ns(); \
    } while (0)

DEVI void tmem_ld_unit(uint32_t tmem_base, int q, int u, float (&v)[32]) {
    uint32_t (&w)[32] = reinterpret_cast<uint32_t (&)[32]>(v);
    tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + u * 32, w);
}

// No K split: drain this warp's 32 accumulator rows unit by unit (the TMEM load of unit u+1 is in flight while unit u is emitted).
// This warp takes units u0, u0 + 2, ... (u0 = 0 or 1: the two warps of a lane quarter).
template <int EPI>
DEVI void epilogue_direct(uint32_t tmem_base, int q, int lane, int u0, int n_valid, const EpiCtx& c, int row0, int n0, uint32_t stg0) {
    if (u0 >= n_valid) return;
    float va[32], vb[32];
    tmem_ld_unit(tmem_base, q, u0, va);
#pragma unroll 1
    for (int u = u0; u < n_valid; u += 4) {
        tmem_ld_wait();
        if (u + 2 < n_valid) tmem_ld_unit(tmem_base, q, u + 2, vb);
        emit_unit<EPI>(va, stg0, c, row0, n0 + u * 32, lane);
        if (u + 2 < n_valid) {
            tmem_ld_wait();
            if (u + 4 < n_valid) tmem_ld_unit(tmem_base, q, u + 4, va);
            emit_unit<EPI>(vb, stg0 + 32 * STG_LD * 4, c, row0, n0 + (u + 2) * 32, lane);
        }
    }
}

// The same with ONE staging buffer per warp (the persistent kernel spends its shared memory on a sixth pipeline stage): a warp
// barrier separates the reads of unit u from the writes of unit u+1.
template <int EPI>
DEVI void epilogue_direct_1buf(uint32_t tmem_base, int q, int lane, int u0, int n_valid, const EpiCtx& c, int row0, int n0,
                               uint32_t stg0) {
#pragma unroll 1
    for (int u = u0; u < n_valid; u += 2) {
        float va[32];
        tmem_ld_unit(tmem_base, q, u, va);
        tmem_ld_wait();
        __syncwarp();
        emit_unit<EPI>(va, stg0, c, row0, n0 + u * 32, lane);
    }
}

// Raw fp32 tile (deferred K-split partial, see GemmArgs::defer), one staging buffer per warp.
DEVI void epilogue_f32_1buf(uint32_t tmem_base, int q, int lane, int u0, int n_valid, float* dst, int M, int N, int row0, int n0,
                            uint32_t stg0) {
#pragma unroll 1
    for (int u = u0; u < n_valid; u += 2) {
        float va[32];
        tmem_ld_unit(tmem_base, q, u, va);
        tmem_ld_wait();
        __syncwarp();
        emit_unit_f32(va, stg0, dst, N, M, N, row0, n0 + u * 32, lane);
    }
}

// K split: sum the S landed partials of each of this CTA's units in split order, then emit.
template <int EPI>
DEVI void epilogue_reduce(uint32_t land, int S, int split, int n_valid, int row, int lane, const EpiCtx& c, int row0, int n0,
                          uint32_t stg0) {
    int j = 0;
#pragma unroll 1
    for (int u = split; u < n_valid; u += S, j++) {
        float acc[32];
#pragma unroll
        for (int e = 0; e < 32; e++) acc[e] = 0.f;
#pragma unroll 1
        for (int z = 0; z < S; z++) {  // fixed order: deterministic sum
            const uint32_t src = land + (uint32_t)((j * S + z) * PAIR_UNIT_FLOATS * 4 + row * 16);
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const float4 t = lds128(src + (uint32_t)e * 128u * 16u);
                acc[4 * e] += t.x; acc[4 * e + 1] += t.y; acc[4 * e + 2] += t.z; acc[4 * e + 3] += t.w;
            }
        }
        emit_unit<EPI>(acc, stg0 + (uint32_t)(j & 1) * (32 * STG_LD * 4), c, row0, n0 + u * 32, lane);
    }
}

__global__ void __launch_bounds__(PAIR_THREADS, 1) pair_gemm_kernel(const __grid_constant__ PairGroup grp) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int n_stages = grp.n_stages;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + PAIR_MAX_STAGES * PAIR_STAGE_BYTES);
    uint64_t* empty_bar = full_bar + PAIR_MAX_STAGES;
    uint64_t* accum_bar = empty_bar + PAIR_MAX_STAGES;
    uint64_t* red_bar = accum_bar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(red_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    const int pr = crank & 1;  // rank inside the CTA pair: 0 = leader (issues the MMAs)
    const int tile = blockIdx.x >> 1;
    int pi = 0;
    while (pi + 1 < grp.n && tile >= grp.p[pi].tile_end) pi++;
    const PairProblem& P = grp.p[pi];
    const GemmArgs& g = P.g;
    const int bn = P.bn;
    const int local_tile = tile - (pi ? grp.p[pi - 1].tile_end : 0);
    const int n0 = local_tile * bn;
    const int m0 = blockIdx.y * 256 + pr * 128;
    const int split = blockIdx.z;
    const int S = g.n_split;
    const int kb_total = (g.K + BK - 1) / BK;
    const int kb0 = split * g.kb_per_split;
    const int n_kb = min(kb_total, kb0 + g.kb_per_split) - kb0;  // >= 1 by construction of the launch
    const int n_units = bn >> 5;
    int n_valid = (g.N - n0 + 31) >> 5;  // units of this tile that hold columns < N (tail tile of a problem)
    if (n_valid > n_units) n_valid = n_units;
    const int epi = g.epi;
    EpiCtx ec;
    ec.out = g.out; ec.bias = g.bias; ec.residual = g.residual; ec.ldo = g.ldo; ec.M = g.M; ec.N = g.N;
    const uint32_t w_half_bytes = (uint32_t)(bn >> 1) * BK * 2;
    PPROBE(0);

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&P.tmA);
        tma_prefetch_desc(&P.tmW);
        for (int s = 0; s < n_stages; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(accum_bar, 1);
        mbar_init(red_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_pair(tmem_slot, PAIR_TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    cluster_arrive_wait();  // the peer's barriers exist before anything is signalled across the pair
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    PPROBE(1);

    if (warp == 0) {
        if (lane == 0) {
            const uint64_t pol_w = policy_evict_first();
            const uint64_t pol_a = policy_evict_last();
            const uint32_t tx = 2u * (128u * BK * 2u + w_half_bytes);  // both CTAs' bytes land on the leader's barrier
            const uint32_t leader = crank & ~1u;
            const int npre = min(n_kb, n_stages);
            // weights first: they do not depend on the previous kernel
            for (int it = 0; it < npre; it++) {
                uint8_t* st = smem + it * PAIR_STAGE_BYTES;
                if (pr == 0) mbar_arrive_expect_tx(&full_bar[it], tx);
                tma_load_2d_pair(st + 128 * BK * 2, &P.tmW, (kb0 + it) * BK, n0 + pr * (bn >> 1),
                                 mapa_shared(smem_u32(&full_bar[it]), leader), pol_w);
            }
            pdl_wait();
            for (int it = 0; it < npre; it++)
                tma_load_2d_pair(smem + it * PAIR_STAGE_BYTES, &P.tmA, (kb0 + it) * BK, m0,
                                 mapa_shared(smem_u32(&full_bar[it]), leader), pol_a);
            int s = npre % n_stages;  // running (stage, phase) instead of a division per K block
            uint32_t ph = (uint32_t)(npre / n_stages) & 1;
            for (int it = npre; it < n_kb; it++) {
                mbar_wait(&empty_bar[s], ph ^ 1);
                uint8_t* st = smem + s * PAIR_STAGE_BYTES;
                if (pr == 0) mbar_arrive_expect_tx(&full_bar[s], tx);
                const uint32_t bar = mapa_shared(smem_u32(&full_bar[s]), leader);
                tma_load_2d_pair(st, &P.tmA, (kb0 + it) * BK, m0, bar, pol_a);
                tma_load_2d_pair(st + 128 * BK * 2, &P.tmW, (kb0 + it) * BK, n0 + pr * (bn >> 1), bar, pol_w);
                if (++s == n_stages) {
                    s = 0;
                    ph ^= 1;
                }
            }
        } else {
            pdl_wait();
        }
        __syncwarp();
    } else if (warp == 1) {
        pdl_wait();
        if (pr == 0 && lane == 0) {
            const uint32_t idesc = umma_idesc_f16(256, (uint32_t)bn);
            const uint16_t mask = (uint16_t)(3u << (crank & ~1u));
            long long t_wait = 0, t_issue = 0, t_mma = 0;  // debug probe only: cycles this thread waits for data / spends issuing
            int s = 0;
            uint32_t ph = 0;
            for (int it = 0; it < n_kb; it++) {
                const long long c0 = g.dbg ? clock64() : 0;
                mbar_wait(&full_bar[s], ph);
                const long long c1 = g.dbg ? clock64() : 0;
                if (it == 0) PPROBE(2);
                tc_fence_after();
                const uint32_t st = smem_u32(smem + s * PAIR_STAGE_BYTES);
                const uint64_t adesc = umma_desc_sw128(st);
                const uint64_t bdesc = umma_desc_sw128(st + 128 * BK * 2);
#pragma unroll
                for (int k = 0; k < BK / 16; k++)
                    umma_f16_ss_pair(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
                const long long c2 = g.dbg ? clock64() : 0;
                umma_commit_pair(&empty_bar[s], mask);  // frees the stage in both CTAs once these MMAs have read it
                if (g.dbg && it > 0) {
                    t_wait += c1 - c0;
                    t_issue += clock64() - c1;
                    t_mma += c2 - c1;
                }
                if (++s == n_stages) {
                    s = 0;
                    ph ^= 1;
                }
            }
            if (g.dbg)
                g.dbg[((size_t)(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 8 + 7] =
                    ((unsigned long long)t_wait << 32) | (unsigned long long)(t_issue & 0xffffffffll);
            if (g.dbg)  // the peer CTA's slot 7 is free: the share of `issue` spent on the four tcgen05.mma alone
                g.dbg[((size_t)(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x + 1) * 8 + 7] = (unsigned long long)t_mma;
            umma_commit_pair(accum_bar, mask);  // accumulators complete (both CTAs)
        }
        __syncwarp();
    } else {
        pdl_wait();
        const int q = warp & 3;          // TMEM lane quarter of this warp
        const int u0 = (warp - 2) >> 2;  // its units: u0, u0 + 2, ...
        mbar_wait(accum_bar, 0);
        PPROBE(3);
        tc_fence_after();
        const int row = q * 32 + lane;  // this thread's accumulator row inside the CTA's 128
        const uint32_t stg0 = smem_u32(smem) + (uint32_t)(warp - 2) * (2 * 32 * STG_LD * 4);
        if (S == 1) {
            EPI_DISPATCH(epi, epilogue_direct<E>(tmem_base, q, lane, u0, n_valid, ec, m0 + q * 32, n0, stg0));
        } else if (u0 >= n_valid) {
            // nothing for this warp (a tile of one unit)
        } else if (g.defer) {
            // deferred reduction: this split's raw fp32 partial, row-major, for the consumer kernel to sum (glue.h)
            float* dst = g.defer + (size_t)split * g.M * g.N;
            float va[32], vb[32];
            tmem_ld_unit(tmem_base, q, u0, va);
#pragma unroll 1
            for (int u = u0; u < n_valid; u += 4) {
                tmem_ld_wait();
                if (u + 2 < n_valid) tmem_ld_unit(tmem_base, q, u + 2, vb);
                emit_unit_f32(va, stg0, dst, g.N, g.M, g.N, m0 + q * 32, n0 + u * 32, lane);
                if (u + 2 < n_valid) {
                    tmem_ld_wait();
                    if (u + 4 < n_valid) tmem_ld_unit(tmem_base, q, u + 4, va);
                    emit_unit_f32(vb, stg0 + 32 * STG_LD * 4, dst, g.N, g.M, g.N, m0 + q * 32, n0 + (u + 2) * 32, lane);
                }
            }
        } else {
            // park: unit u of (tile, pair rank, split) = 8 column groups x 128 rows x float4, thread-major; the TMEM load of
            // unit u+1 is in flight while unit u is stored
            float4* my_part = reinterpret_cast<float4*>(
                g.partials + ((((size_t)blockIdx.y * (gridDim.x >> 1) + tile) * 2 + pr) * S + split) * (size_t)8 * PAIR_UNIT_FLOATS) + row;
            float va[32], vb[32];
            tmem_ld_unit(tmem_base, q, u0, va);
#pragma unroll 1
            for (int u = u0; u < n_valid; u += 4) {
                tmem_ld_wait();
                if (u + 2 < n_valid) tmem_ld_unit(tmem_base, q, u + 2, vb);
                float4* dst = my_part + (size_t)u * (PAIR_UNIT_FLOATS / 4);
#pragma unroll
                for (int e = 0; e < 8; e++) __stcg(dst + e * 128, make_float4(va[4 * e], va[4 * e + 1], va[4 * e + 2], va[4 * e + 3]));
                if (u + 2 < n_valid) {
                    tmem_ld_wait();
                    if (u + 4 < n_valid) tmem_ld_unit(tmem_base, q, u + 4, va);
                    dst += 2 * (PAIR_UNIT_FLOATS / 4);
#pragma unroll
                    for (int e = 0; e < 8; e++) __stcg(dst + e * 128, make_float4(vb[4 * e], vb[4 * e + 1], vb[4 * e + 2], vb[4 * e + 3]));
                }
            }
            __threadfence();
        }
    }

    if (S > 1 && !g.defer) {
        PPROBE(4);
        // every CTA of the cluster (both pair ranks, all K splits of this tile) has parked its partial
        asm volatile("fence.proxy.async;" ::: "memory");  // generic-proxy global writes above vs the bulk (async-proxy) reads below
        cluster_arrive_wait();
        PPROBE(5);
        // this CTA finishes units u = split, split + S, ... of its own 128 rows: pull the S partials of each into the (idle)
        // pipeline stages with bulk copies, all in flight at once
        const float* base = g.partials + (((size_t)blockIdx.y * (gridDim.x >> 1) + tile) * 2 + pr) * (size_t)S * 8 * PAIR_UNIT_FLOATS;
        float* land = reinterpret_cast<float*>(smem + PAIR_STAGING_BYTES);  // up to 8 units x 16 KiB
        int n_mine = 0;
        for (int u = split; u < n_valid; u += S) n_mine++;
        if (threadIdx.x == 64 && n_mine > 0) {
            asm volatile("fence.proxy.async;" ::: "memory");
            mbar_arrive_expect_tx(red_bar, (uint32_t)n_mine * S * PAIR_UNIT_FLOATS * 4);
            int j = 0;
            for (int u = split; u < n_valid; u += S, j++)
                for (int z = 0; z < S; z++)
                    bulk_load(land + (size_t)(j * S + z) * PAIR_UNIT_FLOATS, base + ((size_t)z * 8 + u) * PAIR_UNIT_FLOATS,
                              PAIR_UNIT_FLOATS * 4, red_bar);
        }
        if (warp >= 2 && warp < 6 && n_mine > 0) {   // four warps: the landing zone takes the rest of the stage memory
            const int q = warp & 3;
            mbar_wait(red_bar, 0);
            EPI_DISPATCH(epi, epilogue_reduce<E>(smem_u32(land), S, split, n_valid, q * 32 + lane, lane, ec, m0 + q * 32, n0,
                                                 smem_u32(smem) + (uint32_t)q * (2 * 32 * STG_LD * 4)));
        }
    }

    PPROBE(6);
    tc_fence_before();
    __syncthreads();
    cluster_arrive_wait();  // neither CTA of a pair may exit (or free TMEM) while the other can still signal / write into it
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, PAIR_TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------------------------------
// Prefill-sized token counts (M > 256: several 256-row M blocks).  Same CTA-pair tile, but PERSISTENT: one pair per SM pair
// walks a list of (M block, N tile) tiles, the fp32 accumulator is double-buffered in TMEM (2 x 256 columns), so the four
// epilogue warps drain tile i while the MMA thread already accumulates tile i+1 and the TMA ring never drains between tiles.
// Tile order: groups of PERSIST_GM consecutive M blocks, N-tile-major inside a group — the pairs of one wave share a few W
// tiles and the group's activation rows stay in L2, so the weights stream from HBM once per group.
constexpr int PERSIST_STAGES = 5;
constexpr int PERSIST_STAGING_BYTES = 8 * 32 * STG_LD * 4;  // one 32 x 32(+4) fp32 buffer per epilogue warp = 36 KiB
constexpr int PERSIST_GM = 8;
constexpr uint32_t PERSIST_TMEM_COLS = 512;
constexpr int PERSIST_SMEM_BYTES = PERSIST_STAGES * PAIR_STAGE_BYTES + PERSIST_STAGING_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;

struct TileCoord {
    int pi, m0, n0, local_tile;
};
DEVI TileCoord persist_tile(const PairGroup& grp, int tile, int NT, int gy) {
    const int per_group = PERSIST_GM * NT;
    const int grp_i = tile / per_group;
    const int r = tile - grp_i * per_group;
    const int gm = min(PERSIST_GM, gy - grp_i * PERSIST_GM);
    const int nt = r / gm;
    const int mb = grp_i * PERSIST_GM + (r - nt * gm);
    int pi = 0;
    while (pi + 1 < grp.n && nt >= grp.p[pi].tile_end) pi++;
    TileCoord c;
    c.pi = pi;
    c.local_tile = nt - (pi ? grp.p[pi - 1].tile_end : 0);
    c.n0 = c.local_tile * grp.p[pi].bn;
    c.m0 = mb * 256;
    return c;
}
// Producer side of a chained tile: wait until the tiles this one reads from have been stored by their epilogue warps
// (generic-proxy stores, released by a fence + atomic), then order the TMA (async-proxy) reads behind that.  The last of the
// `n_consumers` waiters puts both counters back to zero for the next launch.
DEVI void chain_wait(int* flags, int dep_flag, int dep_need, int n_consumers) {
    volatile int* f = flags + dep_flag;
    while (*f < dep_need) __nanosleep(40);
    __threadfence();
    asm volatile("fence.proxy.async;" ::: "memory");
    const int old = atomicAdd(flags + 8 + dep_flag, 1);
    if (old == n_consumers - 1) {
        flags[8 + dep_flag] = 0;
        flags[dep_flag] = 0;
        __threadfence();
    }
}
DEVI void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

__global__ void __launch_bounds__(PERSIST_THREADS, 1) pair_gemm_persist_kernel(const __grid_constant__ PairGroup grp, int gy) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* staging = smem + PERSIST_STAGES * PAIR_STAGE_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + PERSIST_STAGING_BYTES);
    uint64_t* empty_bar = full_bar + PERSIST_STAGES;
    uint64_t* tfull_bar = empty_bar + PERSIST_STAGES;   // [2] accumulator buffer complete (arrives in both CTAs)
    uint64_t* tempty_bar = tfull_bar + 2;               // [2] accumulator buffer drained by all 16 epilogue warps of the pair (leader's)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    const int pr = crank & 1;
    const int pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    const int NT = grp.p[grp.n - 1].tile_end;
    const int total = NT * gy;

    if (threadIdx.x == 0) {
        for (int i = 0; i < grp.n; i++) {
            tma_prefetch_desc(&grp.p[i].tmA);
            tma_prefetch_desc(&grp.p[i].tmW);
        }
        for (int s = 0; s < PERSIST_STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; b++) {
            mbar_init(&tfull_bar[b], 1);
            mbar_init(&tempty_bar[b], 16);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_pair(tmem_slot, PERSIST_TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    cluster_arrive_wait();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            const uint64_t pol_w = gy > 1 ? policy_evict_last() : policy_evict_first();  // W tiles are re-read only across M blocks
            const uint64_t pol_a = policy_evict_last();
            const uint32_t leader = crank & ~1u;
            int s = 0;
            uint32_t ph = 0;
            // the first tile's first ring of weight tiles does not depend on the preceding kernel: request it before the wait
            int npre = 0;
            if (pair_id < total) {
                const TileCoord c = persist_tile(grp, pair_id, NT, gy);
                const PairProblem& P = grp.p[c.pi];
                const int bn = P.bn;
                const int n_kb = P.g.kb_per_split;
                const uint32_t tx = 2u * (128u * BK * 2u + (uint32_t)(bn >> 1) * BK * 2u);
                npre = min(n_kb, PERSIST_STAGES);
                for (int kb = 0; kb < npre; kb++) {
                    if (pr == 0) mbar_arrive_expect_tx(&full_bar[kb], tx);
                    tma_load_2d_pair(smem + kb * PAIR_STAGE_BYTES + 128 * BK * 2, &P.tmW, (P.kb0 + kb) * BK, c.n0 + pr * (bn >> 1),
                                     mapa_shared(smem_u32(&full_bar[kb]), leader), pol_w);
                }
                pdl_wait();
                if (P.dep_flag >= 0) chain_wait(grp.flags, P.dep_flag, P.dep_need, 2 * (P.tile_end - (c.pi ? grp.p[c.pi - 1].tile_end : 0)));
                for (int kb = 0; kb < npre; kb++)
                    tma_load_2d_pair(smem + kb * PAIR_STAGE_BYTES, &P.tmA, (P.kb0 + kb) * BK, c.m0 + pr * 128,
                                     mapa_shared(smem_u32(&full_bar[kb]), leader), pol_a);
                s = npre % PERSIST_STAGES;
                ph = (uint32_t)(npre / PERSIST_STAGES) & 1;
            } else {
                pdl_wait();
            }
            for (int tile = pair_id; tile < total; tile += n_pairs) {
                const TileCoord c = persist_tile(grp, tile, NT, gy);
                const PairProblem& P = grp.p[c.pi];
                const int bn = P.bn;
                const int n_kb = P.g.kb_per_split;
                const uint32_t tx = 2u * (128u * BK * 2u + (uint32_t)(bn >> 1) * BK * 2u);
                for (int kb = (tile == pair_id ? npre : 0); kb < n_kb; kb++) {
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t* st = smem + s * PAIR_STAGE_BYTES;
                    if (pr == 0) mbar_arrive_expect_tx(&full_bar[s], tx);
                    const uint32_t bar = mapa_shared(smem_u32(&full_bar[s]), leader);
                    tma_load_2d_pair(st + 128 * BK * 2, &P.tmW, (P.kb0 + kb) * BK, c.n0 + pr * (bn >> 1), bar, pol_w);
                    if (kb == 0 && P.dep_flag >= 0)  // the weights above are already on their way
                        chain_wait(grp.flags, P.dep_flag, P.dep_need, 2 * (P.tile_end - (c.pi ? grp.p[c.pi - 1].tile_end : 0)));
                    tma_load_2d_pair(st, &P.tmA, (P.kb0 + kb) * BK, c.m0 + pr * 128, bar, pol_a);
                    if (++s == PERSIST_STAGES) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            }
        } else {
            pdl_wait();
        }
        __syncwarp();
    } else if (warp == 1) {
        pdl_wait();
        if (pr == 0 && lane == 0) {
            const uint16_t mask = (uint16_t)(3u << (crank & ~1u));
            int s = 0;
            uint32_t ph = 0;
            int j = 0;
            for (int tile = pair_id; tile < total; tile += n_pairs, j++) {
                const TileCoord c = persist_tile(grp, tile, NT, gy);
                const PairProblem& P = grp.p[c.pi];
                const int n_kb = P.g.kb_per_split;
                const uint32_t idesc = umma_idesc_f16(256, (uint32_t)P.bn);
                const int buf = j & 1;
                mbar_wait(&tempty_bar[buf], (uint32_t)((j >> 1) & 1) ^ 1);  // the epilogue of tile j-2 has drained this buffer
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)buf * 256u;
                for (int kb = 0; kb < n_kb; kb++) {
                    mbar_wait(&full_bar[s], ph);
                    tc_fence_after();
                    const uint32_t st = smem_u32(smem + s * PAIR_STAGE_BYTES);
                    const uint64_t adesc = umma_desc_sw128(st);
                    const uint64_t bdesc = umma_desc_sw128(st + 128 * BK * 2);
#pragma unroll
                    for (int k = 0; k < BK / 16; k++)
                        umma_f16_ss_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit_pair(&empty_bar[s], mask);
                    if (++s == PERSIST_STAGES) {
                        s = 0;
                        ph ^= 1;
                    }
                }
                umma_commit_pair(&tfull_bar[buf], mask);
            }
        }
        __syncwarp();
    } else {
        pdl_wait();
        const int q = warp & 3;
        const int u0 = (warp - 2) >> 2;
        const uint32_t stg0 = smem_u32(staging) + (uint32_t)(warp - 2) * (32 * STG_LD * 4);
        const uint32_t tempty_leader0 = mapa_shared(smem_u32(&tempty_bar[0]), crank & ~1u);
        int j = 0;
        for (int tile = pair_id; tile < total; tile += n_pairs, j++) {
            const TileCoord c = persist_tile(grp, tile, NT, gy);
            const PairProblem& P = grp.p[c.pi];
            const GemmArgs& g = P.g;
            EpiCtx ec;
            ec.out = g.out; ec.bias = g.bias; ec.residual = g.residual; ec.ldo = g.ldo; ec.M = g.M; ec.N = g.N;
            int n_valid = (g.N - c.n0 + 31) >> 5;
            if (n_valid > (P.bn >> 5)) n_valid = P.bn >> 5;
            const int buf = j & 1;
            mbar_wait(&tfull_bar[buf], (uint32_t)(j >> 1) & 1);
            tc_fence_after();
            const int epi = g.epi;
            if (g.defer) {
                epilogue_f32_1buf(tmem_base + (uint32_t)buf * 256u, q, lane, u0, n_valid, g.defer, g.M, g.N, c.m0 + pr * 128 + q * 32, c.n0, stg0);
            } else {
                EPI_DISPATCH(epi, epilogue_direct_1buf<E>(tmem_base + (uint32_t)buf * 256u, q, lane, u0, n_valid, ec, c.m0 + pr * 128 + q * 32, c.n0, stg0));
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(tempty_leader0 + (uint32_t)buf * 8u);
            if (P.sig_flag >= 0) {  // chained launch: this warp's rows of the tile are stored
                __threadfence();
                __syncwarp();
                if (lane == 0) atomicAdd(grp.flags + P.sig_flag + c.local_tile / P.sig_div, 1);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_arrive_wait();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, PERSIST_TMEM_COLS);
    }
}

constexpr int PAIR_SMEM_BYTES = PAIR_MAX_STAGES * PAIR_STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
static_assert(PAIR_STAGING_BYTES + 8 * PAIR_UNIT_FLOATS * 4 <= PAIR_MAX_STAGES * PAIR_STAGE_BYTES, "landing zone must fit the stages");
static_assert(PAIR_STAGING_BYTES_8 <= PAIR_MAX_STAGES * PAIR_STAGE_BYTES, "eight staging pairs must fit the stages");

// Pairs of CTAs (clusters of 2 x 1 x s) the device can hold at once.
int pair_slots(int s) {
    static int cache[8][PAIR_MAX_SPLIT + 1] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 8) dev = 0;
    if (cache[dev][s]) return cache[dev][s];
    cudaFuncSetAttribute(pair_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PAIR_SMEM_BYTES);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2, 1, s);
    cfg.blockDim = dim3(PAIR_THREADS);
    cfg.dynamicSmemBytes = PAIR_SMEM_BYTES;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = s;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, pair_gemm_kernel, &cfg) != cudaSuccess || n <= 0) {
        cudaGetLastError();
        n = 148 / (2 * s) - (s > 1 ? 2 : 0);  // conservative guess
    }
    cache[dev][s] = n;
    return n;
}

// CTA pairs of the persistent kernel the device can hold at once (one per SM pair).
int persist_slots() {
    static int cache[8] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 8) dev = 0;
    if (cache[dev]) return cache[dev];
    cudaFuncSetAttribute(pair_gemm_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PERSIST_SMEM_BYTES);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2, 1, 1);
    cfg.blockDim = dim3(PERSIST_THREADS);
    cfg.dynamicSmemBytes = PERSIST_SMEM_BYTES;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, pair_gemm_persist_kernel, &cfg) != cudaSuccess || n <= 0) {
        cudaGetLastError();
        n = 72;
    }
    cache[dev] = n;
    return n;
}

// Modelled time (us) of one CTA of a (bn, s) plan with kps K-blocks per split (rates: DESIGN.md §3, profiles/r01_gemm_timeline*).
double pair_cta_cost(int bn, int s, int kps, bool defer = false) {
    const double clk = 1500.0;                                    // MHz under tensor load (power-capped)
    const double mma = 4.0 * (bn / 2.0);                          // cycles per K block: 4 UMMAs of 256 x bn x 16
    const double ingest = (16384.0 + bn * 64.0) / 52.0;           // bytes into this SM per K block at ~52 B/clk
    // (a flat 560-cycle floor per K block — what the timelines show for BN = 64 ... 256 — was tried: widest tiles everywhere,
    // 6.480 -> 6.468 ms/step, i.e. nothing, so the plans that the parity suite has been run on all along are kept)
    const double main_us = kps * (mma > ingest ? mma : ingest) / clk;
    const double epi_us = defer ? 0.5 + 3.0 * bn / 256.0 : (s > 1) ? 1.2 + 4.5 * bn / 256.0 : 0.5 + 2.5 * bn / 256.0;
    return 1.8 + main_us + epi_us;
}

}  // namespace

int pair_slots_query(int s) { return (s >= 1 && s <= PAIR_MAX_SPLIT) ? pair_slots(s) : 0; }

// Prefill-sized M: persistent pairs, double-buffered accumulators (no K split, BN = 256).
static int persist_group_launch(int n, const LinearDesc* d, int M, int gy, cudaStream_t stream, bool pdl) {
    PairGroup grp;
    grp.n = n;
    grp.n_stages = PERSIST_STAGES;
    int end = 0;
    for (int i = 0; i < n; i++) {
        int bn = 256;
        const int n32 = (d[i].N + 31) / 32 * 32;
        if (bn > n32) bn = n32;
        PairProblem& P = grp.p[i];
        int rc = make_tmap_2d_f16(&P.tmA, d[i].A, (uint64_t)M, (uint64_t)d[i].K, (uint64_t)d[i].lda * 2, 128, BK,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
        if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
        rc = make_tmap_2d_f16(&P.tmW, d[i].W, (uint64_t)d[i].N, (uint64_t)d[i].K, (uint64_t)d[i].ldw * 2, bn / 2, BK,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
        if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
        GemmArgs& g = P.g;
        g.M = M; g.N = d[i].N; g.K = d[i].K;
        g.out = static_cast<__half*>(d[i].out);
        g.ldo = d[i].ldo;
        g.bias = static_cast<const __half*>(d[i].bias);
        g.residual = static_cast<const __half*>(d[i].residual);
        g.epi = d[i].epi;
        if (g.epi < B200_EPI_NONE || g.epi > B200_EPI_RESADD || (g.epi == B200_EPI_RESADD && !g.residual)) return B200_ERR_BAD_ARG;
        g.partials = nullptr; g.kb_per_split = (d[i].K + BK - 1) / BK; g.n_split = 1; g.dbg = nullptr; g.defer = nullptr;
        end += (d[i].N + bn - 1) / bn;
        P.tile_end = end;
        P.bn = bn;
        P.kb0 = 0; P.dep_flag = -1; P.dep_need = 0; P.sig_flag = -1; P.sig_div = 1;
    }
    grp.flags = nullptr;
    for (int i = n; i < PAIR_MAX_GROUP; i++) {
        grp.p[i].tile_end = end;
        grp.p[i].bn = 32;
    }
    const long total = (long)end * gy;
    const int slots = persist_slots();
    const int pairs = (int)(total < slots ? total : slots);
    cudaError_t e = cudaFuncSetAttribute(pair_gemm_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PERSIST_SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pairs, 1, 1);
    cfg.blockDim = dim3(PERSIST_THREADS);
    cfg.dynamicSmemBytes = PERSIST_SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 2 : 1;
    return (int)cudaLaunchKernelEx(&cfg, pair_gemm_persist_kernel, grp, gy);
}

// FFN of one decode step in ONE persistent launch (Albatross/rwkv7.py:678-679): the key GEMM (relu^2, -> hidden) and the value
// GEMM, whose K is split S ways over independent tiles that leave raw fp32 partials for the consumer (glue.h: deferred
// reduction).  A value tile of split z starts its activation loads once the F/S hidden columns it reads have been stored
// (flags[z]); its weights are requested before that.  Pairs that hold no key tile take value tiles first and simply wait.
// Same tiles, same K ranges, same summation order as the two separate launches: bit-identical results.
int ffn_chain_launch(int M, int C, int F, const void* kx, const void* Wk, void* hidden, const void* Wv, float* partials,
                     long partials_bytes, int* flags, int* n_split, cudaStream_t stream, bool pdl) {
    *n_split = 0;
    const int S = 4;
    if (M <= 128 || M > 256 || !flags || !partials) return B200_ERR_UNSUPPORTED;
    if ((F % (S * 256)) || (C & 7) || F / S / BK < 2 || (long)S * M * C * 4 > partials_bytes) return B200_ERR_UNSUPPORTED;
    // Opt-in (read per call): measured 0.22 ms/step SLOWER than the two launches (6.63 vs 6.41 ms) although it saves a launch
    // boundary per layer — the persistent kernel's 254 registers keep the next kernel's CTAs off the SMs until it exits, and a
    // 168-register build spills.  Kept, with its bit-exactness test, as the starting point for fusing the step further.
    if (getenv("B200_CHAIN") == nullptr) return B200_ERR_UNSUPPORTED;
    PairGroup grp;
    grp.n = 1 + S;
    grp.n_stages = PERSIST_STAGES;
    grp.flags = flags;
    int end = 0;
    for (int i = 0; i < 1 + S; i++) {
        PairProblem& P = grp.p[i];
        GemmArgs& g = P.g;
        const bool key = i == 0;
        const int N = key ? F : C, K = key ? C : F;
        int bn = 256;
        const int n32 = (N + 31) / 32 * 32;
        if (bn > n32) bn = n32;
        int rc = make_tmap_2d_f16(&P.tmA, key ? kx : hidden, (uint64_t)M, (uint64_t)K, (uint64_t)K * 2, 128, BK,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
        if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
        rc = make_tmap_2d_f16(&P.tmW, key ? Wk : Wv, (uint64_t)N, (uint64_t)K, (uint64_t)K * 2, bn / 2, BK,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
        if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
        g.M = M; g.N = N; g.K = K;
        g.out = key ? static_cast<__half*>(hidden) : nullptr;
        g.ldo = N;
        g.bias = nullptr; g.residual = nullptr;
        g.epi = key ? B200_EPI_RELU_SQ : B200_EPI_NONE;
        g.partials = nullptr; g.n_split = 1; g.dbg = nullptr;
        g.kb_per_split = key ? (C + BK - 1) / BK : F / S / BK;
        g.defer = key ? nullptr : partials + (size_t)(i - 1) * M * C;
        end += (N + bn - 1) / bn;
        P.tile_end = end;
        P.bn = bn;
        P.kb0 = key ? 0 : (i - 1) * (F / S / BK);
        P.dep_flag = key ? -1 : i - 1;
        P.dep_need = key ? 0 : (F / S / 256) * 16;  // every key tile of the split's hidden columns, 16 epilogue warps each
        P.sig_flag = key ? 0 : -1;
        P.sig_div = key ? F / S / 256 : 1;
    }
    for (int i = 1 + S; i < PAIR_MAX_GROUP; i++) {
        grp.p[i].tile_end = end;
        grp.p[i].bn = 32;
        grp.p[i].kb0 = 0; grp.p[i].dep_flag = -1; grp.p[i].dep_need = 0; grp.p[i].sig_flag = -1; grp.p[i].sig_div = 1;
    }
    const int slots = persist_slots();
    // every pair must be resident at once (value tiles wait on key tiles of other pairs), and the pairs beyond the key tiles
    // must not be needed for key tiles: one launch of at most `slots` pairs, each walking its tile list in order
    const int pairs = end < slots ? end : slots;
    cudaError_t e = cudaFuncSetAttribute(pair_gemm_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PERSIST_SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pairs, 1, 1);
    cfg.blockDim = dim3(PERSIST_THREADS);
    cfg.dynamicSmemBytes = PERSIST_SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 2 : 1;
    const int rc = (int)cudaLaunchKernelEx(&cfg, pair_gemm_persist_kernel, grp, 1);
    if (rc == 0) *n_split = S;
    return rc;
}

// Up to PAIR_MAX_GROUP linear layers that share M (> 128) in ONE launch.  Returns B200_ERR_UNSUPPORTED when the group does
// not fit this kernel (the caller then uses gemm.cu).
int pair_group_launch(int n, const LinearDesc* d, int M, void* workspace, long workspace_bytes, cudaStream_t stream, bool pdl,
                      unsigned long long* dbg, int* defer_splits) {
    if (defer_splits) *defer_splits = 0;
    if (n <= 0 || n > PAIR_MAX_GROUP || M <= 128) return B200_ERR_UNSUPPORTED;
    const int gy = (M + 255) / 256;
    int min_kb = 1 << 30;
    for (int i = 0; i < n; i++) {
        if ((d[i].K & 7) || (d[i].N & 7) || (d[i].lda & 7) || (d[i].ldw & 7) || (d[i].ldo & 7)) return B200_ERR_BAD_ARG;
        // the epilogue moves 8 fp16 (16 bytes) per access
        if ((reinterpret_cast<uintptr_t>(d[i].out) | reinterpret_cast<uintptr_t>(d[i].bias) | reinterpret_cast<uintptr_t>(d[i].residual)) & 15)
            return B200_ERR_UNSUPPORTED;
        const int kb = (d[i].K + BK - 1) / BK;
        if (kb < min_kb) min_kb = kb;
    }
    // tile widths and the common K split: minimise waves x (time of the slowest CTA)
    static const int force_bn = getenv("B200_PAIR_BN") ? atoi(getenv("B200_PAIR_BN")) : 0;  // tuning overrides
    static const int force_s = getenv("B200_PAIR_S") ? atoi(getenv("B200_PAIR_S")) : 0;
    const bool fixed256 = force_bn != 0;
    int best_s = 1, best_bn[PAIR_MAX_GROUP];
    double best_cost = 1e30;
    // 224: FFN key (N = 16384) then has 74 tiles = one per CTA pair, each 7/8 as wide, instead of 64 tiles on 64 of 74 pairs
    const int cands[5] = {force_bn ? force_bn : 256, 224, 192, 128, 64};
    for (int s = 1; s <= PAIR_MAX_SPLIT; s++) {
        if (s > 1 && s > min_kb / 2) break;
        if (force_s && s != force_s && force_s <= min_kb / 2) continue;
        bool ok = true;
        for (int i = 0; i < n && ok; i++) {
            const int kb = (d[i].K + BK - 1) / BK, kps = (kb + s - 1) / s;
            if ((kb + kps - 1) / kps != s) ok = false;  // no empty split
        }
        if (!ok) continue;
        const int slots = pair_slots(s);
        for (int ci = 0; ci < (fixed256 ? 1 : 5); ci++) {
            int bns[PAIR_MAX_GROUP];
            long tiles = 0;
            double worst = 0.0;
            for (int i = 0; i < n; i++) {
                int bn = cands[ci];
                const int n32 = (d[i].N + 31) / 32 * 32;
                if (bn > n32) bn = n32;
                bns[i] = bn;
                tiles += (d[i].N + bn - 1) / bn;
                const int kb = (d[i].K + BK - 1) / BK;
                const double c = pair_cta_cost(bn, s, (kb + s - 1) / s);
                if (c > worst) worst = c;
            }
            tiles *= gy;
            if (s > 1 && (workspace == nullptr || tiles * 2 * s * 8 * PAIR_UNIT_FLOATS * 4 > workspace_bytes)) continue;
            const long waves = (tiles + slots - 1) / slots;
            const double cost = waves * worst;
            if (cost < best_cost) {
                best_cost = cost;
                best_s = s;
                for (int i = 0; i < n; i++) best_bn[i] = bns[i];
            }
        }
    }
    // Several M blocks: the persistent kernel wins once every pair has a few tiles to walk (its accumulator double-buffering and
    // never-draining ring pay per tile boundary); with few tiles and long K the one-tile kernel's K split is worth more.
    static const bool persist_env = getenv("B200_NO_PERSIST") == nullptr;
    long tiles256 = 0;
    for (int i = 0; i < n; i++) tiles256 += (long)((d[i].N + 255) / 256) * gy;
    static const bool persist_all = getenv("B200_PERSIST_ALL") != nullptr;  // experiment
    if (persist_all && persist_env && !defer_splits) return persist_group_launch(n, d, M, gy, stream, pdl);
    if ((gy >= 2 || (n == 1 && tiles256 >= 3L * persist_slots())) && persist_env) {  // also one wide problem (the head: 256 tiles)
        long tiles = 0;
        double work = 0.0;  // tile x K-block iterations at BN = 256
        for (int i = 0; i < n; i++) {
            const long t = (long)((d[i].N + 255) / 256) * gy;
            tiles += t;
            work += (double)t * ((d[i].K + BK - 1) / BK);
        }
        const int slots = persist_slots();
        const long pairs = tiles < slots ? tiles : slots;
        const double rounds = (double)((tiles + pairs - 1) / pairs);
        const double persist_cost = 5.0 + rounds * (work / tiles) * 0.40;
        if (persist_cost < best_cost) return persist_group_launch(n, d, M, gy, stream, pdl);
    }
    // Deferred reduction (one residual-add problem, consumer sums the partials): the K split is then over independent CTA
    // pairs — no cluster along z, so S is not limited by the 15 clusters of 8 the device co-schedules (FFN value: 16 tiles x 4).
    bool defer = false;
    static const bool defer_env = getenv("B200_NO_DEFER") == nullptr;
    if (defer_splits && defer_env && n == 1 && gy == 1 && d[0].epi == B200_EPI_RESADD && !force_bn && !force_s && workspace) {
        const int kb = (d[0].K + BK - 1) / BK;
        const int slots = pair_slots(1);
        const int dc[3] = {256, 192, 128};
        static const int min_kb_env = getenv("B200_DEFER_MIN_KB") ? atoi(getenv("B200_DEFER_MIN_KB")) : 0;  // tuning overrides
        static const int force_ds = getenv("B200_DEFER_S") ? atoi(getenv("B200_DEFER_S")) : 0;
        for (int s = 2; s <= 8 && s <= kb / 2 && kb >= min_kb_env; s++) {
            if (force_ds && s != force_ds) continue;
            const int kps = (kb + s - 1) / s;
            if ((kb + kps - 1) / kps != s) continue;
            if ((long)s * M * d[0].N * 4 > workspace_bytes) break;
            for (int ci = 0; ci < 3; ci++) {
                int bn = dc[ci];
                const int n32 = (d[0].N + 31) / 32 * 32;
                if (bn > n32) bn = n32;
                const long tiles = (long)((d[0].N + bn - 1) / bn) * s;
                const long waves = (tiles + slots - 1) / slots;
                const double cost = waves * pair_cta_cost(bn, s, kps, true) + 2.0;  // + the consumer's share
                if (cost < best_cost) {
                    best_cost = cost;
                    best_s = s;
                    best_bn[0] = bn;
                    defer = true;
                }
            }
        }
    }
    const int s = best_s;
    PairGroup grp;
    grp.n = n;
    static const int stages_env = getenv("B200_PAIR_STAGES") ? atoi(getenv("B200_PAIR_STAGES")) : PAIR_MAX_STAGES;
    grp.n_stages = (s == 1 && stages_env >= 2 && stages_env <= PAIR_MAX_STAGES) ? stages_env : PAIR_MAX_STAGES;
    int end = 0;
    for (int i = 0; i < n; i++) {
        const int bn = best_bn[i];
        PairProblem& P = grp.p[i];
        int rc = make_tmap_2d_f16(&P.tmA, d[i].A, (uint64_t)M, (uint64_t)d[i].K, (uint64_t)d[i].lda * 2, 128, BK,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
        if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
        rc = make_tmap_2d_f16(&P.tmW, d[i].W, (uint64_t)d[i].N, (uint64_t)d[i].K, (uint64_t)d[i].ldw * 2, bn / 2, BK,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
        if (rc) return rc == -1 ? B200_ERR_NO_DRIVER : rc;
        GemmArgs& g = P.g;
        g.M = M; g.N = d[i].N; g.K = d[i].K;
        g.out = static_cast<__half*>(d[i].out);
        g.ldo = d[i].ldo;
        g.bias = static_cast<const __half*>(d[i].bias);
        g.residual = static_cast<const __half*>(d[i].residual);
        g.epi = d[i].epi;
        if (g.epi < B200_EPI_NONE || g.epi > B200_EPI_RESADD || (g.epi == B200_EPI_RESADD && !g.residual)) return B200_ERR_BAD_ARG;
        g.partials = static_cast<float*>(workspace);
        const int kb = (d[i].K + BK - 1) / BK;
        g.kb_per_split = (kb + s - 1) / s;
        g.n_split = s;
        g.dbg = dbg;
        g.defer = defer ? static_cast<float*>(workspace) : nullptr;
        end += (d[i].N + bn - 1) / bn;
        P.tile_end = end;
        P.bn = bn;
    }
    for (int i = n; i < PAIR_MAX_GROUP; i++) {
        grp.p[i].tile_end = end;
        grp.p[i].bn = 32;
    }
    if (s > 1 && !defer) {
        const long need = (long)end * gy * 2 * s * 8 * PAIR_UNIT_FLOATS * 4;
        if (workspace == nullptr || workspace_bytes < need) return B200_ERR_WORKSPACE;
    }
    cudaError_t e = cudaFuncSetAttribute(pair_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PAIR_SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * end, gy, s);
    cfg.blockDim = dim3(PAIR_THREADS);
    cfg.dynamicSmemBytes = PAIR_SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = defer ? 1 : s;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 2 : 1;
    const int rc = (int)cudaLaunchKernelEx(&cfg, pair_gemm_kernel, grp);
    if (rc == 0 && defer) *defer_splits = s;
    return rc;
}

}  // namespace b200

// Fabric probe: how many bytes per second can TMA land in the SMs' shared memory when part of the stream is the SAME tile for
// every CTA (the activation operand of a decode GEMM: 256 x K, re-read by every N tile) and part is unique (the weight stream)?
// And does .multicast::cluster of the shared part inside a cluster of c CTAs lift that ceiling?  No MMA here: a producer thread
// issues the loads into a ring of stages and a consumer thread frees them, i.e. the pair GEMM's mainloop with infinitely fast
// tensor cores.  Decides whether gemm_pair.cu should multicast A across CTA pairs.
//
// build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I chirrup_b200/csrc -o tools/fabric_probe tools/fabric_probe.cu
// run:    tools/fabric_probe            (prints one line per configuration)
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "common.cuh"
#include "tmap.h"

using namespace b200;

constexpr int STAGES = 6;
constexpr int STAGE_BYTES = 32 * 1024;  // 16 KiB shared tile + 16 KiB private tile
constexpr int BK = 64;

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
DEVI void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
DEVI void mbar_arrive_remote(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
DEVI void tma_load_2d_mc(void* dst, const CUtensorMap* m, int c0, int c1, uint64_t* bar, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%2, %3}], "
        "[%4], %5;" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(c0), "r"(c1), "r"(smem_u32(bar)), "h"(mask)
        : "memory");
}

// mode bit 0: multicast the shared tile inside the cluster (each CTA loads 128/cs of its rows and broadcasts them)
// mode bit 1: skip the private (weight) tile;  mode bit 2: skip the shared tile
__global__ void __launch_bounds__(64, 1) probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmAs,
                                                      const __grid_constant__ CUtensorMap tmW, int n_kb, int cs, int mode,
                                                      int w_rows_total) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    const bool mc = (mode & 1) && cs > 1;
    const bool do_w = !(mode & 2), do_a = !(mode & 4);
    const uint32_t rank = cluster_ctarank();
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], mc ? cs : 1);
        }
        fence_barrier_init();
    }
    __syncthreads();
    cluster_sync_all();
    const uint32_t tx = (do_a ? 128u * BK * 2u : 0u) + (do_w ? 128u * BK * 2u : 0u);
    if (threadIdx.x == 0) {
        const uint64_t pol_w = policy_evict_first(), pol_a = policy_evict_last();
        const int rows_a = 128 / cs;
        int w_row = (int)(((long)blockIdx.x * 128) % w_rows_total);
        for (int it = 0; it < n_kb; it++) {
            const int s = it % STAGES;
            const uint32_t ph = (it / STAGES) & 1;
            mbar_wait(&empty_bar[s], ph ^ 1);
            uint8_t* st = smem + s * STAGE_BYTES;
            mbar_arrive_expect_tx(&full_bar[s], tx);
            if (do_w) tma_load_2d_hint(st + 16384, &tmW, it * BK, w_row, &full_bar[s], pol_w);
            if (do_a) {
                if (mc)
                    tma_load_2d_mc(st + rank * rows_a * 128, &tmAs, it * BK, (int)rank * rows_a, &full_bar[s], (uint16_t)((1u << cs) - 1));
                else
                    tma_load_2d_hint(st, &tmA, it * BK, 0, &full_bar[s], pol_a);
            }
        }
    } else if (threadIdx.x == 32) {
        for (int it = 0; it < n_kb; it++) {
            const int s = it % STAGES;
            const uint32_t ph = (it / STAGES) & 1;
            mbar_wait(&full_bar[s], ph);
            if (mc) {
                for (int r = 0; r < cs; r++) mbar_arrive_remote(mapa_shared(smem_u32(&empty_bar[s]), r));
            } else {
                mbar_arrive(&empty_bar[s]);
            }
        }
    }
    __syncthreads();
    cluster_sync_all();  // no CTA leaves while a peer may still multicast into it or arrive on its barriers
}

#define CK(x)                                                                  \
    do {                                                                       \
        cudaError_t e_ = (x);                                                  \
        if (e_ != cudaSuccess) {                                               \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(1);                                                           \
        }                                                                      \
    } while (0)

int main(int argc, char** argv) {
    const int K = 16384, n_kb = K / BK;
    const int w_rows = 148 * 128;  // 148 private 128-row streams: 592 MiB > L2
    __half *A, *W;
    CK(cudaMalloc(&A, (size_t)128 * K * 2));
    CK(cudaMalloc(&W, (size_t)w_rows * K * 2));
    CK(cudaMemset(A, 0, (size_t)128 * K * 2));
    CK(cudaMemset(W, 0, (size_t)w_rows * K * 2));
    const int smem_bytes = STAGES * STAGE_BYTES + 1024 + 256;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int grids[] = {64, 128, 144};
    const int css[] = {1, 2, 4, 8, 16};
    for (int mode : {0, 1, 2, 3, 4}) {
        for (int cs : css) {
            if (!(mode & 1) && cs > 2) continue;      // unicast: cluster size is irrelevant beyond a sanity pair
            if ((mode & 4) && cs > 1) continue;
            for (int grid : grids) {
                if (grid % cs) continue;
                CUtensorMap tmA, tmAs, tmW;
                if (make_tmap_2d_f16(&tmA, A, 128, K, (uint64_t)K * 2, 128, BK) ||
                    make_tmap_2d_f16(&tmAs, A, 128, K, (uint64_t)K * 2, 128 / cs, BK) ||
                    make_tmap_2d_f16(&tmW, W, w_rows, K, (uint64_t)K * 2, 128, BK, CU_TENSOR_MAP_L2_PROMOTION_L2_256B)) {
                    printf("tensor map failed\n");
                    return 1;
                }
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(grid);
                cfg.blockDim = dim3(64);
                cfg.dynamicSmemBytes = smem_bytes;
                cudaLaunchAttribute attr[1];
                attr[0].id = cudaLaunchAttributeClusterDimension;
                attr[0].val.clusterDim.x = cs;
                attr[0].val.clusterDim.y = 1;
                attr[0].val.clusterDim.z = 1;
                cfg.attrs = attr;
                cfg.numAttrs = 1;
                const int reps = 10;
                bool ok = true;
                for (int r = 0; r < 3 && ok; r++) ok = cudaLaunchKernelEx(&cfg, probe_kernel, tmA, tmAs, tmW, n_kb, cs, mode, w_rows) == cudaSuccess;
                if (!ok) {
                    printf("mode %d cs %2d grid %3d: launch failed (%s)\n", mode, cs, grid, cudaGetErrorString(cudaGetLastError()));
                    continue;
                }
                CK(cudaDeviceSynchronize());
                CK(cudaEventRecord(e0));
                for (int r = 0; r < reps; r++) cudaLaunchKernelEx(&cfg, probe_kernel, tmA, tmAs, tmW, n_kb, cs, mode, w_rows);
                CK(cudaEventRecord(e1));
                CK(cudaDeviceSynchronize());
                float ms = 0;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                const double us = ms * 1000.0 / reps;
                const double landed = (double)grid * n_kb * (((mode & 4) ? 0 : 16384.0) + ((mode & 2) ? 0 : 16384.0));
                const double priv = (mode & 2) ? 0.0 : (double)grid * n_kb * 16384.0;
                printf("mode %d (%s%s%s) cs %2d grid %3d: %8.1f us  landed %6.2f TB/s  private(HBM) %5.2f TB/s\n", mode,
                       (mode & 1) ? "multicast" : "unicast", (mode & 2) ? ", shared tile only" : "", (mode & 4) ? ", private tile only" : "",
                       cs, grid, us, landed / us * 1e-6, priv / us * 1e-6);
            }
        }
    }
    return 0;
}

// b200_linear_f16_static: the dense linear layer of gemm.cu / gemm_pair.cu for callers that cannot allocate device memory
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The split-K GEMM plans park fp32 partial tiles in a scratch buffer the caller of b200_linear_f16 has to provide.  Here that
// scratch is a file-scope __device__ array: B200_LINEAR_STATIC_SCRATCH_BYTES (32 MiB) of every GPU that loads the library,
// held until the process exits.  32 MiB covers every single-wave split plan (148 CTAs x 8 units x 18 KiB = 20.8 MiB); a plan
// that would need more is not chosen (CTA-pair planner) or the call fails with B200_ERR_WORKSPACE — never an allocation.
//
// In host code the array's NAME is a host-side stand-in, not a device address: the address each device holds it at is taken
// once per device with cudaGetSymbolAddress (that device current) and that pointer is what the launcher receives.
//
// Nothing in the scratch has to start at zero: every partial is written by the launch that reads it (ordered by the cluster
// barrier), so the previous call's contents are simply overwritten.  All calls share the one array: calls that can overlap on
// the GPU (different streams, or graphs replayed concurrently) race on it — serialise them on one stream.
#include <cuda_runtime.h>

#include "glue.h"
#include "../../include/chirrup_b200.h"

namespace b200 {
namespace {

static __device__ __align__(256) unsigned char g_linear_static_scratch[B200_LINEAR_STATIC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
void* g_scratch_ptr[MAX_DEVICES] = {};   // device address of the array on each device, resolved on first use

// Device address of the scratch array on the CURRENT device (cudaGetSymbolAddress resolves the symbol for the current device).
int static_scratch(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= MAX_DEVICES) return B200_ERR_UNSUPPORTED;
    if (g_scratch_ptr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_linear_static_scratch);
        if (e != cudaSuccess) return (int)e;
        if (p == nullptr) return B200_ERR_UNSUPPORTED;
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return 0;
}

}  // namespace
}  // namespace b200

extern "C" long b200_linear_static_scratch_bytes(void) { return B200_LINEAR_STATIC_SCRATCH_BYTES; }

extern "C" int b200_linear_f16_static(int M, int N, int K, const void* A, long lda, const void* W, long ldw, void* out, long ldo,
                                      const void* bias, const void* residual, int epilogue, void* stream) {
    // stated maxima (include/chirrup_b200.h): beyond them the call is refused, whatever the plan would have needed
    if (M > B200_LINEAR_STATIC_MAX_M || N > B200_LINEAR_STATIC_MAX_N || K > B200_LINEAR_STATIC_MAX_K) return B200_ERR_TOO_LARGE;
    void* scratch = nullptr;
    const int rc = b200::static_scratch(&scratch);
    if (rc) return rc;
    return b200::linear_f16_launch(M, N, K, A, lda, W, ldw, out, ldo, bias, residual, epilogue, scratch,
                                   (long)B200_LINEAR_STATIC_SCRATCH_BYTES, static_cast<cudaStream_t>(stream), /*pdl=*/false);
}

// noalloc.cu -- somar_norm_noalloc: the norm reduction (k_reduce, util.cu -- CH_XD::norm over valid cells,
// MappedAMRPoissonOp.H:698-702) as an entry point that ALLOCATES NOTHING, for callers that cannot call cudaMalloc
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid allocation).  It has no counterpart in
// the reference.
//
// k_reduce keeps its per-block partial results and its "last block folds" ticket in scratch memory that
// somar_level_create allocates per level.  This variant keeps them in a file-scope `static __device__` array instead:
//   * size: NOALLOC_MAX_BLOCKS partials + one ticket line = 128 KiB + 256 B, held by every GPU that loads the
//     library until the process exits;
//   * limit: a block reduces NOALLOC_CELLS_PER_BLOCK = 4096 doubles, so n <= 16384 * 4096 = 67 108 864 doubles
//     (SOMAR_NOALLOC_MAX_N in the header); larger n is rejected with SOMAR_ERR_ARG -- never a fallback allocation;
//   * in host code the array's NAME is a host-side stand-in, not a device address: the device address is taken ONCE
//     PER DEVICE, with that device current, by cudaGetSymbolAddress, and that pointer is what the kernel gets;
//   * every call shares the array.  It keeps the last call's contents: the ticket is cleared on the caller's stream
//     before each launch, the partials are fully rewritten by the blocks of the launch that reads them.  Calls that
//     OVERLAP on one GPU (different streams, or several host threads) race on it -- serialise them.
// The entry point calls no cudaMalloc / cudaFree / allocating library routine: one cudaMemsetAsync, one launch.
#include "../../include/somar_b200.h"
#include "level.hpp"

#include <mutex>

namespace somar {

namespace {

constexpr int NOALLOC_TPB = 256;
constexpr int NOALLOC_CELLS_PER_BLOCK = 4096;
constexpr int NOALLOC_MAX_BLOCKS = 16384;
constexpr int NOALLOC_TICKET_WORDS = 32;   // one 256-byte line after the partials
static_assert((long long)NOALLOC_MAX_BLOCKS * NOALLOC_CELLS_PER_BLOCK == SOMAR_NOALLOC_MAX_N, "header and kernel disagree");

// partials [0, NOALLOC_MAX_BLOCKS), then the ticket (the first 4 bytes of the last line)
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_MAX_BLOCKS + NOALLOC_TICKET_WORDS];

__device__ __forceinline__ double na_warp_red(double v, bool isMax) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double w = __shfl_down_sync(0xffffffffu, v, o);
        v = isMax ? fmax(v, w) : v + w;
    }
    return v;
}
__device__ __forceinline__ double na_block_red(double v, bool isMax) {
    __shared__ double sh[32];
    __syncthreads();
    v = na_warp_red(v, isMax);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        double w = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0.0;
        v = na_warp_red(w, isMax);
    }
    return v;   // valid in thread 0
}

// kind: 0 max |x|, 1 sum |x|, 2 sum x^2 (the square root is taken by the folding block), as k_reduce's kinds 0..2.
// Block b reduces x[b * 4096, (b + 1) * 4096); the last block to finish folds the partials in index order, so the
// result is deterministic for a given n.
__global__ void __launch_bounds__(NOALLOC_TPB)
k_norm_noalloc(const double* __restrict__ x, long long n, int kind, double* scratch, double* __restrict__ out) {
    const bool isMax = kind == 0;
    unsigned int* const ticket = reinterpret_cast<unsigned int*>(scratch + NOALLOC_MAX_BLOCKS);
    const long long b0 = (long long)blockIdx.x * NOALLOC_CELLS_PER_BLOCK;
    double acc = 0.0;
    double v[NOALLOC_CELLS_PER_BLOCK / NOALLOC_TPB / 4][4];
#pragma unroll
    for (int r = 0; r < NOALLOC_CELLS_PER_BLOCK / NOALLOC_TPB / 4; ++r) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {   // four independent loads in flight per thread
            const long long i = b0 + (long long)(r * 4 + u) * NOALLOC_TPB + threadIdx.x;
            v[r][u] = (i < n) ? x[i] : 0.0;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const double w = v[r][u];
            if (kind == 0) acc = fmax(acc, fabs(w));
            else if (kind == 1) acc += fabs(w);
            else acc += w * w;
        }
    }
    acc = na_block_red(acc, isMax);
    __shared__ bool last;
    if (threadIdx.x == 0) {
        scratch[blockIdx.x] = acc;
        __threadfence();
        last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    double s = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
        const double w = ((volatile double*)scratch)[b];
        s = isMax ? fmax(s, w) : s + w;
    }
    s = na_block_red(s, isMax);
    if (threadIdx.x == 0) {
        out[0] = (kind == 2) ? sqrt(s) : s;
        *ticket = 0u;   // leaves the array ready even if the next caller's memset were skipped
    }
}

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;
double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};   // device address of g_noalloc_scratch, per device ordinal

// the device address of the scratch array on the CURRENT device (taken once per device)
int noalloc_scratch(double** out) {
    int dev = -1;
    SOMAR_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) { set_error("somar_norm_noalloc: device ordinal out of range"); return SOMAR_ERR_ARG; }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[dev]) {
        double* p = nullptr;
        SOMAR_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));   // NOT &g_noalloc_scratch: that is a host stand-in
        if (!p) { set_error("somar_norm_noalloc: cudaGetSymbolAddress returned a null pointer"); return SOMAR_ERR_CUDA; }
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return SOMAR_OK;
}

}  // namespace
}  // namespace somar

using namespace somar;

extern "C" int somar_norm_noalloc(const double* d_x, long long n, int ord, double* d_out, void* cuda_stream) {
    if (!d_x || !d_out) { set_error("somar_norm_noalloc: null pointer"); return SOMAR_ERR_ARG; }
    if (ord < 0 || ord > 2) { set_error("somar_norm_noalloc: ord must be 0 (max), 1 or 2"); return SOMAR_ERR_ARG; }
    if (n < 1) { set_error("somar_norm_noalloc: n must be positive"); return SOMAR_ERR_ARG; }
    if (n > (long long)SOMAR_NOALLOC_MAX_N) {
        set_error("somar_norm_noalloc: n = " + std::to_string(n) + " exceeds SOMAR_NOALLOC_MAX_N = " +
                  std::to_string((long long)SOMAR_NOALLOC_MAX_N) +
                  " (the static scratch array holds one partial per 4096 values; this entry point never allocates)");
        return SOMAR_ERR_ARG;
    }
    double* scratch = nullptr;
    int rc = noalloc_scratch(&scratch);
    if (rc) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    // the ticket must start at zero (the array keeps whatever the last call -- possibly an aborted one -- left)
    SOMAR_CUDA(cudaMemsetAsync(scratch + NOALLOC_MAX_BLOCKS, 0, sizeof(double) * NOALLOC_TICKET_WORDS, s));
    const int blocks = (int)((n + NOALLOC_CELLS_PER_BLOCK - 1) / NOALLOC_CELLS_PER_BLOCK);
    k_norm_noalloc<<<blocks, NOALLOC_TPB, 0, s>>>(d_x, n, ord, scratch, d_out);
    SOMAR_CUDA(cudaGetLastError());
    return SOMAR_OK;
}

// cells.cuh -- shared device helpers of the streaming kernels: the valid-cell iteration macros, the launch
// shape that goes with them, and the deterministic two-stage reduction (warp shuffle -> block -> last block).
#pragma once
#include "kernels.hpp"

namespace somar {

#define TPB 256
#define MAX_RED_BLOCKS 1184   // 148 SMs x 8

static inline int row_blocks(const Geom& g, int sms) {
    const long long rows = (long long)g.n[1] * g.n[2];
    const long long perBlock = TPB / 32;                   // one warp per row slice
    long long nb = (rows + perBlock - 1) / perBlock;
    const long long cap = (long long)sms * 8;
    return (int)(nb < 1 ? 1 : (nb > cap ? cap : nb));
}

// Iterate valid cells: each warp takes rows (j,k) round-robin, lanes stride x.
#define FOR_VALID_CELLS(g, c, i, j, k, BODY)                                                              \
    {                                                                                                     \
        const int lane_ = threadIdx.x & 31;                                                               \
        const long long warp_ = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);           \
        const long long nwarps_ = (long long)gridDim.x * (blockDim.x >> 5);                               \
        const long long rows_ = (long long)(g).n[1] * (g).n[2];                                           \
        for (long long r_ = warp_; r_ < rows_; r_ += nwarps_) {                                           \
            const int k = (int)(r_ / (g).n[1]);                                                           \
            const int j = (int)(r_ - (long long)k * (g).n[1]);                                            \
            const long long row0_ = (g).idx(0, j, k);                                                     \
            for (int i = lane_; i < (g).n[0]; i += 32) {                                                  \
                const long long c = row0_ + i;                                                            \
                BODY                                                                                      \
            }                                                                                             \
        }                                                                                                 \
    }

// The same iteration with U cells of the row per lane and step (i = lane, lane+32, ...): LOAD is expanded for all
// U before USE, so a thread keeps U independent loads per array in flight -- a streaming kernel with one 8-byte
// load per thread in flight reaches less than half of the HBM bandwidth, and a latency-bound pass over an
// L2-resident level is U times shorter.  The order in which a lane visits its cells is the same as in
// FOR_VALID_CELLS for every U (reductions stay bit-identical).  i, j, k, c are visible in LOAD and USE.
#define FOR_VALID_CELLSN(g, c, u, U, LOAD, USE)                                                           \
    {                                                                                                     \
        const int lane_ = threadIdx.x & 31;                                                               \
        const long long warp_ = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);           \
        const long long nwarps_ = (long long)gridDim.x * (blockDim.x >> 5);                               \
        const long long rows_ = (long long)(g).n[1] * (g).n[2];                                           \
        for (long long r_ = warp_; r_ < rows_; r_ += nwarps_) {                                           \
            const int k = (int)(r_ / (g).n[1]);                                                           \
            const int j = (int)(r_ - (long long)k * (g).n[1]);                                            \
            const long long row0_ = (g).idx(0, j, k);                                                     \
            (void)k; (void)j;                                                                             \
            for (int i0_ = lane_; i0_ < (g).n[0]; i0_ += 32 * (U)) {                                      \
                _Pragma("unroll") for (int u = 0; u < (U); ++u) {                                         \
                    const int i = i0_ + 32 * u;                                                           \
                    const long long c = row0_ + i;                                                        \
                    (void)c;                                                                              \
                    if (i < (g).n[0]) { LOAD }                                                            \
                }                                                                                         \
                _Pragma("unroll") for (int u = 0; u < (U); ++u) {                                         \
                    const int i = i0_ + 32 * u;                                                           \
                    const long long c = row0_ + i;                                                        \
                    (void)c;                                                                              \
                    if (i < (g).n[0]) { USE }                                                             \
                }                                                                                         \
            }                                                                                             \
        }                                                                                                 \
    }
#define FOR_VALID_CELLS4(g, c, u, LOAD, USE) FOR_VALID_CELLSN(g, c, u, 4, LOAD, USE)

// ---- reductions -------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_red(double v, bool isMax) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double w = __shfl_down_sync(0xffffffffu, v, o);
        v = isMax ? fmax(v, w) : v + w;
    }
    return v;
}
__device__ __forceinline__ double block_red(double v, bool isMax, double ident) {
    __shared__ double sh[32];
    __syncthreads();
    v = warp_red(v, isMax);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        double w = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : ident;
        w = warp_red(w, isMax);
        v = w;
    }
    return v;   // valid in thread 0
}

// Per-block partial results, then the LAST block to finish (atomic ticket) folds them in a fixed order:
// deterministic for a given grid, one launch.
static __device__ void finish_reduce(double part, bool isMax, double ident, double* __restrict__ scratch,
                              double* __restrict__ out, unsigned int* __restrict__ g_ticket, int slot,
                              bool accumulate) {
    __shared__ bool last;
    if (threadIdx.x == 0) {
        scratch[blockIdx.x] = part;
        __threadfence();
        const unsigned t = atomicAdd(&g_ticket[slot], 1u);
        last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (last) {
        __threadfence();
        double v = ident;
        for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
            const double w = ((volatile double*)scratch)[b];
            v = isMax ? fmax(v, w) : v + w;
        }
        v = block_red(v, isMax, ident);
        if (threadIdx.x == 0) {
            out[0] = accumulate ? out[0] + v : v;
            g_ticket[slot] = 0;
        }
    }
}

}  // namespace somar

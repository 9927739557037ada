// gsrb_fused.cu -- one full red-black Gauss-Seidel ITERATION (red pass, then black pass) of the diagonal-metric
// operator in ONE sweep over HBM, bit-identical to two k_gsrb_ortho colour passes.
//
// Why: a colour pass updates every other cell, but DRAM moves 32-byte sectors that hold two cells of each colour,
// so two passes move every array twice (~128 B/cell/iteration against 96 B of algorithmic traffic).
// Here each CTA marches a tile of columns through z with a 5-plane ring of phi in shared memory:
//   step k:  red update of plane k on the tile + 1-cell ring   (needs OLD black of planes k-1, k, k+1)
//            black update of plane k-1 on the tile             (needs NEW red of planes k-2, k-1, k)
//            plane k-1 is final -> written to HBM (both colours, full sectors); plane k+2 is prefetched.
// The ring cells' red values are recomputed by the neighbouring CTA (a 1-cell halo of redundant work), phi is
// read once (+ halo, mostly L2 hits) and written once, every coefficient array is read once: ~64 B/cell/iteration.
// Thread (t, r) owns the x-adjacent cell PAIR (2t, 2t+1) of row r of the extended tile: exactly one of the two is red
// in any plane, and the black cell it updates in plane k-1 sits in the same column as its red cell of plane k,
// so every thread does one red and one black update per step.
//
// Out of place (phiIn -> phiOut): neighbouring CTAs read each other's halo cells of the PRE-iteration state while
// they run concurrently, so the new values go to a second brick (the level's ping-pong partner); the traffic is
// the same as in place (read 8 B, write 8 B per cell).
// Used when no halo exchange is needed between the two colour passes (single rank, no periodic direction);
// otherwise the two-pass kernel with an exchange before each pass is used (LevelGSRB::relax, GSRB.cpp:58-98).
#include "kernels.hpp"
#include "stencil_ortho.cuh"

namespace somar {

#define FPX 32                 // pairs per row of the extended tile  -> 64 cells
#define FEY 16                 // rows of the extended tile
#define FEX (2 * FPX)
#define FCX (FEX - 2)          // core tile: 62 x 14
#define FCY (FEY - 2)
#define FSX (FEX + 2)          // shared-memory plane: extended tile + 1 ring = 66 x 18
#define FSY (FEY + 2)
#define FSLOTS 5
#define FTHREADS (FPX * FEY)
#define FPLANE (FSX * FSY)
#define FLOADS ((FPLANE + FTHREADS - 1) / FTHREADS)

__device__ __forceinline__ int fslot(int plane) { return ((plane % FSLOTS) + FSLOTS) % FSLOTS; }

__global__ void __launch_bounds__(FTHREADS, 2) k_gsrb_fused(const Geom g, const Metric m,
                                                          double* __restrict__ phiOut,
                                                          const double* __restrict__ phi,
                                                          const double* __restrict__ rhs, int kc) {
    __shared__ double sm[FSLOTS][FSY][FSX];
    const int tid = threadIdx.x;
    const int t = tid % FPX, r = tid / FPX;
    const int x0 = blockIdx.x * FCX - 1;      // local x of extended column 0
    const int y0 = blockIdx.y * FCY - 1;
    const int k0 = blockIdx.z * kc;
    const int k1 = min(k0 + kc, g.n[2]);

    // which plane elements this thread stages (shared-memory plane index e -> (sy, sx))
    auto load_plane = [&](int plane, double (&v)[FLOADS]) {
        const bool zin = (plane >= -g.gh[2]) && (plane < g.n[2] + g.gh[2]);
#pragma unroll
        for (int q = 0; q < FLOADS; ++q) {
            const int e = tid + q * FTHREADS;
            v[q] = 0.0;
            if (e < FPLANE && zin) {
                const int sy = e / FSX, sx = e - sy * FSX;
                const int x = x0 - 1 + sx, y = y0 - 1 + sy;
                if (x >= -2 && x <= g.n[0] + 1 && y >= -2 && y <= g.n[1] + 1) v[q] = phi[g.idx(x, y, plane)];
            }
        }
    };
    auto store_plane = [&](int plane, const double (&v)[FLOADS]) {
        const int s = fslot(plane);
#pragma unroll
        for (int q = 0; q < FLOADS; ++q) {
            const int e = tid + q * FTHREADS;
            if (e < FPLANE) (&sm[s][0][0])[e] = v[q];
        }
    };

    double pre[FLOADS];
    for (int p = k0 - 2; p <= k0; ++p) {
        load_plane(p, pre);
        store_plane(p, pre);
    }
    __syncthreads();

    const int y = y0 + r;                      // local row of this thread
    const bool yin = (y >= 0) && (y < g.n[1]);
    const bool ycore = (r >= 1) && (r <= FCY);
    for (int k = k0 - 1; k <= k1; ++k) {
        load_plane(k + 2, pre);                // prefetch: consumed at step k+1
        // ---- red update of plane k (extended tile) ----
        const int par = (g.g0[0] + x0 + g.g0[1] + y + g.g0[2] + k) & 1;   // parity of extended column 0 in this row/plane
        const int ex = 2 * t + par;            // this thread's red cell in plane k == its black cell in plane k-1
        const int x = x0 + ex;
        const bool xin = (x >= 0) && (x < g.n[0]);
        if (k >= 0 && k < g.n[2] && xin && yin) {
            const int s = fslot(k), sl = fslot(k - 1), sh = fslot(k + 1);
            double nb[3][2];
            nb[0][0] = sm[s][r + 1][ex];
            nb[0][1] = sm[s][r + 1][ex + 2];
            nb[1][0] = sm[s][r][ex + 1];
            nb[1][1] = sm[s][r + 2][ex + 1];
            nb[2][0] = sm[sl][r + 1][ex + 1];
            nb[2][1] = sm[sh][r + 1][ex + 1];
            const long long c = g.idx(x, y, k);
            sm[s][r + 1][ex + 1] = gsrb_update_ortho<3>(g, m, rhs, x, y, k, c, sm[s][r + 1][ex + 1], nb);
        }
        __syncthreads();
        // ---- black update of plane k-1 (core tile), then write the finished plane ----
        const int kb = k - 1;
        if (kb >= k0 && kb < k1 && yin && ycore) {
            const int s = fslot(kb), sl = fslot(kb - 1), sh = fslot(kb + 1);
            const bool xcore = (ex >= 1) && (ex <= FCX);
            if (xin && xcore) {
                double nb[3][2];
                nb[0][0] = sm[s][r + 1][ex];
                nb[0][1] = sm[s][r + 1][ex + 2];
                nb[1][0] = sm[s][r][ex + 1];
                nb[1][1] = sm[s][r + 2][ex + 1];
                nb[2][0] = sm[sl][r + 1][ex + 1];
                nb[2][1] = sm[sh][r + 1][ex + 1];
                const long long c = g.idx(x, y, kb);
                phiOut[c] = gsrb_update_ortho<3>(g, m, rhs, x, y, kb, c, sm[s][r + 1][ex + 1], nb);
            }
            // the other cell of the pair is red in plane k-1: its new value sits in shared memory
            const int eo = 2 * t + (1 - par);
            const int xo = x0 + eo;
            if (xo >= 0 && xo < g.n[0] && eo >= 1 && eo <= FCX) phiOut[g.idx(xo, y, kb)] = sm[s][r + 1][eo + 1];
        }
        store_plane(k + 2, pre);               // slot of plane k-3: dead since the previous step
        __syncthreads();
    }
}

bool gsrb_fused_usable(const Geom& g, const Metric& m) {
    if (!m.diagonal || g.nd != 3) return false;
    if (g.periodic[0] || g.periodic[1] || g.periodic[2]) return false;
    if (g.nbr[0] >= 0 || g.nbr[1] >= 0) return false;
    return true;
}

void launch_gsrb_fused(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn,
                       const double* rhs) {
    const int kc = (g.n[2] > 64) ? 64 : g.n[2];
    dim3 grid((g.n[0] + FCX - 1) / FCX, (g.n[1] + FCY - 1) / FCY, (g.n[2] + kc - 1) / kc);
    k_gsrb_fused<<<grid, FTHREADS, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, kc);
    cx.launches++;
}

}  // namespace somar

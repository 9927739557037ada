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
// read once (+ halo, mostly L2 hits) and written once, every coefficient array is read once: measured
// 70 B/cell/iteration of DRAM traffic (profiles/).
// Thread (t, r) owns the x-adjacent cell PAIR (2t, 2t+1) of row r of the extended tile: exactly one of the two is red
// in any plane, and the black cell it updates in plane k-1 sits in the same column as its red cell of plane k,
// so every thread does one red and one black update per step.
// Latency and issue slots: the coefficients of step k+1 (9 values per update) and the phi plane k+2 are loaded
// into registers at the top of step k and consumed one step later (software pipeline, two steps unrolled so the
// register sets swap roles without copies); all addresses are per-thread bases advanced by one plane stride per
// step; tiles that do not touch the domain rim skip the boundary-stencil logic (block-uniform branch).
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
#ifndef FEY
#define FEY 16                 // rows of the extended tile
#endif
#define FEX (2 * FPX)
#define FCX (FEX - 2)          // core tile: 62 x 14
#define FCY (FEY - 2)
#define FSX (FEX + 2)          // shared-memory plane: extended tile + 1 ring = 66 x 18
#define FSY (FEY + 2)
#define FSLOTS 5
#define FTHREADS (FPX * FEY)
#define FPLANE (FSX * FSY)
#define FLOADS ((FPLANE + FTHREADS - 1) / FTHREADS)

__device__ __forceinline__ int slot_dec(int s) { return s == 0 ? FSLOTS - 1 : s - 1; }
__device__ __forceinline__ int slot_inc(int s) { return s == FSLOTS - 1 ? 0 : s + 1; }

// global index (at plane 0) of shared-plane element e of the tile whose extended origin is (x0, y0), or -1
__device__ __forceinline__ int stage_off(const Geom& g, int e, int x0, int y0) {
    if (e >= FPLANE) return -1;
    const int sy_ = e / FSX, sx_ = e - sy_ * FSX;
    const int xx = x0 - 1 + sx_, yy = y0 - 1 + sy_;
    if (xx < -2 || xx > g.n[0] + 1 || yy < -2 || yy > g.n[1] + 1) return -1;
    return (int)g.idx(xx, yy, 0);
}

// One step of the sweep, written as a macro so that all state stays in named registers (no closures, no local
// arrays).  (CR, CB) = coefficients of this step's red / black update; (NR, NB) receive those of the next step.
// Offsets are 32-bit element indices into the brick (the launcher checks that the brick has < 2^31 elements).
#define FUSED_STEP(K, CR, NR)                                                                                     \
    {                                                                                                             \
        const int k_ = (K);                                                                                       \
        const int par = (parBase + k_) & 1, parn = par ^ 1;                                                       \
        const int sl = slot_dec(s), sh = slot_inc(s);                                                             \
        /* software pipeline: everything step k+1 needs from HBM is requested now */                              \
        {                                                                                                         \
            const int p2 = k_ + 2;                                                                                \
            const bool zin = (p2 >= -g.gh[2]) && (p2 < g.n[2] + g.gh[2]);                                         \
            const int po = kOff + 2 * isz;                                                                        \
            pre0 = (zin && goff0 >= 0) ? phi[goff0 + po] : 0.0;                                                   \
            pre1 = (zin && goff1 >= 0) ? phi[goff1 + po] : 0.0;                                                   \
            pre2 = (zin && goff2 >= 0) ? phi[goff2 + po] : 0.0;                                                   \
        }                                                                                                         \
        if (k_ + 1 <= k1 && k_ + 1 >= 0 && k_ + 1 < g.n[2] && ((colMask >> parn) & 1))                            \
            NR = gsrb_load_coef<3>(g, m, rhs, cbase + parn + kOff + isz);                                         \
        /* this step's black update (plane k-1, same column as the red one): its sectors were fetched one step   \
           ago for the red cells next to it, so this is an L1/L2 hit that completes during the red phase */       \
        GsrbCoef CB = GsrbCoef();                                                                                 \
        if (k_ - 1 >= k0 && k_ - 1 < k1 && ((coreMask >> par) & 1))                                               \
            CB = gsrb_load_coef<3>(g, m, rhs, cbase + par + kOff - isz);                                          \
        const int ex = 2 * t + par;                                                                               \
        const int x = x0 + ex;                                                                                    \
        /* red update of plane k (extended tile) */                                                               \
        if (k_ >= 0 && k_ < g.n[2] && ((colMask >> par) & 1)) {                                                   \
            const int gk = g.g0[2] + k_;                                                                          \
            const bool fast = !tileRim && gk != 0 && gk != g.gn[2] - 1;                                           \
            double nb[3][2];                                                                                      \
            nb[0][0] = sm[s][r + 1][ex];                                                                          \
            nb[0][1] = sm[s][r + 1][ex + 2];                                                                      \
            nb[1][0] = sm[s][r][ex + 1];                                                                          \
            nb[1][1] = sm[s][r + 2][ex + 1];                                                                      \
            nb[2][0] = sm[sl][r + 1][ex + 1];                                                                     \
            nb[2][1] = sm[sh][r + 1][ex + 1];                                                                     \
            sm[s][r + 1][ex + 1] = gsrb_update_vals<3>(g, CR, x, y, k_, sm[s][r + 1][ex + 1], nb, fast);          \
        }                                                                                                         \
        __syncthreads();                                                                                          \
        /* black update of plane k-1 (core tile), then write the finished plane */                                \
        if (k_ - 1 >= k0 && k_ - 1 < k1) {                                                                        \
            const int kb = k_ - 1;                                                                                \
            const int sll = slot_dec(sl);                                                                         \
            const int cb0 = cbase + kOff - isz;                                                                   \
            if ((coreMask >> par) & 1) {                                                                          \
                const int gk = g.g0[2] + kb;                                                                      \
                const bool fast = !tileRim && gk != 0 && gk != g.gn[2] - 1;                                       \
                double nb[3][2];                                                                                  \
                nb[0][0] = sm[sl][r + 1][ex];                                                                     \
                nb[0][1] = sm[sl][r + 1][ex + 2];                                                                 \
                nb[1][0] = sm[sl][r][ex + 1];                                                                     \
                nb[1][1] = sm[sl][r + 2][ex + 1];                                                                 \
                nb[2][0] = sm[sll][r + 1][ex + 1];                                                                \
                nb[2][1] = sm[s][r + 1][ex + 1];                                                                  \
                phiOut[cb0 + par] = gsrb_update_vals<3>(g, CB, x, y, kb, sm[sl][r + 1][ex + 1], nb, fast);        \
            }                                                                                                     \
            /* the other cell of the pair is red in plane k-1: its new value sits in shared memory */             \
            if ((coreMask >> parn) & 1) phiOut[cb0 + parn] = sm[sl][r + 1][2 * t + parn + 1];                     \
        }                                                                                                         \
        /* plane k+2 -> the slot of plane k-3 (dead since the previous step) */                                   \
        {                                                                                                         \
            double* dst = smflat + slot_inc(sh) * FPLANE + tid;                                                   \
            dst[0] = pre0;                                                                                        \
            dst[FTHREADS] = pre1;                                                                                 \
            if (tid + 2 * FTHREADS < FPLANE) dst[2 * FTHREADS] = pre2;                                            \
        }                                                                                                         \
        s = sh;                                                                                                   \
        kOff += isz;                                                                                              \
        __syncthreads();                                                                                          \
    }

static_assert(FLOADS == 3, "staging code assumes three elements per thread");

__global__ void __launch_bounds__(FTHREADS, (FTHREADS <= 256 ? 2 : 1)) k_gsrb_fused(const Geom g, const Metric m,
                                                             double* __restrict__ phiOut,
                                                             const double* __restrict__ phi,
                                                             const double* __restrict__ rhs, int kc) {
    __shared__ double sm[FSLOTS][FSY][FSX];
    double* const smflat = &sm[0][0][0];
    const int tid = threadIdx.x;
    const int t = tid % FPX, r = tid / FPX;
    const int x0 = blockIdx.x * FCX - 1;      // local x of extended column 0
    const int y0 = blockIdx.y * FCY - 1;
    const int k0 = blockIdx.z * kc;
    const int k1 = min(k0 + kc, g.n[2]);
    const int isz = (int)g.sz;

    // ---- per-thread constants -------------------------------------------------------------------------------
    // phi staging: global index (at plane 0) of the three plane elements this thread copies, or -1
    const int goff0 = stage_off(g, tid, x0, y0);
    const int goff1 = stage_off(g, tid + FTHREADS, x0, y0);
    const int goff2 = stage_off(g, tid + 2 * FTHREADS, x0, y0);
    const int y = y0 + r;                      // local row of this thread
    const bool yin = (y >= 0) && (y < g.n[1]);
    const bool ycore = (r >= 1) && (r <= FCY);
    const int parBase = g.g0[0] + x0 + g.g0[1] + y + g.g0[2];
    const int cbase = (int)g.idx(x0 + 2 * t, y, 0);
    int colMask = 0, coreMask = 0;
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int ex = 2 * t + p, x = x0 + ex;
        const bool col = yin && x >= 0 && x < g.n[0];
        if (col) colMask |= 1 << p;
        if (col && ycore && ex >= 1 && ex <= FCX) coreMask |= 1 << p;
    }
    // block-uniform: does the extended tile touch the domain rim in x or y?
    const bool tileRim = (g.g0[0] + x0 <= 0) || (g.g0[0] + x0 + FEX - 1 >= g.gn[0] - 1) || (g.g0[1] + y0 <= 0) ||
                         (g.g0[1] + y0 + FEY - 1 >= g.gn[1] - 1);

    // ---- prologue: planes k0-2 .. k0 into the ring, coefficients of the first step ---------------------------
    double pre0, pre1, pre2;
    int s = ((k0 - 2) % FSLOTS + FSLOTS) % FSLOTS;
    for (int p = k0 - 2; p <= k0; ++p) {
        const bool zin = (p >= -g.gh[2]) && (p < g.n[2] + g.gh[2]);
        const int po = p * isz;
        pre0 = (zin && goff0 >= 0) ? phi[goff0 + po] : 0.0;
        pre1 = (zin && goff1 >= 0) ? phi[goff1 + po] : 0.0;
        pre2 = (zin && goff2 >= 0) ? phi[goff2 + po] : 0.0;
        double* dst = smflat + s * FPLANE + tid;
        dst[0] = pre0;
        dst[FTHREADS] = pre1;
        if (tid + 2 * FTHREADS < FPLANE) dst[2 * FTHREADS] = pre2;
        s = slot_inc(s);
    }
    s = slot_dec(slot_dec(s));                 // slot of plane k0-1 = first red plane
    int kOff = (k0 - 1) * isz;                 // plane offset of the current step
    GsrbCoef A_r = GsrbCoef(), B_r = GsrbCoef();
    {
        const int par = (parBase + k0 - 1) & 1;
        if (k0 - 1 >= 0 && ((colMask >> par) & 1)) A_r = gsrb_load_coef<3>(g, m, rhs, cbase + par + kOff);
    }
    __syncthreads();
    for (int k = k0 - 1; k <= k1; k += 2) {
        FUSED_STEP(k, A_r, B_r)
        if (k + 1 <= k1) FUSED_STEP(k + 1, B_r, A_r)
    }
}

bool gsrb_fused_usable(const Geom& g, const Metric& m) {
    if (!m.diagonal || g.nd != 3) return false;
    if (g.periodic[0] || g.periodic[1] || g.periodic[2]) return false;
    if (g.nbr[0] >= 0 || g.nbr[1] >= 0) return false;
    if (g.total >= (1LL << 31)) return false;  // 32-bit brick indices
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

// bicg_resident.cu -- the bottom visit of the V-cycle with the whole bottom problem RESIDENT ON CHIP.
//
// MappedMultiGrid<T>::cycle at the bottom depth (src/calculus/AMRElliptic/MappedMultiGrid.H:566-573): num_smooth_bottom
// relaxations, then Chombo's BiCGStabSolver (provenance in bicgstab.cu) preconditioned by diagonal scaling +
// num_smooth_precond red-black relaxations (MappedAMRPoissonOp::preCond, MappedAMRPoissonOp.cpp:684-734).
// With the bottom solver given enough iterations to actually solve the 256 x 256 x 4 bottom problem of the headline
// (bottom.imax 1000: ~110 iterations per visit) the visit is 16 % of a V-cycle on one GPU and, being replicated, the
// largest single item on eight.  k_bicg_persistent (bicg_persistent.cuh) spends ~125 us per iteration: every pass
// streams its operands from L2 (nine coefficient loads per relaxed cell) and the iteration has nine grid barriers.
// Here one CTA per SM owns a 32 x 16 x nz (nz <= 4) tile for the whole visit:
//   * the tile's coefficients (J g^xx, J g^yy, J g^zz faces and 1/J, core + 5-cell halo) are loaded into shared
//     memory ONCE per visit (149 KB);
//   * a thread owns one z-column of the core: r, r~, e, p, v and phi of its (up to) four cells live in REGISTERS for
//     the whole visit -- the BLAS-1 updates, dot products and norms of BiCGStab touch no memory at all;
//   * preCond + operator ("relax n times, then apply L") is one pass over a shared-memory tile with a halo of
//     2n+1 cells: the CTA loads the vector it is applied to (the only array that crosses CTAs, staged in global
//     memory = L2), relaxes it in shared memory -- every colour pass valid on a region one cell smaller, halo cells
//     relaxed redundantly with the statements and operands of the owning CTA -- and applies the operator on the core;
//   * per iteration: 2 staged vectors, 5 grid barriers (was 9), no coefficient traffic.
// The arithmetic per cell is gsrb_update_vals / apply_vals_ortho3 (the statements of GSRBITER3DORTHO and
// MAPPEDGETFLUXORTHO + MAPPEDFLUXDIVERGENCE3D), so preCond and L are bit-identical to the other bottom modes; the
// SUMS (dot products, norms) are folded per tile and then over tiles, a different -- still deterministic, and the same
// on every rank of a replicated bottom -- order than k_reduce's, so this mode agrees with modes 0-2 to rounding, not bit
// for bit (tests/test_gpu_solve.py::test_resident_bottom_matches_the_other_modes).
// Usable for: 3-D, nz <= 4, diagonal metric, no periodic direction, point GSRB preconditioner with <= 2 relaxations,
// <= 2 bottom relaxations, one tile per SM at most.  Anything else takes k_bicg_persistent (mode 2) as before.
#include "bicg_persistent.cuh"
#include "level.hpp"

#include <cstdio>
#include <cstdlib>

namespace somar {

namespace {

constexpr int RB_CX = 32, RB_CY = 16, RB_NZ = 4, RB_H = 5;     // core tile, max z cells, max halo
constexpr int RB_W = RB_CX + 2 * RB_H, RB_HH = RB_CY + 2 * RB_H; // 42 x 26 region
constexpr int RB_PL = RB_W * RB_HH;                              // cells of a region plane (1092)
constexpr int RB_CELLS = RB_PL * RB_NZ;                          // 4368
constexpr int RB_TPB = RB_CX * RB_CY;                            // 512: one thread per core column
// shared memory (doubles): JX, JY, JI: RB_CELLS each; JZ: one more plane; G (working tile), R (right-hand side)
constexpr int RB_OFF_JX = 0, RB_OFF_JY = RB_OFF_JX + RB_CELLS, RB_OFF_JI = RB_OFF_JY + RB_CELLS,
              RB_OFF_JZ = RB_OFF_JI + RB_CELLS, RB_OFF_G = RB_OFF_JZ + RB_CELLS + RB_PL, RB_OFF_R = RB_OFF_G + RB_CELLS,
              RB_SMEM_DOUBLES = RB_OFF_R + RB_CELLS;
constexpr int RB_SMEM_BYTES = RB_SMEM_DOUBLES * 8;               // 218 400 B
static_assert(RB_SMEM_BYTES <= 227 * 1024, "shared memory budget");

// development probe (SOMAR_B200_BOTTOM_PROF=1): nanoseconds block 0 spends in [0] tile-pass load, [1] colour passes,
// [2] apply, [3] grid barriers incl. waiting, [4] everything else; [5] iterations
__device__ unsigned long long g_rb_prof[8];
__device__ __forceinline__ unsigned long long rb_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define RB_PROF(slot, t0) do { if (prof && blockIdx.x == 0 && threadIdx.x == 0) { const unsigned long long t1_ = rb_now(); g_rb_prof[slot] += t1_ - (t0); (t0) = t1_; } } while (0)

struct RbTile {
    int cx0, cy0, ox, oy;   // core origin, region origin (local cell indices; the region may stick out of the domain)
    int x, y;               // my core column
    bool own;               // my column is inside the domain
    int lcol;               // region index of my column in plane 0
};

// "relax nrelax times, then apply L" on my tile.  G <- src (scaled by the diagonal when `scaled`: the first statement
// of preCond, DIAGPRECOND), R <- rhsSrc (scaled: the same array), both over core + h clipped to the domain;
// then 2*nrelax colour passes in shared memory and the operator on the core.
//   out: pc[k] = relaxed value, lp[k] = L(relaxed) of my column's cells (garbage where !own or k >= nz)
__device__ __forceinline__ void rb_tile_pass(const Geom& g, const Metric& m, double* __restrict__ sm, const RbTile& T,
                                             const double* src, const double* rhsSrc, bool scaled, int nrelax,
                                             double (&pc)[RB_NZ], double (&lp)[RB_NZ], bool prof, unsigned long long& tp) {
    const int nx = g.n[0], ny = g.n[1], nz = g.n[2];
    const int h = 2 * nrelax + 1;
    const int tid = threadIdx.x;
    double* const G = sm + RB_OFF_G;
    double* const R = sm + RB_OFF_R;
    const double* const JX = sm + RB_OFF_JX;
    const double* const JY = sm + RB_OFF_JY;
    const double* const JZ = sm + RB_OFF_JZ;
    const double* const JI = sm + RB_OFF_JI;
    __syncthreads();   // the previous pass's readers are done with G and R
    RB_PROF(4, tp);
    {
        const int xa = max(0, T.cx0 - h), xb = min(nx, T.cx0 + RB_CX + h), ya = max(0, T.cy0 - h), yb = min(ny, T.cy0 + RB_CY + h);
        const int wx = xb - xa, wy = yb - ya;
        for (int q = tid; q < wx * wy * nz; q += RB_TPB) {
            const int k = q / (wx * wy), r = q - k * wx * wy;
            const int jj = r / wx;
            const int j = ya + jj, i = xa + (r - jj * wx);
            const long long c = g.idx(i, j, k);
            const int l = (i - T.ox) + RB_W * (j - T.oy) + RB_PL * k;
            const double v = src[c];
            if (scaled) {
                G[l] = v / (g.alpha + g.beta * m.lapdiag[c]);
                R[l] = v;
            } else {
                G[l] = v;
                if (nrelax > 0) R[l] = rhsSrc[c];
            }
        }
    }
    __syncthreads();
    RB_PROF(0, tp);
    int sdone = 0;
    for (int it = 0; it < nrelax; ++it)
        for (int pass = 0; pass < 2; ++pass) {
            ++sdone;
            const int e = h - sdone;   // this colour pass is valid on core + e
            const int xa = max(0, T.cx0 - e), xb = min(nx, T.cx0 + RB_CX + e), ya = max(0, T.cy0 - e), yb = min(ny, T.cy0 + RB_CY + e);
            // Cell mapping of a colour pass: a warp takes 32 CONSECUTIVE columns of a PAIR of rows, lane by lane
            // alternating between the two rows (the cell of the pass's colour in each column).  Shared-memory words of
            // adjacent lanes then differ by 1 or by 1 + RB_W (= 43: odd), so a half-warp's 16 eight-byte accesses hit 16
            // different banks -- a row-wise mapping strides by 2 cells and is a 2-way conflict on every access (the
            // passes were LDS-bound at 3.4 us each).  The decode uses float reciprocals (exact for these small integers)
            // instead of two runtime integer divisions per cell.
            const int wx = xb - xa, wy = yb - ya;
            const int npair = (wy + 1) >> 1;
            const int perPlane = wx * npair, total = perPlane * nz;
            const float rPlane = 1.0f / (float)perPlane, rWx = 1.0f / (float)wx;
            for (int q = tid; q < total; q += RB_TPB) {
                const int k = (int)(((float)q + 0.5f) * rPlane);
                const int r = q - k * perPlane;
                const int rp = (int)(((float)r + 0.5f) * rWx);
                const int i = xa + (r - rp * wx);
                // the row of the pair whose cell in column i has this pass's colour: (g0 + i + j + k + pass) even
                const int j = ya + 2 * rp + ((g.g0[0] + i + g.g0[1] + ya + g.g0[2] + k + pass) & 1);
                if (j < yb) {
                    const int l = (i - T.ox) + RB_W * (j - T.oy) + RB_PL * k;
                    double nb[3][2];
                    nb[0][0] = G[l - 1];
                    nb[0][1] = G[l + 1];
                    nb[1][0] = G[l - RB_W];
                    nb[1][1] = G[l + RB_W];
                    nb[2][0] = (k > 0) ? G[l - RB_PL] : 0.0;
                    nb[2][1] = (k < nz - 1) ? G[l + RB_PL] : 0.0;
                    GsrbCoef qc;
                    qc.glo[0] = JX[l]; qc.ghi[0] = JX[l + 1];
                    qc.glo[1] = JY[l]; qc.ghi[1] = JY[l + RB_W];
                    qc.glo[2] = JZ[l]; qc.ghi[2] = JZ[l + RB_PL];
                    qc.ji = JI[l];
                    qc.lap = 0.0;   // formed from the faces (LAPFLY): the bits of the stored lapDiag
                    qc.rhs = R[l];
                    G[l] = gsrb_update_vals<3, true>(g, qc, i, j, k, G[l], nb);
                }
            }
            __syncthreads();
        }
    RB_PROF(1, tp);
    if (T.own) {
#pragma unroll
        for (int k = 0; k < RB_NZ; ++k) {
            if (k < nz) {
                const int l = T.lcol + RB_PL * k;
                const double c0 = G[l];
                double nb[3][2];
                nb[0][0] = G[l - 1];
                nb[0][1] = G[l + 1];
                nb[1][0] = G[l - RB_W];
                nb[1][1] = G[l + RB_W];
                nb[2][0] = (k > 0) ? G[l - RB_PL] : 0.0;
                nb[2][1] = (k < nz - 1) ? G[l + RB_PL] : 0.0;
                const double glo[3] = {JX[l], JY[l], JZ[l]};
                const double ghi[3] = {JX[l + 1], JY[l + RB_W], JZ[l + RB_PL]};
                lp[k] = apply_vals_ortho3(g, side_mask(g, T.x, T.y, k), true, c0, nb, glo, ghi, JI[l]);
                pc[k] = c0;
            }
        }
    }
    RB_PROF(2, tp);
}

}  // namespace

// One CTA per tile, cooperative launch (grid barriers).  `part`: 4 * MAX_RED_BLOCKS doubles of reduction partials.
__global__ void __launch_bounds__(RB_TPB, 1)
k_bicg_resident(const Geom g, const Metric m, const BicgFields f, const BicgArgs a, double* part, int* itersAccum,
                const int tilesX, unsigned long long* gbar) {
    cg::grid_group grid = cg::this_grid();
    // Grid-wide barrier.  gbar == nullptr: cooperative groups' grid barrier (every CTA's thread 0 adds to ONE word and
    // polls it: ~128 serialised atomics).  gbar != nullptr (the launch has thread-block clusters): the hardware cluster
    // barrier inside each cluster, one atomic per CLUSTER on a counter that only grows (zeroed before the launch), the
    // cluster barrier again.  Visibility: barrier.cluster is release/acquire at cluster scope, the leader's
    // __threadfence() at gpu scope is cumulative over what it observed through the cluster barrier.
    unsigned bepoch = 0;
    const bool prof = a.tcx == 1234;   // (tcx / tcy belong to the persistent kernel's tile pass; 1234 = probe on)
    unsigned long long tp = prof ? rb_now() : 0ull;
    auto gsync = [&]() {
        if (gbar == nullptr) {
            RB_PROF(4, tp);
            grid.sync();
            RB_PROF(3, tp);
            return;
        }
        cg::cluster_group cl = cg::this_cluster();
        cl.sync();
        if (cl.block_rank() == 0 && threadIdx.x == 0) {
            ++bepoch;
            const unsigned long long want = (unsigned long long)bepoch * (gridDim.x / cl.num_blocks());
            __threadfence();
            atomicAdd(gbar, 1ULL);
            const long long t0 = clock64();
            while (*((volatile unsigned long long*)gbar) < want) {
                if (clock64() - t0 > (20ll << 30)) __trap();   // ~10 s: a bug must not hang the GPU
            }
            __threadfence();
        }
        cl.sync();
    };
    extern __shared__ __align__(16) double sm[];
    __shared__ BicgState s;
    __shared__ double rhoNext;
    const int tid = threadIdx.x;
    const int nx = g.n[0], ny = g.n[1], nz = g.n[2];
    const int nb = gridDim.x;
    const int kind = a.normKind;
    const bool nmax = kind == 0;
    int buf = 0;

    RbTile T;
    {
        const int ty = blockIdx.x / tilesX, tx = blockIdx.x - ty * tilesX;
        T.cx0 = tx * RB_CX; T.cy0 = ty * RB_CY;
        T.ox = T.cx0 - RB_H; T.oy = T.cy0 - RB_H;
        T.x = T.cx0 + (tid & (RB_CX - 1)); T.y = T.cy0 + tid / RB_CX;
        T.own = T.x < nx && T.y < ny;
        T.lcol = (T.x - T.ox) + RB_W * (T.y - T.oy);
    }
    // ---- coefficients of core + halo, once per visit.  A face array holds the LOW faces of the cells; the high face
    //      of the last cell of a row / column / plane sits in the first ghost slot of the brick, so x runs to nx
    //      (y to ny, z to nz) inclusive where the region reaches that far.
    {
        const int xa = max(0, T.ox), xb = min(nx + 1, T.ox + RB_W), ya = max(0, T.oy), yb = min(ny + 1, T.oy + RB_HH);
        const int wx = xb - xa, wy = yb - ya;
        for (int q = tid; q < wx * wy * (nz + 1); q += RB_TPB) {
            const int k = q / (wx * wy), r = q - k * wx * wy;
            const int jj = r / wx;
            const int j = ya + jj, i = xa + (r - jj * wx);
            const long long c = g.idx(i, j, k);
            const int l = (i - T.ox) + RB_W * (j - T.oy) + RB_PL * k;
            if (j < ny && i < nx) sm[RB_OFF_JZ + l] = m.jg[2][2][c];
            if (k < nz) {
                if (j < ny) sm[RB_OFF_JX + l] = m.jg[0][0][c];
                if (i < nx) sm[RB_OFF_JY + l] = m.jg[1][1][c];
                if (j < ny && i < nx) sm[RB_OFF_JI + l] = m.jinv[c];
            }
        }
    }
    // (rb_tile_pass starts with a __syncthreads)

    // Per-thread partial results -> s.red[0..n-1] in every block; lane 0 a sum or (max norm) a maximum, lane 1 a sum.
    // Consecutive reductions alternate between two partial buffers: a block can be one barrier ahead, never two.
    auto reduce = [&](double p0, double p1, int n, bool isMax0) {
        p0 = block_red(p0, isMax0, 0.0);
        if (n > 1) p1 = block_red(p1, false, 0.0);
        double* pb = part + (size_t)buf * 2 * MAX_RED_BLOCKS;
        if (tid == 0) {
            pb[blockIdx.x] = p0;
            if (n > 1) pb[MAX_RED_BLOCKS + blockIdx.x] = p1;
        }
        gsync();
        for (int q = 0; q < n; ++q) {
            const bool isMax = isMax0 && q == 0;
            double v = 0.0;
            for (int b = tid; b < nb; b += RB_TPB) {
                const double w = __ldcg(pb + q * MAX_RED_BLOCKS + b);
                v = isMax ? fmax(v, w) : v + w;
            }
            v = block_red(v, isMax, 0.0);
            if (tid == 0) s.red[q] = v;
        }
        __syncthreads();
        buf ^= 1;
    };
    auto nacc = [&](double acc, double v) {
        if (kind == 0) return fmax(acc, fabs(v));
        if (kind == 1) return acc + fabs(v);
        return acc + v * v;
    };

    double r[RB_NZ], rt[RB_NZ], e[RB_NZ], pv[RB_NZ], v[RB_NZ], phi[RB_NZ];
    double pc[RB_NZ], lp[RB_NZ];
    long long cidx[RB_NZ];
#pragma unroll
    for (int k = 0; k < RB_NZ; ++k) {
        r[k] = rt[k] = e[k] = pv[k] = v[k] = phi[k] = pc[k] = lp[k] = 0.0;
        cidx[k] = (T.own && k < nz) ? g.idx(T.x, T.y, k) : -1;
    }
    // norm(r) and <r~, r> of the register-resident vectors -> s.red[0], rhoNext
    auto norm_and_rho = [&]() {
        double acc = 0.0, rho = 0.0;
#pragma unroll
        for (int k = 0; k < RB_NZ; ++k)
            if (cidx[k] >= 0) {
                acc = nacc(acc, r[k]);
                rho += rt[k] * r[k];
            }
        reduce(acc, rho, 2, nmax);
        if (tid == 0) rhoNext = s.red[1];
    };
    // r = rhs - L phi, r~ = r, e = 0 from the pass's L(phi) in lp
    auto residual_from_lp = [&]() {
#pragma unroll
        for (int k = 0; k < RB_NZ; ++k)
            if (cidx[k] >= 0) {
                r[k] = f.rhs[cidx[k]] - lp[k];
                rt[k] = r[k];
                e[k] = 0.0;
            }
    };

    // ---- num_smooth_bottom relaxations and the true residual: one tile pass straight from the multigrid fields
    rb_tile_pass(g, m, sm, T, f.phi, f.rhs, false, a.presmooth, pc, lp, prof, tp);
#pragma unroll
    for (int k = 0; k < RB_NZ; ++k) phi[k] = pc[k];
    residual_from_lp();
    norm_and_rho();
    if (tid == 0) {
        const double n0 = norm_finish(s.red[0], kind);
        BicgState z = {};
        z.nrm[0] = z.nrm[1] = n0;
        z.initial_norm = (a.bottomMetric > 0) ? a.bottomMetric : n0;
        z.initial_rnorm = n0;
        z.eps = a.eps; z.reps = a.reps; z.small = a.small; z.one_minus_hang = a.one_minus_hang;
        z.imax = a.imax; z.init = 1; z.normKind = kind;
        z.status = loop_top(&z) ? 0 : 1;
        s = z;
    }
    __syncthreads();
    const bool pre = a.nprecond > 0;
    int restarts = 0;
    while (s.status == 0) {
        __syncthreads();   // every thread has read the loop condition before ctl1 may change it
        if (tid == 0) {
            s.red[0] = rhoNext;
            ctl1(&s);
        }
        __syncthreads();
        if (s.status != 0) break;
        {   // p = r  or  ((beta p) + (-beta omega) v) + r;  staged for the neighbours' halos
            const bool init = s.init != 0;
            const double b = s.bet[1], cv = s.c_pv;
#pragma unroll
            for (int k = 0; k < RB_NZ; ++k)
                if (cidx[k] >= 0) {
                    double pn = r[k];
                    if (!init) {
                        const double t1 = pv[k] * b;
                        const double t2 = t1 + cv * v[k];
                        pn = t2 + 1.0 * pn;
                    }
                    pv[k] = pn;
                    f.pv[cidx[k]] = pn;
                }
            gsync();
        }
        // p~ = preCond(p) (diagonal scaling + relaxations), v = L p~
        rb_tile_pass(g, m, sm, T, f.pv, f.pv, pre, a.nprecond, pc, lp, prof, tp);
        {
            double acc = 0.0;
#pragma unroll
            for (int k = 0; k < RB_NZ; ++k)
                if (cidx[k] >= 0) {
                    v[k] = lp[k];
                    acc += rt[k] * v[k];
                }
            reduce(acc, 0.0, 1, false);
        }
        if (tid == 0) ctl2(&s);
        __syncthreads();
        {   // r -= alpha v, e += alpha p~ (or r = 0); the new r is staged; norm(r), next rho
            const bool on = s.doA != 0;
            const double na = s.neg_alpha, al = s.alph[0];
#pragma unroll
            for (int k = 0; k < RB_NZ; ++k)
                if (cidx[k] >= 0) {
                    double nv = 0.0;
                    if (on) {
                        nv = r[k] + na * v[k];
                        e[k] = e[k] + al * pc[k];
                    }
                    r[k] = nv;
                    f.r[cidx[k]] = nv;
                }
            norm_and_rho();   // its grid barrier also publishes the staged r
        }
        if (tid == 0) ctl3(&s);
        __syncthreads();
        if (s.doB) {
            // s~ = preCond(r), t = L s~
            rb_tile_pass(g, m, sm, T, f.r, f.r, pre, a.nprecond, pc, lp, prof, tp);
            {
                double a0 = 0.0, a1 = 0.0;
#pragma unroll
                for (int k = 0; k < RB_NZ; ++k)
                    if (cidx[k] >= 0) {
                        a0 += lp[k] * r[k];
                        a1 += lp[k] * lp[k];
                    }
                reduce(a0, a1, 2, false);
            }
            if (tid == 0) ctl4(&s);
            __syncthreads();
            {
                const double om = s.omega[0], no = s.neg_omega;
#pragma unroll
                for (int k = 0; k < RB_NZ; ++k)
                    if (cidx[k] >= 0) {
                        e[k] = e[k] + om * pc[k];
                        r[k] = r[k] + no * lp[k];
                    }
                norm_and_rho();
            }
        }
        if (tid == 0) ctl5(&s);
        __syncthreads();
        if (s.status == 2) {   // restart: fold the correction into phi and begin again from the true residual
#pragma unroll
            for (int k = 0; k < RB_NZ; ++k)
                if (cidx[k] >= 0) {
                    phi[k] = phi[k] + 1.0 * e[k];
                    e[k] = 0.0;
                }
            if (restarts == a.numRestarts) break;
#pragma unroll
            for (int k = 0; k < RB_NZ; ++k)
                if (cidx[k] >= 0) f.pt[cidx[k]] = phi[k];   // staged in p~'s array (not otherwise used by this kernel)
            gsync();
            rb_tile_pass(g, m, sm, T, f.pt, f.rhs, false, 0, pc, lp, prof, tp);
            residual_from_lp();
            norm_and_rho();
            if (tid == 0) {
                s.nrm[0] = norm_finish(s.red[0], kind);
                for (int q = 0; q < 4; ++q) s.rho[q] = 0.0;
                s.alph[0] = s.alph[1] = s.bet[0] = s.bet[1] = s.omega[0] = s.omega[1] = 0.0;
                s.init = 1;
                s.status = loop_top(&s) ? 0 : 1;
            }
            ++restarts;
            __syncthreads();
        }
    }
    // phi += e.  Every read of f.phi by another CTA (the first tile pass) lies behind at least one grid barrier.
#pragma unroll
    for (int k = 0; k < RB_NZ; ++k)
        if (cidx[k] >= 0) f.phi[cidx[k]] = phi[k] + 1.0 * e[k];
    RB_PROF(4, tp);
    if (prof && blockIdx.x == 0 && tid == 0) g_rb_prof[5] += s.i;
    if (blockIdx.x == 0 && tid == 0) *itersAccum += s.i;
}

// ---- eligibility and launch ---------------------------------------------------------------------------------
// single-rank diagonal-metric geometry g (the rank's own bottom depth, or the gathered one of bottom_gather.cu)
bool bottom_resident_usable(SomarLevel* L, const Geom& g, const Metric& m) {
    if (L->bottom_mode != 3) return false;
    const SomarSolverParams& p = L->prm;
    if (!m.diagonal || g.nd != 3 || g.n[2] > RB_NZ) return false;
    if (g.periodic[0] || g.periodic[1] || g.periodic[2] || g.nbr[0] >= 0 || g.nbr[1] >= 0) return false;
    if (p.relaxMode != SOMAR_RELAX_LEVEL_GSRB || p.precondMode == 2) return false;
    const int nprecond = (p.num_smooth_precond == 0 || p.precondMode == 0) ? 0 : p.num_smooth_precond;
    if (2 * nprecond + 1 > RB_H || 2 * p.num_smooth_bottom + 1 > RB_H || p.num_smooth_bottom < 0) return false;
    if ((long long)g.gn[0] * g.gn[1] * g.gn[2] == 1) return false;
    if (g.gh[0] < 1 || g.gh[1] < 1 || g.gh[2] < 1) return false;   // the high faces sit in the first ghost slot
    if (L->bottom_resident_blocks < 0) {
        int dev = 0, coop = 0, perSm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        if (cudaFuncSetAttribute(k_bicg_resident, cudaFuncAttributeMaxDynamicSharedMemorySize, RB_SMEM_BYTES) != cudaSuccess) {
            (void)cudaGetLastError();
            coop = 0;
        }
        if (coop) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_bicg_resident, RB_TPB, RB_SMEM_BYTES);
        L->bottom_resident_blocks = coop ? perSm * L->cx.sms : 0;
    }
    const int tiles = ((g.n[0] + RB_CX - 1) / RB_CX) * ((g.n[1] + RB_CY - 1) / RB_CY);
    return tiles <= L->bottom_resident_blocks && tiles <= MAX_RED_BLOCKS;
}

int launch_bottom_resident(SomarLevel* L, const Geom& g_, const Metric& m_, double* phi, double* rhs,
                           double* const work[8], void* bicgBar, double* part, int* itersAccum) {
    const SomarSolverParams& p = L->prm;
    BicgFields f;
    f.phi = phi; f.rhs = rhs;
    f.r = work[0]; f.rt = work[1]; f.e = work[2]; f.pv = work[3]; f.pt = work[4];
    f.stl = work[5]; f.t = work[6]; f.v = work[7];
    BicgArgs a;
    a.presmooth = p.num_smooth_bottom;
    a.nprecond = (p.num_smooth_precond == 0 || p.precondMode == 0) ? 0 : p.num_smooth_precond;
    a.numRestarts = p.bottom_numRestarts;
    a.bottomMetric = L->bottomMetric;
    a.eps = L->bottomEps; a.reps = p.bottom_reps; a.small = p.bottom_small; a.one_minus_hang = 1.0 - p.bottom_hang;
    a.imax = p.bottom_imax;
    a.normKind = (p.bottom_normType == 0) ? 0 : (p.bottom_normType == 1 ? 1 : 2);
    a.tcx = a.tcy = 0;
    static const bool profOn = []() { const char* e = getenv("SOMAR_B200_BOTTOM_PROF"); return e && e[0] == '1'; }();
    if (profOn) a.tcx = 1234;
    Geom g = g_;
    Metric m = m_;
    int tilesX = (g.n[0] + RB_CX - 1) / RB_CX;
    const int tiles = tilesX * ((g.n[1] + RB_CY - 1) / RB_CY);
    // SOMAR_B200_BOTTOM_CLUSTER=8|4|2: thread-block clusters for the grid barrier (gsync in the kernel) -- the largest
    // cluster size <= the request that divides the tile count and whose clusters are all co-resident.  OFF by default:
    // measured on B200 at 256x256x4 (2059 iterations per solve) 75.6 us per iteration with clusters of 8 or 4 against
    // 70.8 with cooperative groups' barrier -- as with round 1's counter tree, the barrier itself is not what the
    // iteration waits for.
    int csize = 0;
    {
        const char* e = getenv("SOMAR_B200_BOTTOM_CLUSTER");
        const int want = e ? atoi(e) : 0;
        for (int c = want; c >= 2 && !csize; c >>= 1) {
            if (tiles % c) continue;
            cudaLaunchConfig_t q = {};
            q.gridDim = dim3(tiles); q.blockDim = dim3(RB_TPB); q.dynamicSmemBytes = RB_SMEM_BYTES;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = c; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            q.attrs = at; q.numAttrs = 1;
            int nclusters = 0;
            if (cudaOccupancyMaxActiveClusters(&nclusters, k_bicg_resident, &q) == cudaSuccess && nclusters * c >= tiles) csize = c;
            else (void)cudaGetLastError();
        }
    }
    unsigned long long* gbar = nullptr;
    if (csize) {
        gbar = static_cast<unsigned long long*>(bicgBar);
        SOMAR_CUDA(cudaMemsetAsync(gbar, 0, sizeof(unsigned long long), L->cx.stream));
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(tiles); cfg.blockDim = dim3(RB_TPB); cfg.dynamicSmemBytes = RB_SMEM_BYTES; cfg.stream = L->cx.stream;
        cudaLaunchAttribute at[2];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = csize; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        at[1].id = cudaLaunchAttributeCooperative;
        at[1].val.cooperative = 1;
        cfg.attrs = at; cfg.numAttrs = 2;
        SOMAR_CUDA(cudaLaunchKernelEx(&cfg, k_bicg_resident, g, m, f, a, part, itersAccum, tilesX, gbar));
    } else {
        void* args[] = {&g, &m, &f, &a, &part, &itersAccum, &tilesX, &gbar};
        SOMAR_CUDA(cudaLaunchCooperativeKernel((const void*)k_bicg_resident, dim3(tiles), dim3(RB_TPB), args, RB_SMEM_BYTES,
                                               L->cx.stream));
    }
    L->cx.launches++;
    L->bottom_iters_pending = true;
    if (profOn) {   // development probe: synchronises
        static int visits = 0;
        if (++visits % 5 == 0) {
            unsigned long long h[8];
            cudaStreamSynchronize(L->cx.stream);
            cudaMemcpyFromSymbol(h, g_rb_prof, sizeof(h));
            const double it = h[5] ? (double)h[5] : 1.0;
            fprintf(stderr, "k_bicg_resident probe (block 0, per iteration, us): load %.2f  colour passes %.2f  apply %.2f  "
                            "grid barriers %.2f  rest %.2f  (%llu iterations)\n",
                    h[0] / it * 1e-3, h[1] / it * 1e-3, h[2] / it * 1e-3, h[3] / it * 1e-3, h[4] / it * 1e-3, h[5]);
        }
    }
    return SOMAR_OK;
}

}  // namespace somar

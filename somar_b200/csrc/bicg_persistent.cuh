// bicg_persistent.cuh -- the whole bottom visit of the V-cycle as ONE persistent cooperative kernel.
//
// MappedMultiGrid<T>::cycle at the bottom depth (src/calculus/AMRElliptic/MappedMultiGrid.H:566-573) is
// num_smooth_bottom relaxations followed by Chombo's BiCGStabSolver (bicgstab.cu has the algorithm's provenance),
// whose every iteration is two preconditioner applications (diagonal scaling + point GSRB relaxations,
// MappedAMRPoissonOp.cpp:684-734), two operator applications and five global reductions.  The bottom depth holds
// a few hundred thousand cells -- L2 resident, a microsecond of work per pass -- so as separate kernels the
// iteration is launch latency, and even as one kernel it is INSTRUCTION latency unless the passes are spread over
// many warps.  Hence:
//   * dependent passes are separated by grid.sync() instead of kernel boundaries (15 per iteration);
//   * a pass hands 32-cell pieces of rows ("items") to the warps of the whole grid round-robin, one cell per
//     lane, with as many co-resident warps as the register budget allows (__launch_bounds__(256, 4));
//     colour passes use 64-cell items (lane l owns cell 2l + parity) so that no lane idles;
//   * every block keeps a copy of the scalar state in shared memory and advances it with the same ctl1..ctl5
//     steps as the launch-per-step versions;
//   * the sums keep the order of k_reduce / finish_reduce: the light passes (axpy-like updates, dot products,
//     norms) run in the "canonical" mapping -- the first cb blocks iterate exactly as a k_reduce grid of cb
//     blocks would -- and every block folds the cb partials in finish_reduce's order.  The unconverged BiCGStab
//     of a singular bottom problem amplifies rounding differences enough to move the V-cycle count, so this is
//     what keeps the solve bit-identical to the launch-per-step modes (and to the parity established for them);
//   * <r_tilde, r> of the next iteration is accumulated by the pass that produces r.
// The fields are deliberately NOT const/__restrict__: they are written and re-read across grid.sync() by other
// SMs, so no load may go through the non-coherent path.
// Usable for: single rank, point GSRB smoother/preconditioner, no periodic direction.  The full metric (FULL) runs the
// pass-per-barrier formulation with the cell statements of stencil_full.cuh dispatched by cell kind; its colour passes are
// out of place (the cross terms read the pre-pass state of BOTH colours) and ping-pong with the t field, which is free
// whenever a relaxation runs.
#pragma once
#include "cells.cuh"
#include "stencil_ortho.cuh"
#include "stencil_full.cuh"
#include <cooperative_groups.h>

namespace somar {

namespace cg = cooperative_groups;

struct BicgFields {
    double *phi, *rhs;                          // the bottom depth's correction and residual (MG fields)
    double *r, *rt, *e, *pv, *pt, *stl, *t, *v; // BiCGStab work fields
};
struct BicgArgs {
    int presmooth;       // num_smooth_bottom
    int nprecond;        // relaxations in preCond (0: preCond is a copy)
    int numRestarts;
    double bottomMetric; // > 0: convergence metric set by the outer solver
    double eps, reps, small, one_minus_hang;
    int imax, normKind;
    int tcx, tcy;        // core of the shared-memory tile of the fused preconditioner+operator pass (0: not used)
};

// ---- scalar state of the iteration and its steps (shared with the launch-per-step version in bicgstab.cu) ----
struct BicgState {
    double red[4];                 // landing slots of the reductions
    double rho[4], nrm[2], alph[2], bet[2], omega[2];
    double c_pv;                   // -beta * omega_old        (coefficient of v in the p update)
    double neg_alpha, neg_omega;
    double initial_norm, initial_rnorm, eps, reps, small, one_minus_hang;
    int i, imax, init, recount;
    int status;                    // 0 iterating, 1 finished, 2 restart requested
    int doA, doB, normKind;
};

__device__ __forceinline__ double norm_finish(double v, int kind) { return kind == 2 ? sqrt(v) : v; }

// ---- the scalar steps of one iteration (shared by the one-thread control kernels and the persistent kernel) ----
// after  red[0] = <r_tilde, r>
__device__ __forceinline__ void ctl1(BicgState* st) {
    ++st->i;
    st->nrm[1] = st->nrm[0];
    st->alph[1] = st->alph[0]; st->bet[1] = st->bet[0]; st->omega[1] = st->omega[0];
    st->rho[3] = st->rho[2]; st->rho[2] = st->rho[1];
    st->rho[1] = st->red[0];
    if (st->rho[1] == 0.0) { st->status = 1; return; }
    if (!st->init) {
        st->bet[1] = (st->rho[1] / st->rho[2]) * (st->alph[1] / st->omega[1]);
        st->c_pv = -st->bet[1] * st->omega[1];
    }
}
// after  red[0] = <r_tilde, v>
__device__ __forceinline__ void ctl2(BicgState* st) {
    st->init = 0;
    const double m = st->red[0];
    st->alph[0] = st->rho[1] / m;
    st->neg_alpha = -st->alph[0];
    st->doA = fabs(m) > st->small * fabs(st->rho[1]);
}
// after the alpha step's norm landed in red[0]
__device__ __forceinline__ void ctl3(BicgState* st) {
    st->nrm[0] = st->doA ? norm_finish(st->red[0], st->normKind) : 0.0;
    st->doB = st->nrm[0] > st->eps * st->initial_norm && st->nrm[0] > st->reps * st->initial_rnorm;
}
// after  red[0] = <t, r>, red[1] = <t, t>
__device__ __forceinline__ void ctl4(BicgState* st) {
    if (!st->doB) return;
    st->omega[0] = st->red[0] / st->red[1];
    st->neg_omega = -st->omega[0];
}
// end of the iteration: convergence, hang / restart bookkeeping, and the loop condition of the next one
__device__ __forceinline__ bool loop_top(const BicgState* st) {
    return (st->i < st->imax && st->nrm[0] > st->eps * st->nrm[1]) && (st->nrm[1] > 0);
}
__device__ __forceinline__ void ctl5(BicgState* st) {
    if (st->doB) st->nrm[0] = norm_finish(st->red[0], st->normKind);
    const double n0 = st->nrm[0], n1 = st->nrm[1];
    if (n0 <= st->eps * st->initial_norm || n0 <= st->reps * st->initial_rnorm) { st->status = 1; return; }
    if (st->omega[0] == 0.0 || n0 > st->one_minus_hang * n1) {
        if (st->recount == 0) {
            st->recount = 1;
        } else {
            st->recount = 0;
            st->status = 2;
            return;
        }
    }
    if (!loop_top(st)) st->status = 1;
}
#define BP_TPB 256
#ifndef BP_MINBLOCKS
#define BP_MINBLOCKS 4
#endif

// items of W cells of one row; i0 = first cell of the item.  j, k, i0 visible in BODY.
#define BP_FOR_ITEMS(g, W, BODY)                                                                           \
    {                                                                                                      \
        const int segs_ = ((g).n[0] + (W) - 1) / (W);                                                      \
        const int items_ = (g).n[1] * (g).n[2] * segs_;                                                    \
        for (int it_ = gwarp; it_ < items_; it_ += nwarps) {                                               \
            const int r_ = it_ / segs_;                                                                    \
            const int i0 = (it_ - r_ * segs_) * (W);                                                       \
            const int k = r_ / (g).n[1];                                                                   \
            const int j = r_ - k * (g).n[1];                                                               \
            BODY                                                                                           \
        }                                                                                                  \
    }
// one cell per lane
#define BP_FOR_CELLS(g, c, BODY)                                                                           \
    BP_FOR_ITEMS(g, 32, {                                                                                  \
        const int i = i0 + lane;                                                                           \
        if (i < (g).n[0]) {                                                                                \
            const long long c = (g).idx(i, j, k);                                                          \
            BODY                                                                                           \
        }                                                                                                  \
    })

// The "canonical" iteration: exactly FOR_VALID_CELLSN as a grid of cb blocks of TPB threads would run it (the
// launch shape of k_reduce and of the BLAS-1 kernels), executed by the first cb blocks of this grid.  Sums
// accumulated in it, block-reduced and folded over cb partials are bit-identical to k_reduce / finish_reduce.
#define BP_CANON_CELLS(g, cb, c, u, U, LOAD, USE)                                                          \
    if ((int)blockIdx.x < (cb)) {                                                                          \
        const long long cwarp_ = (long long)blockIdx.x * (BP_TPB >> 5) + (threadIdx.x >> 5);               \
        const long long cnw_ = (long long)(cb) * (BP_TPB >> 5);                                            \
        const long long rows_ = (long long)(g).n[1] * (g).n[2];                                            \
        for (long long r_ = cwarp_; r_ < rows_; r_ += cnw_) {                                              \
            const int k = (int)(r_ / (g).n[1]);                                                            \
            const int j = (int)(r_ - (long long)k * (g).n[1]);                                             \
            const long long row0_ = (g).idx(0, j, k);                                                      \
            for (int i0_ = lane; i0_ < (g).n[0]; i0_ += 32 * (U)) {                                        \
                _Pragma("unroll") for (int u = 0; u < (U); ++u) {                                          \
                    const int i = i0_ + 32 * u;                                                            \
                    const long long c = row0_ + i;                                                         \
                    (void)c;                                                                               \
                    if (i < (g).n[0]) { LOAD }                                                             \
                }                                                                                          \
                _Pragma("unroll") for (int u = 0; u < (U); ++u) {                                          \
                    const int i = i0_ + 32 * u;                                                            \
                    const long long c = row0_ + i;                                                         \
                    (void)c;                                                                               \
                    if (i < (g).n[0]) { USE }                                                              \
                }                                                                                          \
            }                                                                                              \
        }                                                                                                  \
    }
// Grid-wide barrier for a cooperatively launched (co-resident) grid: arrivals are spread over BP_BAR_GROUPS padded
// counters, the last arrival of a group bumps a top counter, the last group publishes the epoch; everybody polls one
// flag.  Counters only grow (zeroed by the host before the launch), so there is no reset race.  Same fence /
// __syncthreads structure as cooperative groups' software barrier, i.e. the same visibility guarantees.
#define BP_BAR_GROUPS 8
#define BP_BAR_WORDS (16 * (BP_BAR_GROUPS + 2))
__device__ __forceinline__ void tree_barrier(unsigned long long* bar, unsigned& epoch) {
    __syncthreads();
    if (threadIdx.x == 0) {
        ++epoch;
        const unsigned nb = gridDim.x;
        const unsigned grp = blockIdx.x % BP_BAR_GROUPS;
        const unsigned members = (nb - grp + BP_BAR_GROUPS - 1) / BP_BAR_GROUPS;
        const unsigned ngroups = nb < BP_BAR_GROUPS ? nb : BP_BAR_GROUPS;
        volatile unsigned long long* flag = bar + 16 * (BP_BAR_GROUPS + 1);
        __threadfence();
        const unsigned long long old = atomicAdd(bar + 16 * grp, 1ULL);
        if (old + 1 == (unsigned long long)members * epoch) {
            __threadfence();
            const unsigned long long t = atomicAdd(bar + 16 * BP_BAR_GROUPS, 1ULL);
            if (t + 1 == (unsigned long long)ngroups * epoch) {
                __threadfence();
                *flag = epoch;
            }
        }
        const long long t0 = clock64();
        while (*flag < epoch) {
            if (clock64() - t0 > (20ll << 30)) __trap();   // ~10 s: a bug must not hang the GPU
        }
        __threadfence();
    }
    __syncthreads();
}

#define BP_U 8

// ---- fused "relaxations + operator" pass on shared-memory tiles ------------------------------------------------
// preCond's nprecond red-black relaxations followed by L(.) are 2*nprecond + 1 dependent stencil passes, each a
// grid barrier in the pass-per-barrier formulation.  Here a block loads a tile of the input with a halo of
// h = 2*nprecond + 1 cells (all z-planes), relaxes it in shared memory -- every colour pass is valid on a region one
// cell smaller, block barriers only -- and applies the operator on the core: ONE grid barrier instead of 2*nprecond+1.
// Halo cells are relaxed redundantly by the neighbouring blocks with the statements, operands and order of the
// owning block (gsrb_update_vals / apply_vals_ortho3 are the value forms of gsrb_cell_ortho / apply_cell_ortho), so
// the result is bit-identical to the pass-per-barrier version.  Input and output are different arrays (a block
// reads halo cells that another block owns).  3-D only; the tile must hold (tcx+2h) x (tcy+2h) x nz cells.
#define BP_TILE_CAP 1872   // 26 x 18 x 4
__device__ __forceinline__ void bp_tile_relax_apply(const Geom& g, const Metric& m, double* __restrict__ tile,
                                                    const double* __restrict__ in, const double* __restrict__ rhs,
                                                    double* __restrict__ outPhi, double* __restrict__ outL, int nrelax,
                                                    int tcx, int tcy) {
    const int h = 2 * nrelax + 1;
    const int nx = g.n[0], ny = g.n[1], nz = g.n[2];
    const int W = tcx + 2 * h, Hh = tcy + 2 * h;
    const int tilesX = (nx + tcx - 1) / tcx, tilesY = (ny + tcy - 1) / tcy;
    const int tid = threadIdx.x;
    for (int t = blockIdx.x; t < tilesX * tilesY; t += gridDim.x) {
        const int ty = t / tilesX, tx = t - ty * tilesX;
        const int cx0 = tx * tcx, cy0 = ty * tcy;              // core origin
        const int ox = cx0 - h, oy = cy0 - h;                  // tile origin (may lie outside the domain)
        __syncthreads();                                       // the previous tile's readers are done
        {   // load core + h, clipped to the domain
            const int xa = max(0, ox), xb = min(nx, cx0 + tcx + h), ya = max(0, oy), yb = min(ny, cy0 + tcy + h);
            const int wx = xb - xa, wy = yb - ya;
            for (int q = tid; q < wx * wy * nz; q += BP_TPB) {
                const int k = q / (wx * wy), r = q - k * wx * wy;
                const int j = ya + r / wx, i = xa + (r - (r / wx) * wx);
                tile[(i - ox) + W * ((j - oy) + Hh * k)] = in[g.idx(i, j, k)];
            }
        }
        __syncthreads();
        int sdone = 0;
        for (int it = 0; it < nrelax; ++it)
            for (int pass = 0; pass < 2; ++pass) {
                ++sdone;
                const int e = h - sdone;                       // this colour pass is valid on core + e
                const int xa = max(0, cx0 - e), xb = min(nx, cx0 + tcx + e), ya = max(0, cy0 - e), yb = min(ny, cy0 + tcy + e);
                const int half = (xb - xa + 1) >> 1, wy = yb - ya;
                for (int q = tid; q < half * wy * nz; q += BP_TPB) {
                    const int k = q / (half * wy), r = q - k * half * wy;
                    const int j = ya + r / half;
                    const int par = (g.g0[0] + g.g0[1] + j + g.g0[2] + k + pass) & 1;
                    const int i = xa + ((xa ^ par) & 1) + 2 * (r - (r / half) * half);
                    if (i < xb) {
                        const int l = (i - ox) + W * ((j - oy) + Hh * k);
                        double nb[3][2];
                        nb[0][0] = tile[l - 1];
                        nb[0][1] = tile[l + 1];
                        nb[1][0] = tile[l - W];
                        nb[1][1] = tile[l + W];
                        nb[2][0] = (k > 0) ? tile[l - W * Hh] : 0.0;
                        nb[2][1] = (k < nz - 1) ? tile[l + W * Hh] : 0.0;
                        const GsrbCoef qc = gsrb_load_coef<3>(g, m, rhs, g.idx(i, j, k));
                        tile[l] = gsrb_update_vals<3>(g, qc, i, j, k, tile[l], nb);
                    }
                }
                __syncthreads();
            }
        {   // the operator on the core, and the relaxed values
            const int xb = min(nx, cx0 + tcx), yb = min(ny, cy0 + tcy);
            const int wx = xb - cx0, wy = yb - cy0;
            for (int q = tid; q < wx * wy * nz; q += BP_TPB) {
                const int k = q / (wx * wy), r = q - k * wx * wy;
                const int j = cy0 + r / wx, i = cx0 + (r - (r / wx) * wx);
                const int l = (i - ox) + W * ((j - oy) + Hh * k);
                const long long c = g.idx(i, j, k);
                const double pc = tile[l];
                double nb[3][2];
                nb[0][0] = tile[l - 1];
                nb[0][1] = tile[l + 1];
                nb[1][0] = tile[l - W];
                nb[1][1] = tile[l + W];
                nb[2][0] = (k > 0) ? tile[l - W * Hh] : 0.0;
                nb[2][1] = (k < nz - 1) ? tile[l + W * Hh] : 0.0;
                const double glo[3] = {m.jg[0][0][c], m.jg[1][1][c], m.jg[2][2][c]};
                const double ghi[3] = {m.jg[0][0][c + 1], m.jg[1][1][c + g.sy], m.jg[2][2][c + g.sz]};
                outL[c] = apply_vals_ortho3(g, side_mask(g, i, j, k), true, pc, nb, glo, ghi, m.jinv[c]);
                outPhi[c] = pc;
            }
        }
    }
}

// ---- full metric: one colour pass (out of place) and the operator, one copy of each in the kernel -------------------
#ifndef BP_MINBLOCKS_FULL
#define BP_MINBLOCKS_FULL 2
#endif
template <int ND>
__device__ __noinline__ void bp_full_colour_pass(const Geom& g, const Metric& m, double* out, const double* in,
                                                 const double* rhs, int pass, int gwarp, int nwarps, int lane) {
    BP_FOR_ITEMS(g, 64, {
        const int par = (g.g0[0] + g.g0[1] + j + ((ND > 2) ? (g.g0[2] + k) : 0) + pass) & 1;
        const int ia = i0 + 2 * lane;
        const int i = ia + par;                                  // the lane's cell of this pass's colour
        if (ia < g.n[0]) {
            const long long ca = g.idx(ia, j, k);
            double va = in[ca];
            double vb = (ia + 1 < g.n[0]) ? in[ca + 1] : 0.0;
            if (i < g.n[0]) {
                const double v = gsrb_cell_full_kind<ND>(g, m, in, rhs, i, j, k, ca + par, cell_kind<ND>(g, i, j, k));
                if (par) vb = v; else va = v;
            }
            out[ca] = va;
            if (ia + 1 < g.n[0]) out[ca + 1] = vb;
        }
    })
}
template <int ND>
__device__ __noinline__ void bp_full_apply(const Geom& g, const Metric& m, double* out, const double* in, int gwarp,
                                           int nwarps, int lane) {
    BP_FOR_CELLS(g, c, { out[c] = apply_cell_full_kind<ND>(g, m, in, i, j, k, c, true, cell_kind<ND>(g, i, j, k)); })
}

template <int ND, bool FULL = false>
__global__ void __launch_bounds__(BP_TPB, FULL ? BP_MINBLOCKS_FULL : BP_MINBLOCKS)
    k_bicg_persistent(const Geom g, const Metric m, const BicgFields f, const BicgArgs a, double* part,
                      int* itersAccum, const int cb, unsigned long long* bar) {
    cg::grid_group grid = cg::this_grid();
    unsigned epoch = 0;
    // bar != nullptr: the two-level counter barrier below instead of cg::grid_group::sync() (one atomic word for all
    // blocks: ~600 serialised arrivals per barrier at 4 blocks/SM)
    auto gsync = [&]() {
        if (bar) tree_barrier(bar, epoch);
        else grid.sync();
    };
    __shared__ BicgState s;
    __shared__ double rhoNext;   // <r_tilde, r> formed by the pass that wrote r
    __shared__ double tile[(ND > 2 && !FULL) ? BP_TILE_CAP : 1];
    // fused preconditioner + operator (bp_tile_relax_apply): the diagonally scaled input of preCond(p) is staged in
    // stl's array and that of preCond(s) in pt's array -- each is free at that point of the iteration
    const bool tiled = !FULL && (ND > 2) && a.tcx > 0 && a.nprecond > 0;
    int buf = 0;
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int gwarp = blockIdx.x * (BP_TPB / 32) + (tid >> 5);
    const int nwarps = gridDim.x * (BP_TPB / 32);
    const int kind = a.normKind;
    const bool nmax = kind == 0;

    // Two per-thread partial results -> s.red[0..n-1] in every block.  Lane 0 is a sum or (max norm) a maximum,
    // lane 1 always a sum.  The partials of consecutive reductions alternate between two buffers: a block can be
    // one grid.sync() ahead of another, never two.
    auto reduce = [&](double p0, double p1, int n, bool isMax0) {
        p0 = block_red(p0, isMax0, 0.0);
        if (n > 1) p1 = block_red(p1, false, 0.0);
        double* pb = part + (size_t)buf * 2 * MAX_RED_BLOCKS;
        if (tid == 0 && (int)blockIdx.x < cb) {
            pb[blockIdx.x] = p0;
            if (n > 1) pb[MAX_RED_BLOCKS + blockIdx.x] = p1;
        }
        gsync();
        for (int q = 0; q < n; ++q) {
            const bool isMax = isMax0 && q == 0;
            double v = 0.0;
            for (int b = tid; b < cb; b += BP_TPB) {
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
    // norm(r) and <r_tilde, r> in the canonical order (r complete on entry) -> s.red[0], rhoNext
    auto norm_and_rho = [&]() {
        double acc = 0.0, rho = 0.0;
        double tv[BP_U], rv[BP_U];
        BP_CANON_CELLS(g, cb, c, u, BP_U, { tv[u] = f.rt[c]; rv[u] = f.r[c]; }, {
            acc = nacc(acc, rv[u]);
            rho += tv[u] * rv[u];
        })
        reduce(acc, rho, 2, nmax);
        if (tid == 0) rhoNext = s.red[1];
    };
    // one colour of point Gauss-Seidel, in place (LevelGSRB::relax, GSRB.cpp:58-98): lane l of a 64-cell item owns
    // the cell 2l + parity
    auto colour_pass = [&](double* phi, const double* rhs, int pass) {
        BP_FOR_ITEMS(g, 64, {
            const int par = (g.g0[0] + g.g0[1] + j + ((ND > 2) ? (g.g0[2] + k) : 0) + pass) & 1;
            const int i = i0 + 2 * lane + par;
            if (i < g.n[0]) {
                const long long c = g.idx(i, j, k);
                phi[c] = gsrb_cell_ortho<ND>(g, m, phi, rhs, i, j, k, c);
            }
        })
    };
    auto relax = [&](double* phi, const double* rhs, int iters) {   // enters and leaves with phi complete
        if (FULL) {   // two passes per iteration: the result is back in phi
            double* src = phi;
            double* dst = f.t;
            for (int ps = 0; ps < 2 * iters; ++ps) {
                bp_full_colour_pass<ND>(g, m, dst, src, rhs, ps & 1, gwarp, nwarps, lane);
                gsync();
                double* sw = src; src = dst; dst = sw;
            }
        } else {
            for (int it = 0; it < iters; ++it) {
                colour_pass(phi, rhs, 0);
                gsync();
                colour_pass(phi, rhs, 1);
                gsync();
            }
        }
    };
    // r = rhs - L phi, r_tilde = r, e = 0, norm(r), <r, r>      (phi complete on entry)
    auto true_residual = [&]() {
        if (FULL) bp_full_apply<ND>(g, m, f.r, f.phi, gwarp, nwarps, lane);   // (same cell, same thread below)
        BP_FOR_CELLS(g, c, {
            const double rv = f.rhs[c] - (FULL ? f.r[c] : apply_cell_ortho<ND>(g, m, f.phi, i, j, k, c, true));
            f.r[c] = rv;
            f.rt[c] = rv;
            f.e[c] = 0.0;
        })
        gsync();
        norm_and_rho();
    };

    relax(f.phi, f.rhs, a.presmooth);
    true_residual();
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
    int restarts = 0;
    while (s.status == 0) {
        __syncthreads();   // every thread has read the loop condition before ctl1 may change it
        if (tid == 0) {   // rho = <r_tilde, r> came with the pass that produced r
            s.red[0] = rhoNext;
            ctl1(&s);
        }
        __syncthreads();
        if (s.status != 0) break;
        {   // p = r  or  ((beta p) + (-beta omega) v) + r;   p_tilde = preCond(p): diagonal scaling, then relaxations
            const bool init = s.init != 0;
            const double b = s.bet[1], cv = s.c_pv;
            BP_FOR_CELLS(g, c, {
                double pn = f.r[c];
                if (!init) {
                    const double t1 = f.pv[c] * b;
                    const double t2 = t1 + cv * f.v[c];
                    pn = t2 + 1.0 * pn;
                }
                f.pv[c] = pn;
                (tiled ? f.stl : f.pt)[c] = a.nprecond ? pn / (g.alpha + g.beta * m.lapdiag[c]) : pn;
            })
            gsync();
            if (!tiled) relax(f.pt, f.pv, a.nprecond);
        }
        {   // v = L p_tilde (spread over the grid), then <r_tilde, v> in the canonical order
            if (tiled) {
                if (!FULL && ND > 2) bp_tile_relax_apply(g, m, tile, f.stl, f.pv, f.pt, f.v, a.nprecond, a.tcx, a.tcy);
            } else {
                if (FULL) bp_full_apply<ND>(g, m, f.v, f.pt, gwarp, nwarps, lane);
                else BP_FOR_CELLS(g, c, { f.v[c] = apply_cell_ortho<ND>(g, m, f.pt, i, j, k, c, true); })
            }
            gsync();
            double acc = 0.0;
            double tv[BP_U], vv[BP_U];
            BP_CANON_CELLS(g, cb, c, u, BP_U, { tv[u] = f.rt[c]; vv[u] = f.v[c]; }, { acc += tv[u] * vv[u]; })
            reduce(acc, 0.0, 1, false);
        }
        if (tid == 0) ctl2(&s);
        __syncthreads();
        {   // r -= alpha v, e += alpha p_tilde, s_tilde = diagonal scaling of the new r;  then norm(r), next rho
            const bool on = s.doA != 0;
            const double na = s.neg_alpha, al = s.alph[0];
            BP_FOR_CELLS(g, c, {
                double nv = 0.0;
                if (on) {
                    nv = f.r[c] + na * f.v[c];
                    f.e[c] = f.e[c] + al * f.pt[c];
                }
                f.r[c] = nv;
                (tiled ? f.pt : f.stl)[c] = a.nprecond ? nv / (g.alpha + g.beta * m.lapdiag[c]) : nv;
            })
            gsync();
            norm_and_rho();
        }
        if (tid == 0) ctl3(&s);
        __syncthreads();
        if (s.doB) {
            if (!tiled) relax(f.stl, f.r, a.nprecond);
            {   // t = L s_tilde (spread over the grid), then <t, r>, <t, t> in the canonical order
                if (tiled) {
                    if (!FULL && ND > 2) bp_tile_relax_apply(g, m, tile, f.pt, f.r, f.stl, f.t, a.nprecond, a.tcx, a.tcy);
                } else {
                    if (FULL) bp_full_apply<ND>(g, m, f.t, f.stl, gwarp, nwarps, lane);
                    else BP_FOR_CELLS(g, c, { f.t[c] = apply_cell_ortho<ND>(g, m, f.stl, i, j, k, c, true); })
                }
                gsync();
                double a0 = 0.0, a1 = 0.0;
                double tv[BP_U], rr[BP_U];
                BP_CANON_CELLS(g, cb, c, u, BP_U, { tv[u] = f.t[c]; rr[u] = f.r[c]; }, {
                    a0 += tv[u] * rr[u];
                    a1 += tv[u] * tv[u];
                })
                reduce(a0, a1, 2, false);
            }
            if (tid == 0) ctl4(&s);
            __syncthreads();
            {   // e += omega s_tilde, r -= omega t (norm, next rho)
                const double om = s.omega[0], no = s.neg_omega;
                BP_FOR_CELLS(g, c, {
                    f.e[c] = f.e[c] + om * f.stl[c];
                    f.r[c] = f.r[c] + no * f.t[c];
                })
                gsync();
                norm_and_rho();
            }
        }
        if (tid == 0) ctl5(&s);
        __syncthreads();
        if (s.status == 2) {   // restart: fold the correction into phi and begin again from the true residual
            {
                double pp[BP_U], ee[BP_U];
                BP_CANON_CELLS(g, cb, c, u, BP_U, { pp[u] = f.phi[c]; ee[u] = f.e[c]; }, {
                    f.phi[c] = pp[u] + 1.0 * ee[u];
                    f.e[c] = 0.0;
                })
            }
            if (restarts == a.numRestarts) break;
            gsync();
            true_residual();
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
    {   // phi += e   (e was last written in the canonical mapping; phi before the last grid.sync())
        double pp[BP_U], ee[BP_U];
        BP_CANON_CELLS(g, cb, c, u, BP_U, { pp[u] = f.phi[c]; ee[u] = f.e[c]; }, { f.phi[c] = pp[u] + 1.0 * ee[u]; })
    }
    if (blockIdx.x == 0 && tid == 0) *itersAccum += s.i;
}

}  // namespace somar

// line.cu -- batched vertical line relaxation (LineGSRB::relax, GSRB.cpp:148-330; LineGSRBIter{2,3}D,
// GSRBF.ChF:1529-2042).  Each red/black column solves  T x = B  with the tridiagonal
//   D_k  = alpha/Jinv + lapDiag_k,   DL_k = DU_k = Jg^{vv}(k+1/2) * beta/dz^2,
//   B_k  = -lphi_k + rhs/Jinv        (lphi = horizontal normal terms + all cross terms of the frozen state).
// The reference calls LAPACK dgtsv per column (GSRBF.ChF:1709,2022).  T does not depend on phi, and it is
// diagonally dominant, so dgtsv never pivots: its elimination  fact_k = DL_k / D'_k,  D'_{k+1} = D_{k+1} -
// fact_k DU_k  is done ONCE per (alpha, beta, metric) by k_line_factor and stored (fact, dmod); a relaxation
// pass is then the forward substitution  B'_{k+1} = B_{k+1} - fact_k B'_k  and the back substitution
// x_k = (B'_k - DU_k x_{k+1}) / D'_k  in dgtsv's own operation order.  One thread per column, lanes
// adjacent in x (coalesced planes), marching in k.
#include "kernels.hpp"
#include "stencil_full.cuh"

namespace somar {

#ifndef LINE_UNROLL
#define LINE_UNROLL 1
#endif
#ifndef LINE_MINB
#define LINE_MINB 8
#endif
#define LINE_PRAGMA_(x) _Pragma(#x)
#define LINE_PRAGMA(n) LINE_PRAGMA_(unroll n)

__device__ __forceinline__ double vert_coeff(int bct) { return (bct == BC_DIRI) ? 2.0 : 0.0; }

// diagonal entry of the column matrix at cell (i,j,k) (without alpha/Jinv) -- GSRBF.ChF:1596-1660 / 1800-1880
template <int ND>
__device__ __forceinline__ double line_lapdiag(const Geom& g, const Metric& m, int i, int j, int k, long long c,
                                               unsigned sm) {
    constexpr int VD = ND - 1;
    const long long stv = (VD == 1) ? g.sy : g.sz;
    const int kk = (VD == 1) ? j : k;
    const int N = g.n[VD];
    const double vv = g.beta * 1.0 / (g.dx[VD] * g.dx[VD]);
    const double gLo = m.jg[VD][VD][c], gHi = m.jg[VD][VD][c + stv];
    double lapDiag;
    if (kk == N - 1) lapDiag = -vv * (vert_coeff(g.bct[VD][1]) * gHi + gLo);
    else if (kk == 0) lapDiag = -vv * (gHi + vert_coeff(g.bct[VD][0]) * gLo);
    else lapDiag = -vv * (gLo + gHi);
    const long long st[2] = {1, g.sy};
#pragma unroll
    for (int h = 0; h < VD; ++h) {
        const double aa = g.beta * 1.0 / (g.dx[h] * g.dx[h]);
        const bool bLo = (sm >> (2 * h)) & 1u, bHi = (sm >> (2 * h + 1)) & 1u;
        if (!(bLo && g.bc[h][0] == BC_NEUM)) lapDiag = lapDiag - aa * m.jg[h][h][c];
        if (!(bHi && g.bc[h][1] == BC_NEUM)) lapDiag = lapDiag - aa * m.jg[h][h][c + st[h]];
    }
    return lapDiag;
}

template <int ND>
__global__ void k_line_factor(const Geom g, const Metric m, double* __restrict__ fact, double* __restrict__ dmod) {
    constexpr int VD = ND - 1;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = (ND == 3) ? blockIdx.y * blockDim.y + threadIdx.y : 0;
    if (i >= g.n[0] || (ND == 3 && j >= g.n[1])) return;
    const long long stv = (VD == 1) ? g.sy : g.sz;
    const int N = g.n[VD];
    const double vv = g.beta * 1.0 / (g.dx[VD] * g.dx[VD]);
    double dprev = 0.0, dlprev = 0.0;
    for (int r = 0; r < N; ++r) {
        const int jj = (ND == 3) ? j : r, kk = (ND == 3) ? r : 0;
        const long long c = g.idx(i, jj, kk);
        const unsigned sm = side_mask(g, i, jj, kk);
        double d = g.alpha / m.jinv[c] + line_lapdiag<ND>(g, m, i, jj, kk, c, sm);
        if (r > 0) {
            const double f = dlprev / dprev;          // dgtsv: fact = dl(i)/d(i)
            fact[c - stv] = f;
            d = d - f * dlprev;                        // d(i+1) = d(i+1) - fact*du(i),  du == dl
        }
        dmod[c] = d;
        dprev = d;
        dlprev = m.jg[VD][VD][c + stv] * vv;           // DL(r) = gHi * vvScale
    }
}

void launch_line_factor(Ctx& cx, const Geom& g, const Metric& m, double* fact, double* dmod) {
    if (g.nd == 3) {
        dim3 block(64, 2), grid((g.n[0] + 63) / 64, (g.n[1] + 1) / 2);
        k_line_factor<3><<<grid, block, 0, cx.stream>>>(g, m, fact, dmod);
    } else {
        k_line_factor<2><<<(g.n[0] + 63) / 64, 64, 0, cx.stream>>>(g, m, fact, dmod);
    }
    cx.launches++;
}

// RIMFREE: the block's columns touch no x/y side of the domain, so the boundary tests of the horizontal terms
// vanish at compile time and the k loops become straight-line code the compiler can unroll and batch loads for.
template <int ND, bool FULL, bool RIMFREE>
__device__ __forceinline__ void line_column(const Geom& g, const Metric& m, double* __restrict__ phiOut,
                                            const double* __restrict__ phiIn, const double* __restrict__ rhs,
                                            const double* __restrict__ fact, const double* __restrict__ dmod,
                                            int pass) {
    constexpr int VD = ND - 1;
    // thread t of a row owns the column i = 2t + (parity of this pass in row j): every lane of a warp is active
    // (a thread-per-cell mapping idles the lanes of the other colour: 16 of 32, measured)
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = (ND == 3) ? blockIdx.y * blockDim.y + threadIdx.y : 0;
    if (ND == 3 && j >= g.n[1]) return;
    const long long stv = (VD == 1) ? g.sy : g.sz;
    const int N = g.n[VD];
    const int par = (g.g0[0] + ((ND == 3) ? (g.g0[1] + j) : 0) + pass) & 1;
    const int i = 2 * t + par;
    if (phiOut != phiIn) {   // other colour: carried over unchanged (full metric, out of place)
        const int io = 2 * t + (par ^ 1);
        if (io < g.n[0])
            for (int r = 0; r < N; ++r) {
                const long long c = (ND == 3) ? g.idx(io, j, r) : g.idx(io, r, 0);
                phiOut[c] = phiIn[c];
            }
    }
    if (i >= g.n[0]) return;
    const double vv = g.beta * 1.0 / (g.dx[VD] * g.dx[VD]);
    const long long st[2] = {1, g.sy};
    double bprev = 0.0;
    LINE_PRAGMA(LINE_UNROLL)
    for (int r = 0; r < N; ++r) {
        const int jj = (ND == 3) ? j : r, kk = (ND == 3) ? r : 0;
        const long long c = g.idx(i, jj, kk);
        const unsigned sm = RIMFREE ? 0u : side_mask(g, i, jj, kk);
        const double pc = RIMFREE ? 0.0 : phiIn[c];   // only Dirichlet ghosts of x/y sides read it
        double JD[2] = {0.0, 0.0};
#pragma unroll
        for (int h = 0; h < VD; ++h) {
            const bool bLo = (sm >> (2 * h)) & 1u, bHi = (sm >> (2 * h + 1)) & 1u;
            if (!(bLo && g.bc[h][0] == BC_NEUM)) {
                double nb = phiIn[c - st[h]];
                if (bLo && g.bc[h][0] == BC_DIRI) nb = diri_ghost_val(pc, phiIn[c + st[h]], 0.0, g.diriOrder);
                JD[h] = JD[h] + m.jg[h][h][c] * nb;
            }
            if (!(bHi && g.bc[h][1] == BC_NEUM)) {
                double nb = phiIn[c + st[h]];
                if (bHi && g.bc[h][1] == BC_DIRI) nb = diri_ghost_val(pc, phiIn[c - st[h]], 0.0, g.diriOrder);
                JD[h] = JD[h] + m.jg[h][h][c + st[h]] * nb;
            }
        }
        double lphi = JD[0] * (g.beta * 1.0 / (g.dx[0] * g.dx[0]));
        if (ND == 3) lphi = lphi + JD[1] * (g.beta * 1.0 / (g.dx[1] * g.dx[1]));
        if (FULL) {
            const ExView<ND> v(g, phiIn, i, jj, kk, EXMODE_SMOOTHER);
            if (ND == 2) {
                const double ab01 = g.beta * 0.25 / (g.dx[0] * g.dx[1]);
                lphi = lphi + (cross_ab<ND>(m, v, i, jj, kk, c, 1, 0, 1) + cross_ab<ND>(m, v, i, jj, kk, c, g.sy, 1, 0)) * ab01;
            } else {
                const double ab12 = g.beta * 0.25 / (g.dx[1] * g.dx[2]);
                const double ab20 = g.beta * 0.25 / (g.dx[2] * g.dx[0]);
                const double ab01 = g.beta * 0.25 / (g.dx[0] * g.dx[1]);
                lphi = lphi + (cross_ab<ND>(m, v, i, jj, kk, c, g.sy, 1, 2) + cross_ab<ND>(m, v, i, jj, kk, c, g.sz, 2, 1)) * ab12;
                lphi = lphi + (cross_ab<ND>(m, v, i, jj, kk, c, g.sz, 2, 0) + cross_ab<ND>(m, v, i, jj, kk, c, 1, 0, 2)) * ab20;
                lphi = lphi + (cross_ab<ND>(m, v, i, jj, kk, c, 1, 0, 1) + cross_ab<ND>(m, v, i, jj, kk, c, g.sy, 1, 0)) * ab01;
            }
        }
        double b = -lphi + rhs[c] / m.jinv[c];
        if (r > 0) b = b - fact[c - stv] * bprev;      // b(i+1) = b(i+1) - fact*b(i)
        phiOut[c] = b;
        bprev = b;
    }
    // back substitution
    long long c = (ND == 3) ? g.idx(i, j, N - 1) : g.idx(i, N - 1, 0);
    double x = phiOut[c] / dmod[c];
    phiOut[c] = x;
    LINE_PRAGMA(LINE_UNROLL)
    for (int r = N - 2; r >= 0; --r) {
        c -= stv;
        const double du = m.jg[VD][VD][c + stv] * vv;
        x = (phiOut[c] - du * x) / dmod[c];
        phiOut[c] = x;
    }
}

template <int ND, bool FULL>
__global__ void __launch_bounds__(128, LINE_MINB) k_line_gsrb(const Geom g, const Metric m, double* __restrict__ phiOut,
                                                   const double* __restrict__ phiIn, const double* __restrict__ rhs,
                                                   const double* __restrict__ fact, const double* __restrict__ dmod,
                                                   int pass) {
    // block-uniform: do this block's cells (both colours of its x range, its rows) stay off the x/y domain sides?
    const int iLo = 2 * (int)(blockIdx.x * blockDim.x), iHi = iLo + 2 * (int)blockDim.x - 1;
    bool rimfree = !FULL && (iLo > 0 || g.bc[0][0] == BC_NONE) && (iHi < g.n[0] - 1 || g.bc[0][1] == BC_NONE);
    if (ND == 3) {
        const int jLo = blockIdx.y * blockDim.y, jHi = jLo + blockDim.y - 1;
        rimfree = rimfree && (jLo > 0 || g.bc[1][0] == BC_NONE) && (jHi < g.n[1] - 1 || g.bc[1][1] == BC_NONE);
    }
    if (rimfree) line_column<ND, FULL, true>(g, m, phiOut, phiIn, rhs, fact, dmod, pass);
    else line_column<ND, FULL, false>(g, m, phiOut, phiIn, rhs, fact, dmod, pass);
}

void launch_line_gsrb(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn,
                      const double* rhs, const double* fact, const double* dmod, int pass) {
    const int half = (g.n[0] + 1) / 2;
    if (g.nd == 3) {
        dim3 block(32, 4), grid((half + 31) / 32, (g.n[1] + 3) / 4);
        if (m.diagonal) k_line_gsrb<3, false><<<grid, block, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, fact, dmod, pass);
        else k_line_gsrb<3, true><<<grid, block, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, fact, dmod, pass);
    } else {
        const int nb = (half + 63) / 64;
        if (m.diagonal) k_line_gsrb<2, false><<<nb, 64, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, fact, dmod, pass);
        else k_line_gsrb<2, true><<<nb, 64, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, fact, dmod, pass);
    }
    cx.launches++;
}

}  // namespace somar

// line_compact.cu -- LineGSRB (LineGSRB::relax, GSRB.cpp:148-330; LineGSRBIter3D + dgtsv, GSRBF.ChF:1730-2042) in a
// COLOUR-SPLIT layout.
//
// k_line_gsrb (line.cu) moves 88 B per cell and colour pass where 44 are needed: columns are coloured by i + j, x is the
// fastest index, so a pass touches every 32-byte sector of every array and uses half of it.  A fused red+black kernel
// cannot fix that (the black columns need the red columns' COMPLETE solution: four z-sweeps through a working set far
// beyond L2).  What can: everything a pass touches only at its own cells is stored per colour, x-compacted:
//     element (t, j, k) of a pass-p array  <->  cell (2t + par(j, p), j, k),   par = (g0x + g0y + j + p) & 1
//   * once per (alpha, beta, metric):  XL, XH = J g^xx on the low / high x-face, YL, YH = J g^yy likewise, FACT, DMOD of
//     the pre-factored column matrix (k_line_factor), DU = J g^zz(k + 1/2) beta / dz^2 -- 7 arrays per pass;
//   * once per relax call:  RJ = rhs / Jinv, and phi itself (PHI[p]) -- phi stays compact for all the iterations of the
//     call and is merged back into the brick at the end;
//   * scratch:  BP, the forward-substituted right-hand side.
// A pass reads PHI[1 - p] (the four neighbours of a column are of the other colour: x-neighbours at t-1 / t or t / t+1,
// y-neighbours at the same t of rows j -+ 1), XL, XH, YL, YH, RJ, FACT of pass p, writes BP; then reads BP, DMOD, DU and
// writes PHI[p]: 96 B per active cell, every sector fully used, against ~190.  The statements and their order are
// line_column's (RIMFREE or not), with the same operands, so the results are the same bits.
// Built for: 3-D, diagonal metric, one rank, x and y not periodic (anything else keeps k_line_gsrb).
#include "level.hpp"
#include "stencil_ortho.cuh"

#include <cstdlib>

namespace somar {

namespace {

struct LineC {
    int nxh;              // columns per row of one colour = ceil(nx / 2)
    long long pitch;      // compact row pitch (multiple of 4 doubles)
    long long plane;      // pitch * ny
    long long total;      // plane * nz
};

struct LineCArrays {
    double *XL, *XH, *YL, *YH, *FACT, *DMOD, *DU;   // coefficients of pass p
    double* RJ;
    double* PHI;          // this pass's phi (in/out)
    const double* PHIO;   // the other colour's phi
    double* BP;
};

__device__ __forceinline__ int par_of(const Geom& g, int j, int pass) { return (g.g0[0] + g.g0[1] + j + pass) & 1; }

// MODE 0: coefficient set-up of pass `pass` (after k_line_factor); 1: pack phi and rhs/Jinv; 2: unpack phi
template <int MODE>
__global__ void __launch_bounds__(256) k_line_pack(const Geom g, const Metric m, const LineC lc, int pass, LineCArrays a,
                                                    double* __restrict__ phi, const double* __restrict__ rhs,
                                                    const double* __restrict__ fact, const double* __restrict__ dmod) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (t >= lc.nxh || j >= g.n[1]) return;
    const int i = 2 * t + par_of(g, j, pass);
    if (i >= g.n[0]) return;
    const long long c = g.idx(i, j, k);
    const long long q = t + lc.pitch * j + lc.plane * k;
    if (MODE == 0) {
        const double vv = g.beta * 1.0 / (g.dx[2] * g.dx[2]);
        a.XL[q] = m.jg[0][0][c];
        a.XH[q] = m.jg[0][0][c + 1];
        a.YL[q] = m.jg[1][1][c];
        a.YH[q] = m.jg[1][1][c + g.sy];
        a.FACT[q] = fact[c];
        a.DMOD[q] = dmod[c];
        a.DU[q] = m.jg[2][2][c + g.sz] * vv;
    } else if (MODE == 1) {
        a.PHI[q] = phi[c];
        a.RJ[q] = rhs[c] / m.jinv[c];
    } else {
        phi[c] = a.PHI[q];
    }
}

// one colour pass: thread = one column (t, j), marching k up (forward substitution) and down (back substitution)
template <bool RIMFREE>
__device__ __forceinline__ void line_column_c(const Geom& g, const LineC& lc, const LineCArrays& a, int pass) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    if (t >= lc.nxh || j >= g.n[1]) return;
    const int par = par_of(g, j, pass);
    const int i = 2 * t + par;
    if (i >= g.n[0]) return;
    const int N = g.n[2];
    // the other colour's columns next to mine: x-neighbours i-1, i+1; y-neighbours (i, j -+ 1) sit at my own t
    const long long q0 = t + lc.pitch * j;
    const long long qxl = q0 - (par ? 0 : 1), qxh = q0 + (par ? 1 : 0);
    const long long qyl = q0 - lc.pitch, qyh = q0 + lc.pitch;
    const double ax = g.beta * 1.0 / (g.dx[0] * g.dx[0]);
    const double ay = g.beta * 1.0 / (g.dx[1] * g.dx[1]);
    // x / y sides of the domain this column touches (block-uniformly none when RIMFREE)
    const bool bxl = !RIMFREE && i == 0 && g.bc[0][0] != BC_NONE, bxh = !RIMFREE && i == g.n[0] - 1 && g.bc[0][1] != BC_NONE;
    const bool byl = !RIMFREE && j == 0 && g.bc[1][0] != BC_NONE, byh = !RIMFREE && j == g.n[1] - 1 && g.bc[1][1] != BC_NONE;
    double bprev = 0.0;
    for (int r = 0; r < N; ++r) {
        const long long o = lc.plane * r;
        const double pc = RIMFREE ? 0.0 : a.PHI[q0 + o];   // only Dirichlet ghosts of x/y sides read it
        double JD0 = 0.0, JD1 = 0.0;
        if (!(bxl && g.bc[0][0] == BC_NEUM)) {
            double nb = bxl ? 0.0 : a.PHIO[qxl + o];
            if (bxl && g.bc[0][0] == BC_DIRI) nb = diri_ghost_val(pc, a.PHIO[qxh + o], 0.0, g.diriOrder);
            JD0 = JD0 + a.XL[q0 + o] * nb;
        }
        if (!(bxh && g.bc[0][1] == BC_NEUM)) {
            double nb = bxh ? 0.0 : a.PHIO[qxh + o];
            if (bxh && g.bc[0][1] == BC_DIRI) nb = diri_ghost_val(pc, a.PHIO[qxl + o], 0.0, g.diriOrder);
            JD0 = JD0 + a.XH[q0 + o] * nb;
        }
        if (!(byl && g.bc[1][0] == BC_NEUM)) {
            double nb = byl ? 0.0 : a.PHIO[qyl + o];
            if (byl && g.bc[1][0] == BC_DIRI) nb = diri_ghost_val(pc, a.PHIO[qyh + o], 0.0, g.diriOrder);
            JD1 = JD1 + a.YL[q0 + o] * nb;
        }
        if (!(byh && g.bc[1][1] == BC_NEUM)) {
            double nb = byh ? 0.0 : a.PHIO[qyh + o];
            if (byh && g.bc[1][1] == BC_DIRI) nb = diri_ghost_val(pc, a.PHIO[qyl + o], 0.0, g.diriOrder);
            JD1 = JD1 + a.YH[q0 + o] * nb;
        }
        double lphi = JD0 * ax;
        lphi = lphi + JD1 * ay;
        double b = -lphi + a.RJ[q0 + o];
        if (r > 0) b = b - a.FACT[q0 + o - lc.plane] * bprev;   // b(i+1) = b(i+1) - fact*b(i)
        a.BP[q0 + o] = b;
        bprev = b;
    }
    long long q = q0 + lc.plane * (N - 1);
    double x = bprev / a.DMOD[q];
    a.PHI[q] = x;
    for (int r = N - 2; r >= 0; --r) {
        q -= lc.plane;
        x = (a.BP[q] - a.DU[q] * x) / a.DMOD[q];
        a.PHI[q] = x;
    }
}

__global__ void __launch_bounds__(128, 8) k_line_pass_c(const Geom g, const LineC lc, const LineCArrays a, int pass) {
    const int iLo = 2 * (int)(blockIdx.x * blockDim.x), iHi = iLo + 2 * (int)blockDim.x - 1;
    const int jLo = blockIdx.y * blockDim.y, jHi = jLo + blockDim.y - 1;
    const bool rimfree = (iLo > 0 || g.bc[0][0] == BC_NONE) && (iHi < g.n[0] - 1 || g.bc[0][1] == BC_NONE) &&
                         (jLo > 0 || g.bc[1][0] == BC_NONE) && (jHi < g.n[1] - 1 || g.bc[1][1] == BC_NONE);
    if (rimfree) line_column_c<true>(g, lc, a, pass);
    else line_column_c<false>(g, lc, a, pass);
}

}  // namespace

// ---- host side ------------------------------------------------------------------------------------------------
struct LineCompact {
    LineC lc;
    double* coef[2][7] = {{nullptr}};   // XL XH YL YH FACT DMOD DU per pass
    double* rj[2] = {nullptr, nullptr};
    double* phi[2] = {nullptr, nullptr};
    double* bp = nullptr;
    bool coef_valid = false;
};

bool line_compact_usable(const SomarLevel* L, const Depth& dp) {
    const char* e = getenv("SOMAR_B200_LINE_COMPACT");   // read per call: the tests switch it within one process
    const bool on = !(e && e[0] == '0');
    const Geom& g = dp.g;
    // (small depths -- the bottom solver's preconditioner -- keep the brick kernel: two pack and two merge launches per
    //  call cost more than the sectors they save)
    return on && dp.m.diagonal && g.nd == 3 && L->desc.nranks == 1 && !g.periodic[0] && !g.periodic[1] && g.n[0] >= 2 &&
           (g.cells() >= (1ll << 20) || (e && e[0] == '2'));   // SOMAR_B200_LINE_COMPACT=2: at every size (tests)
}

void line_compact_free(Depth& dp) {
    LineCompact* C = static_cast<LineCompact*>(dp.line_compact);
    if (!C) return;
    for (int p = 0; p < 2; ++p) {
        for (double* x : C->coef[p]) cudaFree(x);
        cudaFree(C->rj[p]);
        cudaFree(C->phi[p]);
    }
    cudaFree(C->bp);
    delete C;
    dp.line_compact = nullptr;
}

static LineCArrays arrays_of(LineCompact* C, int pass) {
    LineCArrays a;
    a.XL = C->coef[pass][0]; a.XH = C->coef[pass][1]; a.YL = C->coef[pass][2]; a.YH = C->coef[pass][3];
    a.FACT = C->coef[pass][4]; a.DMOD = C->coef[pass][5]; a.DU = C->coef[pass][6];
    a.RJ = C->rj[pass];
    a.PHI = C->phi[pass];
    a.PHIO = C->phi[1 - pass];
    a.BP = C->bp;
    return a;
}

// `iters` LineGSRB iterations on phi (brick) with right-hand side rhs (brick); fact / dmod: k_line_factor's arrays.
// coefStale: alpha, beta or the metric changed since the compact coefficients were built.
int line_compact_relax(SomarLevel* L, Depth& dp, double* phi, const double* rhs, int iters, bool coefStale, int profPhase) {
    const Geom& g = dp.g;
    LineCompact* C = static_cast<LineCompact*>(dp.line_compact);
    if (!C) {
        C = new LineCompact();
        dp.line_compact = C;
        C->lc.nxh = (g.n[0] + 1) / 2;
        C->lc.pitch = ((C->lc.nxh + 3) / 4) * 4;
        C->lc.plane = C->lc.pitch * g.n[1];
        C->lc.total = C->lc.plane * g.n[2];
        const size_t bytes = sizeof(double) * (size_t)C->lc.total;
        // every array starts at zero: the pad columns of a row and, with an odd row length, the last column of one
        // colour are never written, and a value selected away must not be a NaN that a 0 * x would keep
        auto zalloc = [&](double*& x) {
            if (cudaMalloc(&x, bytes) != cudaSuccess) return false;
            return cudaMemsetAsync(x, 0, bytes, L->cx.stream) == cudaSuccess;
        };
        bool ok = true;
        for (int p = 0; p < 2; ++p) {
            for (double*& x : C->coef[p]) ok = ok && zalloc(x);
            ok = ok && zalloc(C->rj[p]) && zalloc(C->phi[p]);
        }
        ok = ok && zalloc(C->bp);
        if (!ok) {
            set_error(std::string("line_compact: ") + cudaGetErrorString(cudaGetLastError()));
            return SOMAR_ERR_CUDA;
        }
    }
    const LineC lc = C->lc;
    const dim3 pb(64, 4), pg((lc.nxh + 63) / 64, (g.n[1] + 3) / 4, g.n[2]);
    if (!C->coef_valid || coefStale) {
        for (int p = 0; p < 2; ++p) {
            k_line_pack<0><<<pg, pb, 0, L->cx.stream>>>(g, dp.m, lc, p, arrays_of(C, p), nullptr, nullptr, dp.line_fact, dp.line_dmod);
            L->cx.launches++;
        }
        C->coef_valid = true;
    }
    for (int p = 0; p < 2; ++p) {
        k_line_pack<1><<<pg, pb, 0, L->cx.stream>>>(g, dp.m, lc, p, arrays_of(C, p), phi, rhs, nullptr, nullptr);
        L->cx.launches++;
    }
    const dim3 kb(32, 4), kg((lc.nxh + 31) / 32, (g.n[1] + 3) / 4);
    for (int it = 0; it < iters; ++it)
        for (int pass = 0; pass < 2; ++pass) {
            ProfScope ps(L, profPhase);
            k_line_pass_c<<<kg, kb, 0, L->cx.stream>>>(g, lc, arrays_of(C, pass), pass);
            L->cx.launches++;
        }
    for (int p = 0; p < 2; ++p) {
        k_line_pack<2><<<pg, pb, 0, L->cx.stream>>>(g, dp.m, lc, p, arrays_of(C, p), phi, nullptr, nullptr, nullptr);
        L->cx.launches++;
    }
    return SOMAR_OK;
}

}  // namespace somar

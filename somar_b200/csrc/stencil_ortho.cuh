// stencil_ortho.cuh -- per-cell device functions for the DIAGONAL-metric operator (J g^{ij} = 0, i != j).
// The arithmetic order follows the reference's Fortran statement by statement so that, compiled with
// -fmad=false, results are bit-identical to a non-contracting CPU build of the same statements:
//   flux      MAPPEDGETFLUXORTHO        src/calculus/AMRElliptic/MappedAMRPoissonOpOrthoF.ChF:33-82
//   BC flux   EllipticConstNeumBCFluxClass  src/calculus/BCInterface/EllipticBCUtils.cpp:491-546
//   BC ghost  ELLIPTICCONSTDIRIBCGHOST  src/calculus/BCInterface/EllipticBCUtilsF.ChF:29-120
//   div       MAPPEDFLUXDIVERGENCE{2,3}D src/calculus/DivCurlGrad/DivCurlGradF.ChF:1034-1214
//   axby      AXBYIP                    src/calculus/AMRElliptic/MappedAMRPoissonOpF.ChF:62-84
//   GSRB      GSRBITER{2,3}DORTHO       src/calculus/AMRElliptic/RelaxationMethods/GSRBF.ChF:442-705
//   GSRB bdry GSRBBOUNDARYITER{2,3}DORTHO  GSRBF.ChF:1233-1505
#pragma once
#include "geom.cuh"

namespace somar {

// every function of this header also compiles for the host (tests/cuda/*_host.cu run the statements on the CPU)
#ifndef SOMAR_HD
#define SOMAR_HD __host__ __device__
#endif
#ifndef SOMAR_UNROLL
#ifdef __CUDA_ARCH__
#define SOMAR_UNROLL _Pragma("unroll")
#else
#define SOMAR_UNROLL   // (the host pass does not know the pragma)
#endif
#endif

SOMAR_HD __forceinline__ double diri_ghost_val(double nearV, double farV, double bcval, int order) {
    if (order == 0) return bcval;
    if (order == 1) return (bcval == 0.0) ? -nearV : 2.0 * bcval - nearV;
    return (bcval == 0.0) ? (1.0 / 3.0) * farV - 2.0 * nearV : (8.0 * bcval - 6.0 * nearV + farV) / 3.0;
}

// Which physical-boundary sides does local cell (i,j,k) touch?  bit (2*d+s).
SOMAR_HD __forceinline__ unsigned side_mask(const Geom& g, int i, int j, int k) {
    unsigned m = 0;
    if (i == 0 && g.bc[0][0] != BC_NONE) m |= 1u;
    if (i == g.n[0] - 1 && g.bc[0][1] != BC_NONE) m |= 2u;
    if (j == 0 && g.bc[1][0] != BC_NONE) m |= 4u;
    if (j == g.n[1] - 1 && g.bc[1][1] != BC_NONE) m |= 8u;
    if (g.nd > 2) {
        if (k == 0 && g.bc[2][0] != BC_NONE) m |= 16u;
        if (k == g.n[2] - 1 && g.bc[2][1] != BC_NONE) m |= 32u;
    }
    return m;
}

// beta * flux through the low (S=0) / high (S=1) face of the cell at address c in direction with stride st.
// pc = phi(c); the neighbour / ghost is resolved here.
template <int S>
SOMAR_HD __forceinline__ double face_flux_ortho(const Geom& g, const double* __restrict__ phi,
                                                  const double* __restrict__ jgd, long long c, long long st, int d,
                                                  bool atBdry, bool homog, double pc) {
    const double beta_dx = g.rdx[d];
    const long long f = S ? c + st : c;            // face address
    double nb;                                      // value on the other side of the face
    if (atBdry) {
        const int bct = g.bc[d][S];
        if (bct == BC_NEUM) return (homog ? 0.0 : g.bcval[d][S]) * g.beta;
        // Dirichlet: ghost from the interior values
        const double farV = S ? phi[c - st] : phi[c + st];
        if (bct == BC_EXTRAP2) {   // ELLIPTICEXTRAPBCGHOST order 2, EllipticBCUtilsF.ChF:165-173
            const double far2 = S ? phi[c - 2 * st] : phi[c + 2 * st];
            nb = 3.0 * (pc - farV) + far2;
        } else {
            nb = diri_ghost_val(pc, farV, homog ? 0.0 : g.bcval[d][S], g.diriOrder);
        }
    } else {
        nb = S ? phi[c + st] : phi[c - st];
    }
    const double diff = S ? (nb - pc) : (pc - nb);
    return (jgd[f] * beta_dx * diff) * g.beta;
}

// (alpha I + beta L) phi at one cell, diagonal metric.
template <int ND>
SOMAR_HD __forceinline__ double apply_cell_ortho(const Geom& g, const Metric& m, const double* __restrict__ phi,
                                                   int i, int j, int k, long long c, bool homog) {
    const unsigned sm = side_mask(g, i, j, k);
    const double pc = phi[c];
    const double f0l = face_flux_ortho<0>(g, phi, m.jg[0][0], c, 1, 0, sm & 1u, homog, pc);
    const double f0h = face_flux_ortho<1>(g, phi, m.jg[0][0], c, 1, 0, sm & 2u, homog, pc);
    const double f1l = face_flux_ortho<0>(g, phi, m.jg[1][1], c, g.sy, 1, sm & 4u, homog, pc);
    const double f1h = face_flux_ortho<1>(g, phi, m.jg[1][1], c, g.sy, 1, sm & 8u, homog, pc);
    double s = (f0h - f0l) * g.rdx[0];
    s = s + (f1h - f1l) * g.rdx[1];
    if (ND > 2) {
        const double f2l = face_flux_ortho<0>(g, phi, m.jg[2][2], c, g.sz, 2, sm & 16u, homog, pc);
        const double f2h = face_flux_ortho<1>(g, phi, m.jg[2][2], c, g.sz, 2, sm & 32u, homog, pc);
        s = s + (f2h - f2l) * g.rdx[2];
    }
    double r = m.jinv[c] * s;
    if (g.alpha != 0.0) r = g.alpha * pc + 1.0 * r;
    return r;
}

// The same two functions from VALUES (for kernels that march through z with phi and J g^zz in registers): nbv = phi
// across the face, farV = phi across the opposite face (Dirichlet ghosts), jgf = J g^{dd} on the face.  Statement
// for statement the arithmetic of face_flux_ortho / apply_cell_ortho.
template <int S>
SOMAR_HD __forceinline__ double face_flux_vals(const Geom& g, int d, bool atBdry, bool homog, double pc, double nbv,
                                                 double farV, double jgf) {
    const double beta_dx = g.rdx[d];
    double nb = nbv;
    if (atBdry) {
        const int bct = g.bc[d][S];
        if (bct == BC_NEUM) return (homog ? 0.0 : g.bcval[d][S]) * g.beta;
        nb = diri_ghost_val(pc, farV, homog ? 0.0 : g.bcval[d][S], g.diriOrder);
    }
    const double diff = S ? (nb - pc) : (pc - nb);
    return (jgf * beta_dx * diff) * g.beta;
}

SOMAR_HD __forceinline__ double apply_vals_ortho3(const Geom& g, unsigned sm, bool homog, double pc,
                                                    const double (&nb)[3][2], const double (&glo)[3],
                                                    const double (&ghi)[3], double ji) {
    const double f0l = face_flux_vals<0>(g, 0, sm & 1u, homog, pc, nb[0][0], nb[0][1], glo[0]);
    const double f0h = face_flux_vals<1>(g, 0, sm & 2u, homog, pc, nb[0][1], nb[0][0], ghi[0]);
    const double f1l = face_flux_vals<0>(g, 1, sm & 4u, homog, pc, nb[1][0], nb[1][1], glo[1]);
    const double f1h = face_flux_vals<1>(g, 1, sm & 8u, homog, pc, nb[1][1], nb[1][0], ghi[1]);
    double s = (f0h - f0l) * g.rdx[0];
    s = s + (f1h - f1l) * g.rdx[1];
    const double f2l = face_flux_vals<0>(g, 2, sm & 16u, homog, pc, nb[2][0], nb[2][1], glo[2]);
    const double f2h = face_flux_vals<1>(g, 2, sm & 32u, homog, pc, nb[2][1], nb[2][0], ghi[2]);
    s = s + (f2h - f2l) * g.rdx[2];
    double r = ji * s;
    if (g.alpha != 0.0) r = g.alpha * pc + 1.0 * r;
    return r;
}

// Is the cell on the rim of the DOMAIN (any active direction, periodic or not)?  Such cells are relaxed by the
// boundary kernels of the reference (BaseGSRB::boundaryGSRB, GSRB.cpp:489-653; domInterior = grow(domain,-1)).
template <int ND>
SOMAR_HD __forceinline__ bool on_domain_rim(const Geom& g, int i, int j, int k) {
    const int gi = i + g.g0[0], gk = k + g.g0[2];
    int gj = j + g.g0[1];
    // a ghost row beyond a periodic y side is the image of a row at the other end of the domain (the fused smoother
    // relaxes the red cells of the first ghost row exactly as their owner does)
    if (g.periodic[1]) gj = (gj < 0) ? gj + g.gn[1] : (gj >= g.gn[1] ? gj - g.gn[1] : gj);
    bool r = (gi == 0) | (gi == g.gn[0] - 1) | (gj == 0) | (gj == g.gn[1] - 1);
    if (ND > 2) r = r | (gk == 0) | (gk == g.gn[2] - 1);
    return r;
}

// Coefficients one red-black update reads, as VALUES (so that a kernel can prefetch them into registers).
struct GsrbCoef {
    double glo[3], ghi[3];   // J g^{dd} on the low / high face of the cell
    double ji, lap, rhs;     // 1/J, lapDiag, right-hand side
};

template <int ND>
SOMAR_HD __forceinline__ GsrbCoef gsrb_load_coef(const Geom& g, const Metric& m, const double* __restrict__ rhs,
                                                   long long c) {
    GsrbCoef q;
    q.glo[0] = m.jg[0][0][c];
    q.ghi[0] = m.jg[0][0][c + 1];
    q.glo[1] = m.jg[1][1][c];
    q.ghi[1] = m.jg[1][1][c + g.sy];
    q.glo[2] = (ND > 2) ? m.jg[2][2][c] : 0.0;
    q.ghi[2] = (ND > 2) ? m.jg[2][2][c + g.sz] : 0.0;
    q.ji = m.jinv[c];
    q.lap = m.lapdiag[c];
    q.rhs = rhs[c];
    return q;
}

// One red-black update of one cell, diagonal metric, from VALUES: nb[d][s] = phi of the neighbour across side s of
// direction d (any number where that side is a physical boundary), pc = the cell's own pre-pass value (only used
// for Dirichlet ghosts, which the reference fills from the pre-pass state).  Shared by the colour-pass kernel
// (values from global memory) and the fused red+black kernel (phi from the shared-memory planes, coefficients
// prefetched), so that both execute the same statements in the same order.
// LAPFLY: do not read q.lap; form lapDiag from the six face coefficients and 1/J the update holds anyway, with the
// statements of FILLMAPPEDLAPDIAG (k_lapdiag) -- the same bits, one array less to stream (8 of 64 B/cell).
template <int ND, bool LAPFLY = false>
SOMAR_HD __forceinline__ double gsrb_update_vals(const Geom& g, const GsrbCoef& q, int i, int j, int k, double pc,
                                                   const double (&nb)[3][2], bool knownInterior = false,
                                                   bool forceRim = false) {
    const double sc0 = g.rdx2[0];
    const double sc1 = g.rdx2[1];
    const double sc2 = (ND > 2) ? g.rdx2[2] : 0.0;
    const double ji = q.ji;
    // knownInterior: the caller has established (block-uniformly) that the cell is not on the domain rim
    // forceRim: use the boundary kernel's statements even off the domain rim (LooseGSRB relaxes every BOX rim with them)
    if (!forceRim && (knownInterior || !on_domain_rim<ND>(g, i, j, k))) {
        const double JDxx = sc0 * (q.ghi[0] * nb[0][1] + q.glo[0] * nb[0][0]);
        const double JDyy = sc1 * (q.ghi[1] * nb[1][1] + q.glo[1] * nb[1][0]);
        double lphi;
        if (ND > 2) {
            const double JDzz = sc2 * (q.ghi[2] * nb[2][1] + q.glo[2] * nb[2][0]);
            lphi = g.beta * ji * (JDxx + JDyy + JDzz);
        } else {
            lphi = g.beta * (JDxx + JDyy) * ji;   // GSRBITER2DORTHO's own association (GSRBF.ChF:442-543), not the 3-D kernel's
        }
        double lap = q.lap;
        if (LAPFLY) {
            if (ND > 2) lap = -ji * ((q.ghi[0] + q.glo[0]) * sc0 + (q.ghi[1] + q.glo[1]) * sc1 + (q.ghi[2] + q.glo[2]) * sc2);
            else lap = -ji * ((q.ghi[0] + q.glo[0]) * sc0 + (q.ghi[1] + q.glo[1]) * sc1);
        }
        return (q.rhs - lphi) / (g.alpha + g.beta * lap);
    }
    // rim cell: Neumann sides are dropped from the stencil and from the diagonal
    const unsigned sm = side_mask(g, i, j, k);
    double JDlo[3] = {0.0, 0.0, 0.0}, JDhi[3] = {0.0, 0.0, 0.0};
    double lapDiag = 0.0;
    const double sc[3] = {sc0, sc1, sc2};
    // the diagonal is accumulated in the Fortran's order: 3-D lo x, lo y, lo z, hi x, hi y, hi z (GSRBBOUNDARYITER3DORTHO);
    // 2-D lo x, hi x, lo y, hi y (GSRBBOUNDARYITER2DORTHO, GSRBF.ChF:1233-1361)
    auto lo_term = [&](int d) {
        const bool atB = (sm >> (2 * d)) & 1u;
        if (!(atB && g.bc[d][0] == BC_NEUM)) {
            double v = nb[d][0];
            if (atB && g.bc[d][0] == BC_DIRI) v = diri_ghost_val(pc, nb[d][1], 0.0, g.diriOrder);
            JDlo[d] = q.glo[d] * v;
            lapDiag = lapDiag - sc[d] * q.glo[d];
        }
    };
    auto hi_term = [&](int d) {
        const bool atB = (sm >> (2 * d + 1)) & 1u;
        if (!(atB && g.bc[d][1] == BC_NEUM)) {
            double v = nb[d][1];
            if (atB && g.bc[d][1] == BC_DIRI) v = diri_ghost_val(pc, nb[d][0], 0.0, g.diriOrder);
            JDhi[d] = q.ghi[d] * v;
            lapDiag = lapDiag - sc[d] * q.ghi[d];
        }
    };
    if (ND > 2) {
SOMAR_UNROLL
        for (int d = 0; d < ND; ++d) lo_term(d);
SOMAR_UNROLL
        for (int d = 0; d < ND; ++d) hi_term(d);
    } else {
SOMAR_UNROLL
        for (int d = 0; d < ND; ++d) {
            lo_term(d);
            hi_term(d);
        }
    }
    lapDiag = lapDiag * ji;
    double s = (JDlo[0] + JDhi[0]) * sc0 + (JDlo[1] + JDhi[1]) * sc1;
    if (ND > 2) s = s + (JDlo[2] + JDhi[2]) * sc2;
    const double lphi = g.beta * ji * s;
    return (q.rhs - lphi) / (g.alpha + g.beta * lapDiag);
}

// Same, reading everything from global memory (in-place colour pass).
template <int ND>
SOMAR_HD __forceinline__ double gsrb_cell_ortho(const Geom& g, const Metric& m, const double* __restrict__ phi,
                                                  const double* __restrict__ rhs, int i, int j, int k, long long c) {
    double nb[3][2];
    nb[0][0] = phi[c - 1];
    nb[0][1] = phi[c + 1];
    nb[1][0] = phi[c - g.sy];
    nb[1][1] = phi[c + g.sy];
    nb[2][0] = (ND > 2) ? phi[c - g.sz] : 0.0;
    nb[2][1] = (ND > 2) ? phi[c + g.sz] : 0.0;
    const GsrbCoef q = gsrb_load_coef<ND>(g, m, rhs, c);
    return gsrb_update_vals<ND>(g, q, i, j, k, phi[c], nb);
}

}  // namespace somar

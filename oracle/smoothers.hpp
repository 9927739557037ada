// oracle/smoothers.hpp -- TEST INFRASTRUCTURE ONLY (CPU oracle; never the product path).
//
// Restatement of the red-black Gauss-Seidel / Jacobi Fortran kernels
//   src/calculus/AMRElliptic/RelaxationMethods/GSRBF.ChF, JacobiF.ChF
// Red/black parity uses absolute indices: imin = lo + |mod(lo + j + k + redBlack, 2)|  (GSRBF.ChF:381-385).
#pragma once
#include "kernels.hpp"

namespace orc {

inline int rb_imin(int ilo, int j, int k, int redBlack, int nd) {
    const int indtot = ilo + j + (nd > 2 ? k : 0);
    return ilo + std::abs((indtot + redBlack) % 2);
}

// GSRBITER3DORTHO -- GSRBF.ChF:545-705, GSRBITER2DORTHO -- :442-543.
// Fast path (alpha==0, beta==1, ncomp==1) accumulates vJGrad term by term, then divides by lapDiag.
inline void gsrb_ortho(FAB& phi, const FAB& rhs, const FAB* Jg, const FAB& Jinv, const FAB& lapDiag,
                       const Box& region, const double* dx, double alpha, double beta, int redBlack, int nd) {
    const double xx = 1.0 / (dx[0] * dx[0]);
    const double yy = 1.0 / (dx[1] * dx[1]);
    const double zz = (nd > 2) ? 1.0 / (dx[2] * dx[2]) : 0.0;
    const bool fast = (alpha == 0.0 && beta == 1.0 && phi.nc == 1);
    for (int n = 0; n < phi.nc; ++n)
        for (int k = region.lo[2]; k <= region.hi[2]; ++k)
            for (int j = region.lo[1]; j <= region.hi[1]; ++j) {
                const int imin = rb_imin(region.lo[0], j, k, redBlack, nd);
                for (int i = imin; i <= region.hi[0]; i += 2) {
                    if (fast) {
                        double v = xx * (Jg[0](i + 1, j, k, 0) * phi(i + 1, j, k) + Jg[0](i, j, k, 0) * phi(i - 1, j, k));
                        v = v + yy * (Jg[1](i, j + 1, k, 1) * phi(i, j + 1, k) + Jg[1](i, j, k, 1) * phi(i, j - 1, k));
                        if (nd > 2)
                            v = v + zz * (Jg[2](i, j, k + 1, 2) * phi(i, j, k + 1) + Jg[2](i, j, k, 2) * phi(i, j, k - 1));
                        double p = rhs(i, j, k) - Jinv(i, j, k) * v;
                        phi(i, j, k) = p / lapDiag(i, j, k);
                    } else {
                        const double JDxx =
                            xx * (Jg[0](i + 1, j, k, 0) * phi(i + 1, j, k, n) + Jg[0](i, j, k, 0) * phi(i - 1, j, k, n));
                        const double JDyy =
                            yy * (Jg[1](i, j + 1, k, 1) * phi(i, j + 1, k, n) + Jg[1](i, j, k, 1) * phi(i, j - 1, k, n));
                        double lphi;
                        if (nd > 2) {
                            const double JDzz = zz * (Jg[2](i, j, k + 1, 2) * phi(i, j, k + 1, n) +
                                                      Jg[2](i, j, k, 2) * phi(i, j, k - 1, n));
                            lphi = beta * Jinv(i, j, k) * (JDxx + JDyy + JDzz);
                        } else {
                            // GSRBITER2DORTHO (GSRBF.ChF:442-543) associates differently from the 3-D kernel:
                            // lphi = beta * (JDxx + JDyy) * Jinv  (found by running the Fortran itself, oracle/chf2py.py)
                            lphi = beta * (JDxx + JDyy) * Jinv(i, j, k);
                        }
                        phi(i, j, k, n) = (rhs(i, j, k, n) - lphi) / (alpha + beta * lapDiag(i, j, k));
                    }
                }
            }
}

// GSRBITER3D -- GSRBF.ChF:283-440 and GSRBITER2D -- :155-282 (full metric; cross terms read extrap).
inline void gsrb_full(FAB& phi, const FAB& ex, const FAB& rhs, const FAB* Jg, const FAB& Jinv, const FAB& lapDiag,
                      const Box& region, const double* dx, double alpha, double beta, int redBlack, int nd) {
    const double xx = 1.0 / (dx[0] * dx[0]), yy = 1.0 / (dx[1] * dx[1]);
    const double xy = 0.25 / (dx[0] * dx[1]);
    const double zz = (nd > 2) ? 1.0 / (dx[2] * dx[2]) : 0.0;
    const double yz = (nd > 2) ? 0.25 / (dx[1] * dx[2]) : 0.0;
    const double zx = (nd > 2) ? 0.25 / (dx[2] * dx[0]) : 0.0;
    const FAB &G0 = Jg[0], &G1 = Jg[1];
    for (int n = 0; n < phi.nc; ++n)
        for (int k = region.lo[2]; k <= region.hi[2]; ++k)
            for (int j = region.lo[1]; j <= region.hi[1]; ++j) {
                const int imin = rb_imin(region.lo[0], j, k, redBlack, nd);
                for (int i = imin; i <= region.hi[0]; i += 2) {
                    double lphi;
                    if (nd == 2) {
                        const double JDxx = G0(i + 1, j, k, 0) * phi(i + 1, j, k, n) + G0(i, j, k, 0) * phi(i - 1, j, k, n);
                        const double JDxy =
                            G0(i + 1, j, k, 1) * (ex(i + 1, j + 1, k, n) - ex(i + 1, j - 1, k, n) + ex(i, j + 1, k, n) - ex(i, j - 1, k, n)) -
                            G0(i, j, k, 1) * (ex(i, j + 1, k, n) - ex(i, j - 1, k, n) + ex(i - 1, j + 1, k, n) - ex(i - 1, j - 1, k, n));
                        const double JDyx =
                            G1(i, j + 1, k, 0) * (ex(i + 1, j + 1, k, n) - ex(i - 1, j + 1, k, n) + ex(i + 1, j, k, n) - ex(i - 1, j, k, n)) -
                            G1(i, j, k, 0) * (ex(i + 1, j, k, n) - ex(i - 1, j, k, n) + ex(i + 1, j - 1, k, n) - ex(i - 1, j - 1, k, n));
                        const double JDyy = G1(i, j + 1, k, 1) * phi(i, j + 1, k, n) + G1(i, j, k, 1) * phi(i, j - 1, k, n);
                        lphi = beta * Jinv(i, j, k) * (JDxx * xx + JDyy * yy + (JDxy + JDyx) * xy);
                    } else {
                        const FAB& G2 = Jg[2];
                        const double pdx = ex(i + 1, j, k, n) - ex(i - 1, j, k, n);
                        const double pdy = ex(i, j + 1, k, n) - ex(i, j - 1, k, n);
                        const double pdz = ex(i, j, k + 1, n) - ex(i, j, k - 1, n);
                        const double JDxx = G0(i + 1, j, k, 0) * phi(i + 1, j, k, n) + G0(i, j, k, 0) * phi(i - 1, j, k, n);
                        const double JDxy = G0(i + 1, j, k, 1) * (ex(i + 1, j + 1, k, n) - ex(i + 1, j - 1, k, n) + pdy) -
                                            G0(i, j, k, 1) * (pdy + ex(i - 1, j + 1, k, n) - ex(i - 1, j - 1, k, n));
                        const double JDxz = G0(i + 1, j, k, 2) * (ex(i + 1, j, k + 1, n) - ex(i + 1, j, k - 1, n) + pdz) -
                                            G0(i, j, k, 2) * (pdz + ex(i - 1, j, k + 1, n) - ex(i - 1, j, k - 1, n));
                        const double JDyx = G1(i, j + 1, k, 0) * (ex(i + 1, j + 1, k, n) - ex(i - 1, j + 1, k, n) + pdx) -
                                            G1(i, j, k, 0) * (pdx + ex(i + 1, j - 1, k, n) - ex(i - 1, j - 1, k, n));
                        const double JDyy = G1(i, j + 1, k, 1) * phi(i, j + 1, k, n) + G1(i, j, k, 1) * phi(i, j - 1, k, n);
                        const double JDyz = G1(i, j + 1, k, 2) * (ex(i, j + 1, k + 1, n) - ex(i, j + 1, k - 1, n) + pdz) -
                                            G1(i, j, k, 2) * (pdz + ex(i, j - 1, k + 1, n) - ex(i, j - 1, k - 1, n));
                        const double JDzx = G2(i, j, k + 1, 0) * (ex(i + 1, j, k + 1, n) - ex(i - 1, j, k + 1, n) + pdx) -
                                            G2(i, j, k, 0) * (pdx + ex(i + 1, j, k - 1, n) - ex(i - 1, j, k - 1, n));
                        const double JDzy = G2(i, j, k + 1, 1) * (ex(i, j + 1, k + 1, n) - ex(i, j - 1, k + 1, n) + pdy) -
                                            G2(i, j, k, 1) * (pdy + ex(i, j + 1, k - 1, n) - ex(i, j - 1, k - 1, n));
                        const double JDzz = G2(i, j, k + 1, 2) * phi(i, j, k + 1, n) + G2(i, j, k, 2) * phi(i, j, k - 1, n);
                        lphi = beta * Jinv(i, j, k) *
                               (JDxx * xx + JDyy * yy + JDzz * zz + (JDxy + JDyx) * xy + (JDyz + JDzy) * yz + (JDzx + JDxz) * zx);
                    }
                    phi(i, j, k, n) = (rhs(i, j, k, n) - lphi) / (alpha + beta * lapDiag(i, j, k));
                }
            }
}

// GSRBBOUNDARYITER3DORTHO -- GSRBF.ChF:1362-1505, 2DORTHO -- :1233-1361.
// st[d][side] are BCType codes; a Neumann side is dropped from the stencil AND from the diagonal.
inline void gsrb_boundary_ortho(FAB& phi, const FAB& rhs, const FAB* Jg, const FAB& Jinv, const Box& region,
                                const double* dx, double alpha, double beta, const int st[3][2], int redBlack, int nd) {
    const double sc[3] = {1.0 / (dx[0] * dx[0]), 1.0 / (dx[1] * dx[1]), (nd > 2) ? 1.0 / (dx[2] * dx[2]) : 0.0};
    for (int n = 0; n < phi.nc; ++n)
        for (int k = region.lo[2]; k <= region.hi[2]; ++k)
            for (int j = region.lo[1]; j <= region.hi[1]; ++j) {
                const int imin = rb_imin(region.lo[0], j, k, redBlack, nd);
                for (int i = imin; i <= region.hi[0]; i += 2) {
                    double JDlo[3] = {0, 0, 0}, JDhi[3] = {0, 0, 0};
                    double lapDiag = 0.0;
                    // order of accumulation follows the Fortran: 3-D loX, loY, loZ, hiX, hiY, hiZ (GSRBF.ChF:1440-1480);
                    // 2-D loX, hiX, loY, hiY (:1300-1330 -- found by running the Fortran itself, oracle/chf2py.py)
                    auto lo_term = [&](int d) {
                        const int di = d == 0, dj = d == 1, dk = d == 2;
                        if (st[d][Lo] != BC_Neum) {
                            JDlo[d] = Jg[d](i, j, k, d) * phi(i - di, j - dj, k - dk, n);
                            lapDiag = lapDiag - sc[d] * Jg[d](i, j, k, d);
                        }
                    };
                    auto hi_term = [&](int d) {
                        const int di = d == 0, dj = d == 1, dk = d == 2;
                        if (st[d][Hi] != BC_Neum) {
                            JDhi[d] = Jg[d](i + di, j + dj, k + dk, d) * phi(i + di, j + dj, k + dk, n);
                            lapDiag = lapDiag - sc[d] * Jg[d](i + di, j + dj, k + dk, d);
                        }
                    };
                    if (nd > 2) {
                        for (int d = 0; d < nd; ++d) lo_term(d);
                        for (int d = 0; d < nd; ++d) hi_term(d);
                    } else {
                        for (int d = 0; d < nd; ++d) { lo_term(d); hi_term(d); }
                    }
                    lapDiag = lapDiag * Jinv(i, j, k);
                    double s = (JDlo[0] + JDhi[0]) * sc[0] + (JDlo[1] + JDhi[1]) * sc[1];
                    if (nd > 2) s = s + (JDlo[2] + JDhi[2]) * sc[2];
                    const double lphi = beta * Jinv(i, j, k) * s;
                    phi(i, j, k, n) = (rhs(i, j, k, n) - lphi) / (alpha + beta * lapDiag);
                }
            }
}

// Flux through one face of cell (i,j,k) in direction d, side s, as written in
// GSRBBOUNDARYITER3D -- GSRBF.ChF:1024-1232 and GSRBBOUNDARYITER2D -- :858-1022:
//   +/- [ aaScale*Jg_d(face,d)*phi(neighbour) +/- abScale*Jg_d(face,b)*(4-pt extrap diff) ... ]
// The normal term multiplies the NEIGHBOUR value only (the centre value lives in lapDiag).
inline double bdry_face_term(const FAB& phi, const FAB& ex, const FAB* Jg, const double* dx, int i, int j, int k, int n,
                             int d, int side, int nd) {
    const int e[3] = {d == 0, d == 1, d == 2};
    const int fo = (side == Hi) ? 1 : 0;                         // face index offset
    const int fi = i + fo * e[0], fj = j + fo * e[1], fk = k + fo * e[2];
    const int no = (side == Hi) ? 1 : -1;                        // neighbour offset
    const double sgn = (side == Hi) ? 1.0 : -1.0;
    double val = 0.0;
    // The Fortran lists the terms in direction order 0,1,2 (normal term sits at position d).
    for (int t = 0; t < nd; ++t) {
        if (t == d) {
            const double aa = 1.0 / (dx[d] * dx[d]);
            val += aa * Jg[d](fi, fj, fk, d) * phi(i + no * e[0], j + no * e[1], k + no * e[2], n);
        } else {
            const int b[3] = {t == 0, t == 1, t == 2};
            const double ab = 0.25 / (dx[d] * dx[t]);
            // cells on the two sides of the face: "outer" = towards side, "inner" = the other one
            const int oi = (side == Hi) ? i + e[0] : i, oj = (side == Hi) ? j + e[1] : j, ok = (side == Hi) ? k + e[2] : k;
            const int ii = oi - e[0], ij = oj - e[1], ik = ok - e[2];
            // ex(o+b)-ex(o-b)+ex(in+b)-ex(in-b): Hi: o = i+e, in = i.  Lo: o = i, in = i-e.
            const double diff = ex(oi + b[0], oj + b[1], ok + b[2], n) - ex(oi - b[0], oj - b[1], ok - b[2], n) +
                                ex(ii + b[0], ij + b[1], ik + b[2], n) - ex(ii - b[0], ij - b[1], ik - b[2], n);
            val += sgn * ab * Jg[d](fi, fj, fk, t) * diff;
        }
    }
    return val;
}

inline void gsrb_boundary_full(FAB& phi, const FAB& ex, const FAB& rhs, const FAB* Jg, const FAB& Jinv,
                               const Box& region, const double* dx, double alpha, double beta, const int st[3][2],
                               int redBlack, int nd) {
    for (int n = 0; n < phi.nc; ++n)
        for (int k = region.lo[2]; k <= region.hi[2]; ++k)
            for (int j = region.lo[1]; j <= region.hi[1]; ++j) {
                const int imin = rb_imin(region.lo[0], j, k, redBlack, nd);
                for (int i = imin; i <= region.hi[0]; i += 2) {
                    double JD[3][2] = {{0, 0}, {0, 0}, {0, 0}};
                    double lapDiag = 0.0;
                    // Fortran order: loX, hiX, loY, hiY, loZ, hiZ
                    for (int d = 0; d < nd; ++d)
                        for (int s = 0; s < 2; ++s) {
                            if (st[d][s] == BC_Neum) continue;
                            JD[d][s] = bdry_face_term(phi, ex, Jg, dx, i, j, k, n, d, s, nd);
                            const int fo = (s == Hi);
                            lapDiag = lapDiag - (1.0 / (dx[d] * dx[d])) * Jg[d](i + fo * (d == 0), j + fo * (d == 1), k + fo * (d == 2), d);
                        }
                    lapDiag = lapDiag * Jinv(i, j, k);
                    double s = JD[0][Lo] + JD[0][Hi] + JD[1][Lo] + JD[1][Hi];
                    if (nd > 2) s = s + JD[2][Lo] + JD[2][Hi];
                    const double lphi = beta * Jinv(i, j, k) * s;
                    phi(i, j, k, n) = (rhs(i, j, k, n) - lphi) / (alpha + beta * lapDiag);
                }
            }
}

// JACOBIITER -- JacobiF.ChF:31-61.  phi += 0.5 * res / (alpha + beta*lapDiag).
inline void jacobi_iter(FAB& phi, const FAB& res, const FAB& lapDiag, const Box& region, double alpha, double beta) {
    for (int n = 0; n < phi.nc; ++n)
        ORC_LOOP(region, i, j, k)
            phi(i, j, k, n) = phi(i, j, k, n) + 0.5 * res(i, j, k, n) / (alpha + beta * lapDiag(i, j, k));
}

}  // namespace orc

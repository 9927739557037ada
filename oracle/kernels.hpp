// oracle/kernels.hpp -- TEST INFRASTRUCTURE ONLY (CPU oracle; never the product path).
//
// Plain C++ restatement of the Chombo-Fortran kernels on SOMAR's pressure-solve hot path.
// Each function cites the reference file:line it follows (paths relative to /root/reference).
// The reference cannot be compiled here (needs Chombo 3.1 + Fortran + MPI + HDF5) and ships no golden
// vectors, but its Fortran kernels can be EXECUTED here: oracle/chf2py.py translates the .ChF sources mechanically and
// tests/test_reference_kernels.py runs every function of this file, of smoothers.hpp and of line_transfer.hpp next to the
// reference kernel it restates, on the same seeded arrays -- PINNED, bit for bit (fixtures tests/golden/refkernels_*.npz).
//
// "nd" = number of active directions: 3 for a 3-D operator, 2 for a 2-D operator (CH_SPACEDIM==2)
// or for the horizontal operator of a 3-D run (m_horizontalOp).  2-D boxes are one cell thick in
// direction 2 and every index expression below uses k offsets only when nd == 3, exactly as the
// CHF_DTERM/CHF_IX macros drop the third index when CH_SPACEDIM == 2.
#pragma once
#include "chombo_lite.hpp"
#include <cmath>

namespace orc {

// BCDescriptor.H:33-38
enum BCType { BC_Undefined = -2, BC_None = -1, BC_Neum = 0, BC_Diri = 1, BC_Periodic = 2, BC_CF = 3 };

#define ORC_LOOP(bx, i, j, k)                              \
    for (int k = (bx).lo[2]; k <= (bx).hi[2]; ++k)         \
        for (int j = (bx).lo[1]; j <= (bx).hi[1]; ++j)     \
            for (int i = (bx).lo[0]; i <= (bx).hi[0]; ++i)

// SUBTRACTOP -- MappedAMRPoissonOpF.ChF:36-54.  sub = a1 - a2 (out of place).
inline void subtractop(FAB& sub, const FAB& a1, const FAB& a2, const Box& region) {
    for (int n = 0; n < a1.nc; ++n) ORC_LOOP(region, i, j, k) sub(i, j, k, n) = a1(i, j, k, n) - a2(i, j, k, n);
}

// AXBYIP -- MappedAMRPoissonOpF.ChF:62-84.  lhs = alpha*phi + beta*lhs.
inline void axbyip(FAB& lhs, const FAB& phi, double alpha, double beta, const Box& region) {
    for (int n = 0; n < lhs.nc; ++n)
        ORC_LOOP(region, i, j, k) lhs(i, j, k, n) = alpha * phi(i, j, k, n) + beta * lhs(i, j, k, n);
}

// FILLMAPPEDLAPDIAG{1,2,3}D -- MappedAMRPoissonOpF.ChF:91-274.
// Jg[d] is the face-centred FAB in direction d; component d holds J g^{dd}.
inline void fill_lapdiag(FAB& lapDiag, const FAB* Jg, const FAB& Jinv, const Box& region, const double* dx, int nd) {
    const double s0 = 1.0 / (dx[0] * dx[0]);
    const double s1 = (nd > 1) ? 1.0 / (dx[1] * dx[1]) : 0.0;
    const double s2 = (nd > 2) ? 1.0 / (dx[2] * dx[2]) : 0.0;
    ORC_LOOP(region, i, j, k) {
        if (nd == 1) {
            lapDiag(i, j, k) = -Jinv(i, j, k) * s0 * (Jg[0](i + 1, j, k, 0) + Jg[0](i, j, k, 0));
        } else if (nd == 2) {
            lapDiag(i, j, k) = -Jinv(i, j, k) * ((Jg[0](i + 1, j, k, 0) + Jg[0](i, j, k, 0)) * s0 +
                                                 (Jg[1](i, j + 1, k, 1) + Jg[1](i, j, k, 1)) * s1);
        } else {
            lapDiag(i, j, k) = -Jinv(i, j, k) * ((Jg[0](i + 1, j, k, 0) + Jg[0](i, j, k, 0)) * s0 +
                                                 (Jg[1](i, j + 1, k, 1) + Jg[1](i, j, k, 1)) * s1 +
                                                 (Jg[2](i, j, k + 1, 2) + Jg[2](i, j, k, 2)) * s2);
        }
    }
}

// DIAGPRECOND -- MappedAMRPoissonOpF.ChF:284-328.
inline void diagprecond(FAB& phi, const FAB& rhs, const FAB& lapDiag, const Box& region, double alpha, double beta) {
    for (int n = 0; n < phi.nc; ++n)
        ORC_LOOP(region, i, j, k) phi(i, j, k, n) = rhs(i, j, k, n) / (alpha + beta * lapDiag(i, j, k));
}

// MAPPEDGETFLUXORTHO -- MappedAMRPoissonOpOrthoF.ChF:33-82.  flux = Jgaa * beta_dx * (phi(i) - phi(i-e_a)).
inline void getflux_ortho(FAB& flux, const FAB& phi, const FAB& Jga, int adir, const Box& fc, double beta_dx) {
    const int ai = adir == 0, aj = adir == 1, ak = adir == 2;
    for (int n = 0; n < flux.nc; ++n)
        ORC_LOOP(fc, i, j, k)
            flux(i, j, k, n) = Jga(i, j, k, adir) * beta_dx * (phi(i, j, k, n) - phi(i - ai, j - aj, k - ak, n));
}

// MAPPEDGETFLUX -- MappedAMRPoissonOpF.ChF:335-427 (nd == CH_SPACEDIM) and
// MAPPEDGETFLATFLUX3D -- :498-588 (nd == 2 inside a 3-D build: bdir = (adir+1) mod 2, no cdir).
inline void getflux_full(FAB& flux, const FAB& phi, const FAB& ex, const FAB& Jga, int adir, const Box& fc,
                         double beta, const double* dx, int nd) {
    const int bdir = (adir + 1) % nd;
    const int cdir = (adir + 2) % nd;
    const int ai = adir == 0, aj = adir == 1, ak = adir == 2;
    const int bi = bdir == 0, bj = bdir == 1, bk = bdir == 2;
    const int ci = cdir == 0, cj = cdir == 1, ck = cdir == 2;
    const double aS = beta / dx[adir];
    const double bS = 0.25 * beta / dx[bdir];
    const double cS = (nd > 2) ? 0.25 * beta / dx[cdir] : 0.0;
    for (int n = 0; n < flux.nc; ++n)
        ORC_LOOP(fc, i, j, k) {
            double f = aS * Jga(i, j, k, adir) * (phi(i, j, k, n) - phi(i - ai, j - aj, k - ak, n)) +
                       bS * Jga(i, j, k, bdir) *
                           (ex(i + bi, j + bj, k + bk, n) - ex(i - bi, j - bj, k - bk, n) +
                            ex(i + bi - ai, j + bj - aj, k + bk - ak, n) - ex(i - bi - ai, j - bj - aj, k - bk - ak, n));
            if (nd > 2)
                f += cS * Jga(i, j, k, cdir) *
                     (ex(i + ci, j + cj, k + ck, n) - ex(i - ci, j - cj, k - ck, n) +
                      ex(i + ci - ai, j + cj - aj, k + ck - ak, n) - ex(i - ci - ai, j - cj - aj, k - ck - ak, n));
            flux(i, j, k, n) = f;
        }
}

// MAPPEDFLUXDIVERGENCE{1,2,3}D -- DivCurlGradF.ChF:965-1214.  div = Jinv * sum_d (F_d(+) - F_d(-)) / dx_d.
inline void flux_divergence(FAB& div, const FAB* flux, const FAB& Jinv, const Box& region, const double* dx, int nd) {
    const double ix = 1.0 / dx[0];
    const double iy = (nd > 1) ? 1.0 / dx[1] : 0.0;
    const double iz = (nd > 2) ? 1.0 / dx[2] : 0.0;
    for (int n = 0; n < div.nc; ++n)
        ORC_LOOP(region, i, j, k) {
            double s = (flux[0](i + 1, j, k, n) - flux[0](i, j, k, n)) * ix;
            if (nd > 1) s = s + (flux[1](i, j + 1, k, n) - flux[1](i, j, k, n)) * iy;
            if (nd > 2) s = s + (flux[2](i, j, k + 1, n) - flux[2](i, j, k, n)) * iz;
            div(i, j, k, n) = Jinv(i, j, k) * s;
        }
}

// EXTRAPOLATEFACENOEV -- ExtrapolationUtilsF.ChF:35-140 (orders 0..4).
inline void extrapolate_face_noev(FAB& dest, const FAB& src, const Box& toRegion, int dir, int sidesign, int order) {
    const int ii = sidesign * (dir == 0), jj = sidesign * (dir == 1), kk = sidesign * (dir == 2);
    for (int n = 0; n < src.nc; ++n)
        ORC_LOOP(toRegion, i, j, k) {
            auto s = [&](int m) { return src(i - m * ii, j - m * jj, k - m * kk, n); };
            double v;
            switch (order) {
                case 0: v = s(1); break;
                case 1: v = 2.0 * s(1) - s(2); break;
                case 2: v = 3.0 * (s(1) - s(2)) + s(3); break;
                case 3: v = 4.0 * (s(1) + s(3)) - 6.0 * s(2) - s(4); break;
                case 4: v = 5.0 * (s(1) - s(4)) + 10.0 * (s(3) - s(2)) + s(5); break;
                default: fprintf(stderr, "EXTRAPOLATEFACENOEV: order %d is invalid\n", order); abort();
            }
            dest(i, j, k, n) = v;
        }
}

// ELLIPTICCONSTDIRIBCGHOST -- EllipticBCUtilsF.ChF:29-120 (orders 1 and 2, cell-centred ghosts).
//   order 1: ghost = 2*bcval - phi(near);  order 2: ghost = (8/3) bcval - 2 phi(near) + (1/3) phi(far)
inline void diri_ghost(FAB& phi, const Box& ghostBox, double bcval, int dir, int sidesign, int order) {
    const int ii = sidesign * (dir == 0), jj = sidesign * (dir == 1), kk = sidesign * (dir == 2);
    for (int n = 0; n < phi.nc; ++n)
        ORC_LOOP(ghostBox, i, j, k) {
            const double nearVal = phi(i - ii, j - jj, k - kk, n);
            if (order == 0) {
                phi(i, j, k, n) = bcval;
            } else if (order == 1) {
                phi(i, j, k, n) = (bcval == 0.0) ? -nearVal : 2.0 * bcval - nearVal;
            } else {
                const double farVal = phi(i - 2 * ii, j - 2 * jj, k - 2 * kk, n);
                phi(i, j, k, n) = (bcval == 0.0) ? (1.0 / 3.0) * farVal - 2.0 * nearVal
                                                 : (8.0 * bcval - 6.0 * nearVal + farVal) / 3.0;
            }
        }
}

// ELLIPTICCONSTNEUMBCGHOSTORTHO -- EllipticBCUtilsF.ChF:415-455.
inline void neum_ghost_ortho(FAB& phi, const FAB& nhat, int nhatComp, const Box& ghostBox, double bcval, int fdir,
                             int fsign, double dxDir) {
    const double factor = bcval * dxDir;
    const int ai = fsign * (fdir == 0), aj = fsign * (fdir == 1), ak = fsign * (fdir == 2);
    const int fo = (1 - fsign) / 2;
    const int fi = fo * (fdir == 0), fj = fo * (fdir == 1), fk = fo * (fdir == 2);
    for (int n = 0; n < phi.nc; ++n)
        ORC_LOOP(ghostBox, i, j, k)
            phi(i, j, k, n) = phi(i - ai, j - aj, k - ak, n) + factor / nhat(i + fi, j + fj, k + fk, nhatComp);
}

// ELLIPTICCONSTNEUMBCGHOST -- EllipticBCUtilsF.ChF:293-409 (full metric; cross terms from extrap).
inline void neum_ghost_full(FAB& phi, const FAB& ex, const FAB& nhat, const Box& ghostBox, double bcval, int fdir,
                            int fsign, const double* dx, int nd) {
    const int adir = fdir, bdir = (fdir + 1) % nd, cdir = (fdir + 2) % nd;
    const int bi = bdir == 0, bj = bdir == 1, bk = bdir == 2;
    const int ci = cdir == 0, cj = cdir == 1, ck = cdir == 2;
    const int fo = (1 - fsign) / 2;
    const int fio = fo * (adir == 0), fjo = fo * (adir == 1), fko = fo * (adir == 2);
    const int vio = -fsign * (adir == 0), vjo = -fsign * (adir == 1), vko = -fsign * (adir == 2);
    const double idxb = -0.25 / dx[bdir];
    const double idxc = (nd > 2) ? -0.25 / dx[cdir] : 0.0;
    for (int n = 0; n < phi.nc; ++n)
        ORC_LOOP(ghostBox, gi, gj, gk) {
            const int fi = gi + fio, fj = gj + fjo, fk = gk + fko;
            const int vi = gi + vio, vj = gj + vjo, vk = gk + vko;
            double cross = (ex(gi + bi, gj + bj, gk + bk, n) - ex(gi - bi, gj - bj, gk - bk, n) +
                            ex(vi + bi, vj + bj, vk + bk, n) - ex(vi - bi, vj - bj, vk - bk, n)) *
                           nhat(fi, fj, fk, bdir) * idxb;
            if (nd > 2)
                cross += (ex(gi + ci, gj + cj, gk + ck, n) - ex(gi - ci, gj - cj, gk - ck, n) +
                          ex(vi + ci, vj + cj, vk + ck, n) - ex(vi - ci, vj - cj, vk - ck, n)) *
                         nhat(fi, fj, fk, cdir) * idxc;
            phi(gi, gj, gk, n) = phi(vi, vj, vk, n) + (bcval - cross) * dx[adir] / nhat(fi, fj, fk, adir);
        }
}

}  // namespace orc

// oracle/kernel_capi.cpp -- TEST INFRASTRUCTURE ONLY (CPU oracle; never the product path).
// The oracle's hand-written kernel restatements (kernels.hpp, smoothers.hpp, line_transfer.hpp) callable one by one on raw
// arrays, so that tests/test_reference_kernels.py can run each of them next to the reference's own Fortran kernel
// (translated by oracle/chf2py.py) on identical inputs.  A raw FAB is (pointer, lo[3], hi[3], ncomp), x fastest,
// components outermost -- the layout of chombo_lite.hpp's FAB and of a C-ordered numpy array (nc, nz, ny, nx).
#include <cstring>

#include "line_transfer.hpp"
#include "smoothers.hpp"

namespace orc {
void macgrad_ortho(FAB& grad, const FAB& phi, const FAB& Jga, const Box& edgeBox, double dxDir, int dir);
void macgrad_full(FAB& grad, const FAB& phi, const FAB& ex, const FAB& Jga, const Box& edgeBox, const double* dx, int adir,
                  int nd);
}  // namespace orc

using namespace orc;

extern "C" {

struct RawFab {
    double* p;
    int lo[3], hi[3];
    int nc;
};

static FAB load(const RawFab* r) {
    FAB f(Box(IntVect(r->lo[0], r->lo[1], r->lo[2]), IntVect(r->hi[0], r->hi[1], r->hi[2])), r->nc);
    std::memcpy(f.d.data(), r->p, sizeof(double) * f.d.size());
    return f;
}
static void store(const FAB& f, RawFab* r) { std::memcpy(r->p, f.d.data(), sizeof(double) * f.d.size()); }
static Box box3(const int* lo, const int* hi) { return Box(IntVect(lo[0], lo[1], lo[2]), IntVect(hi[0], hi[1], hi[2])); }

void orck_gsrb_ortho(RawFab* phi, const RawFab* rhs, const RawFab* jg0, const RawFab* jg1, const RawFab* jg2,
                     const RawFab* jinv, const RawFab* lapDiag, const int* rlo, const int* rhi, const double* dx, double alpha,
                     double beta, int redBlack, int nd) {
    FAB p = load(phi);
    FAB Jg[3] = {load(jg0), load(jg1), nd > 2 ? load(jg2) : FAB()};
    gsrb_ortho(p, load(rhs), Jg, load(jinv), load(lapDiag), box3(rlo, rhi), dx, alpha, beta, redBlack, nd);
    store(p, phi);
}

void orck_gsrb_full(RawFab* phi, const RawFab* ex, const RawFab* rhs, const RawFab* jg0, const RawFab* jg1, const RawFab* jg2,
                    const RawFab* jinv, const RawFab* lapDiag, const int* rlo, const int* rhi, const double* dx, double alpha,
                    double beta, int redBlack, int nd) {
    FAB p = load(phi);
    FAB Jg[3] = {load(jg0), load(jg1), nd > 2 ? load(jg2) : FAB()};
    gsrb_full(p, load(ex), load(rhs), Jg, load(jinv), load(lapDiag), box3(rlo, rhi), dx, alpha, beta, redBlack, nd);
    store(p, phi);
}

void orck_gsrb_boundary_ortho(RawFab* phi, const RawFab* rhs, const RawFab* jg0, const RawFab* jg1, const RawFab* jg2,
                              const RawFab* jinv, const int* rlo, const int* rhi, const double* dx, double alpha, double beta,
                              const int* st6, int redBlack, int nd) {
    FAB p = load(phi);
    FAB Jg[3] = {load(jg0), load(jg1), nd > 2 ? load(jg2) : FAB()};
    const int st[3][2] = {{st6[0], st6[1]}, {st6[2], st6[3]}, {st6[4], st6[5]}};
    gsrb_boundary_ortho(p, load(rhs), Jg, load(jinv), box3(rlo, rhi), dx, alpha, beta, st, redBlack, nd);
    store(p, phi);
}

void orck_gsrb_boundary_full(RawFab* phi, const RawFab* ex, const RawFab* rhs, const RawFab* jg0, const RawFab* jg1,
                             const RawFab* jg2, const RawFab* jinv, const int* rlo, const int* rhi, const double* dx,
                             double alpha, double beta, const int* st6, int redBlack, int nd) {
    FAB p = load(phi);
    FAB Jg[3] = {load(jg0), load(jg1), nd > 2 ? load(jg2) : FAB()};
    const int st[3][2] = {{st6[0], st6[1]}, {st6[2], st6[3]}, {st6[4], st6[5]}};
    gsrb_boundary_full(p, load(ex), load(rhs), Jg, load(jinv), box3(rlo, rhi), dx, alpha, beta, st, redBlack, nd);
    store(p, phi);
}

void orck_line_gsrb(RawFab* phi, const RawFab* ex, const RawFab* rhs, const RawFab* jg0, const RawFab* jg1, const RawFab* jg2,
                    const RawFab* jinv, const int* rlo, const int* rhi, const double* dx, double dzCrse, double alpha, double beta,
                    int redBlack, const int* st6, int nd, int isDiagonal) {
    FAB p = load(phi);
    FAB Jg[3] = {load(jg0), load(jg1), nd > 2 ? load(jg2) : FAB()};
    const int st[3][2] = {{st6[0], st6[1]}, {st6[2], st6[3]}, {st6[4], st6[5]}};
    line_gsrb(p, load(ex), load(rhs), Jg, load(jinv), box3(rlo, rhi), dx, dzCrse, alpha, beta, redBlack, st, nd, isDiagonal != 0);
    store(p, phi);
}

void orck_jacobi(RawFab* phi, const RawFab* res, const RawFab* lapDiag, const int* rlo, const int* rhi, double alpha, double beta) {
    FAB p = load(phi);
    jacobi_iter(p, load(res), load(lapDiag), box3(rlo, rhi), alpha, beta);
    store(p, phi);
}

void orck_fill_lapdiag(RawFab* lapDiag, const RawFab* jg0, const RawFab* jg1, const RawFab* jg2, const RawFab* jinv,
                       const int* rlo, const int* rhi, const double* dx, int nd) {
    FAB l = load(lapDiag);
    FAB Jg[3] = {load(jg0), nd > 1 ? load(jg1) : FAB(), nd > 2 ? load(jg2) : FAB()};
    fill_lapdiag(l, Jg, load(jinv), box3(rlo, rhi), dx, nd);
    store(l, lapDiag);
}

void orck_diagprecond(RawFab* phi, const RawFab* rhs, const RawFab* lapDiag, const int* rlo, const int* rhi, double alpha,
                      double beta) {
    FAB p = load(phi);
    diagprecond(p, load(rhs), load(lapDiag), box3(rlo, rhi), alpha, beta);
    store(p, phi);
}

void orck_getflux_ortho(RawFab* flux, const RawFab* phi, const RawFab* jga, int adir, const int* rlo, const int* rhi,
                        double beta_dx) {
    FAB f = load(flux);
    getflux_ortho(f, load(phi), load(jga), adir, box3(rlo, rhi), beta_dx);
    store(f, flux);
}

void orck_getflux_full(RawFab* flux, const RawFab* phi, const RawFab* ex, const RawFab* jga, int adir, const int* rlo,
                       const int* rhi, double beta, const double* dx, int nd) {
    FAB f = load(flux);
    getflux_full(f, load(phi), load(ex), load(jga), adir, box3(rlo, rhi), beta, dx, nd);
    store(f, flux);
}

void orck_flux_divergence(RawFab* div, const RawFab* f0, const RawFab* f1, const RawFab* f2, const RawFab* jinv, const int* rlo,
                          const int* rhi, const double* dx, int nd) {
    FAB d = load(div);
    FAB fl[3] = {load(f0), nd > 1 ? load(f1) : FAB(), nd > 2 ? load(f2) : FAB()};
    flux_divergence(d, fl, load(jinv), box3(rlo, rhi), dx, nd);
    store(d, div);
}

void orck_macgrad_ortho(RawFab* grad, const RawFab* phi, const RawFab* jga, const int* rlo, const int* rhi, double dxDir, int dir) {
    FAB g = load(grad);
    macgrad_ortho(g, load(phi), load(jga), box3(rlo, rhi), dxDir, dir);
    store(g, grad);
}

void orck_macgrad_full(RawFab* grad, const RawFab* phi, const RawFab* ex, const RawFab* jga, const int* rlo, const int* rhi,
                       const double* dx, int adir, int nd) {
    FAB g = load(grad);
    macgrad_full(g, load(phi), load(ex), load(jga), box3(rlo, rhi), dx, adir, nd);
    store(g, grad);
}

void orck_extrapolate_face_noev(RawFab* dest, const RawFab* src, const int* rlo, const int* rhi, int dir, int sidesign, int order) {
    FAB d = load(dest);
    extrapolate_face_noev(d, load(src), box3(rlo, rhi), dir, sidesign, order);
    store(d, dest);
}

void orck_diri_ghost(RawFab* phi, const int* rlo, const int* rhi, double bcval, int dir, int sidesign, int order) {
    FAB p = load(phi);
    diri_ghost(p, box3(rlo, rhi), bcval, dir, sidesign, order);
    store(p, phi);
}

void orck_neum_ghost_ortho(RawFab* phi, const RawFab* nhat, int nhatComp, const int* rlo, const int* rhi, double bcval, int fdir,
                           int fsign, double dxDir) {
    FAB p = load(phi);
    neum_ghost_ortho(p, load(nhat), nhatComp, box3(rlo, rhi), bcval, fdir, fsign, dxDir);
    store(p, phi);
}

void orck_neum_ghost_full(RawFab* phi, const RawFab* ex, const RawFab* nhat, const int* rlo, const int* rhi, double bcval,
                          int fdir, int fsign, const double* dx, int nd) {
    FAB p = load(phi);
    neum_ghost_full(p, load(ex), load(nhat), box3(rlo, rhi), bcval, fdir, fsign, dx, nd);
    store(p, phi);
}

void orck_mappedaverage2(RawFab* coarse, const RawFab* fine, const RawFab* fineJinv, const int* clo, const int* chi, const int* r) {
    FAB c = load(coarse);
    mappedaverage2(c, load(fine), load(fineJinv), box3(clo, chi), IntVect(r[0], r[1], r[2]));
    store(c, coarse);
}

void orck_average_harmonic(RawFab* coarse, const RawFab* fine, const int* clo, const int* chi, const int* r) {
    FAB c = load(coarse);
    average_harmonic(c, load(fine), box3(clo, chi), IntVect(r[0], r[1], r[2]));
    store(c, coarse);
}

void orck_average_face(RawFab* coarse, const RawFab* fine, const int* clo, const int* chi, int dir, const int* r) {
    FAB c = load(coarse);
    average_face(c, load(fine), box3(clo, chi), dir, IntVect(r[0], r[1], r[2]));
    store(c, coarse);
}

void orck_const_interp(RawFab* fine, const RawFab* coarse, const int* flo, const int* fhi, const int* m) {
    FAB f = load(fine);
    const_interp(f, load(coarse), box3(flo, fhi), IntVect(m[0], m[1], m[2]));
    store(f, fine);
}

void orck_const_interp_with_avg(RawFab* fine, const RawFab* coarse, const int* flo, const int* fhi, const int* m,
                                const RawFab* jinv, double dxProduct, double* vol, double* sum) {
    FAB f = load(fine);
    const_interp_with_avg(f, load(coarse), box3(flo, fhi), IntVect(m[0], m[1], m[2]), load(jinv), dxProduct, *vol, *sum);
    store(f, fine);
}

}  // extern "C"

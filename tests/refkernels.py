"""Kernel-by-kernel comparison harness (used by tests/test_reference_kernels.py and tools/make_reference_goldens.py).

Every case builds seeded inputs on small boxes with non-zero origins, runs ONE kernel of the reference -- its own
Chombo-Fortran source executed through oracle/chf2py.py -- and the oracle's hand-written restatement of that kernel
(oracle/kernel_capi.cpp for the C++ ones, oracle/advect_ref.py for the numpy ones), and returns both outputs.  The
arguments follow the reference's call sites, which are cited case by case."""
import ctypes as C
import itertools
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")


class RawFab(C.Structure):
    _fields_ = [("p", C.c_void_p), ("lo", C.c_int * 3), ("hi", C.c_int * 3), ("nc", C.c_int)]


def _lib():
    if not hasattr(_lib, "h"):
        _lib.h = C.CDLL(ORACLE_SO)
    return _lib.h


class Arr:
    """a FAB for both sides: numpy (nc, [nz,] ny, nx) over the box [lo, hi] (D entries)"""

    def __init__(self, lo, hi, nc=1, rng=None, fill=None, positive=False):
        self.lo, self.hi, self.D = list(lo), list(hi), len(lo)
        shp = (nc,) + tuple(hi[d] - lo[d] + 1 for d in reversed(range(self.D)))
        if fill is not None:
            self.a = np.full(shp, float(fill))
        elif positive:
            self.a = rng.uniform(0.5, 1.5, size=shp)
        else:
            self.a = rng.uniform(-1.0, 1.0, size=shp)

    def copy(self):
        o = Arr.__new__(Arr)
        o.lo, o.hi, o.D, o.a = list(self.lo), list(self.hi), self.D, self.a.copy()
        return o

    def fab(self, comp0=0):            # for the translated Fortran
        from oracle.chf_runtime import Fab
        return Fab(self.a, self.lo, comp0=comp0)

    def raw(self):                     # for the C++ oracle
        r = RawFab()
        r.p = self.a.ctypes.data
        for d in range(3):
            r.lo[d] = self.lo[d] if d < self.D else 0
            r.hi[d] = self.hi[d] if d < self.D else 0
        r.nc = self.a.shape[0]
        return C.byref(r)


def _i3(v, pad=0):
    v = list(v) + [pad] * (3 - len(v))
    return (C.c_int * 3)(*v)


def _d3(v):
    v = list(v) + [1.0] * (3 - len(v))
    return (C.c_double * 3)(*v)


def _box(lo, hi):
    from oracle.chf_runtime import Box
    return Box(lo, hi)


def _vec(v, integer=False):
    from oracle.chf_runtime import Vec
    return Vec(v, integer=integer)


def _grow(lo, hi, g, dirs=None):
    D = len(lo)
    dirs = range(D) if dirs is None else dirs
    lo, hi = list(lo), list(hi)
    for d in dirs:
        lo[d] -= g
        hi[d] += g
    return lo, hi


def _faces(lo, hi, d):
    hi = list(hi)
    hi[d] += 1
    return list(lo), hi


class Setup:
    """the fields of one box of a level: valid box with non-zero origin, phi / extrap with one ghost layer, face-centred
    J g^{d.} with SpaceDim components (FluxBox), cell-centred 1/J and lapDiag"""

    def __init__(self, D, seed):
        rng = np.random.default_rng(seed)
        self.D = D
        self.lo = [2, -1, 3][:D]
        self.hi = [self.lo[d] + [5, 4, 3][d] for d in range(D)]
        self.dx = [0.5, 0.25, 0.125][:D]
        glo, ghi = _grow(self.lo, self.hi, 1)
        self.phi = Arr(glo, ghi, 1, rng)
        self.ex = Arr(glo, ghi, 1, rng)
        self.rhs = Arr(self.lo, self.hi, 1, rng)
        self.jinv = Arr(glo, ghi, 1, rng, positive=True)
        self.lapdiag = Arr(self.lo, self.hi, 1, rng)
        self.lapdiag.a -= 3.0                              # away from zero (it is a divisor)
        self.jg = []
        for d in range(D):
            flo, fhi = _faces(*_grow(self.lo, self.hi, 1, [e for e in range(D) if e != d]), d)
            a = Arr(flo, fhi, D, rng, positive=True)
            for e in range(D):
                if e != d:
                    a.a[e] = 0.3 * (a.a[e] - 1.0)          # cross terms: smaller, either sign
            self.jg.append(a)
        self.rng = rng


def _jg_raw(s):
    return [s.jg[d].raw() if d < s.D else None for d in range(3)]


# ---------------------------------------------------------------------------------------------------------------
# Each case: name -> function(D, ref_module_or_None) -> (oracle_outputs, reference_outputs_or_None); outputs are dicts of
# numpy arrays.  Call sites cited are the reference's.
# ---------------------------------------------------------------------------------------------------------------
def case_gsrb_ortho(D, ref):
    """GSRBITER{2,3}DORTHO <- BaseGSRB::fullStencilGSRB, GSRB.cpp:405-445"""
    out_o, out_r = {}, ({} if ref else None)
    for (alpha, beta), rb in itertools.product([(0.0, 1.0), (0.3, 0.7)], [0, 1]):
        s = Setup(D, 11)
        key = "a%.1f_b%.1f_rb%d" % (alpha, beta, rb)
        p = s.phi.copy()
        _lib().orck_gsrb_ortho(p.raw(), s.rhs.raw(), *_jg_raw(s), s.jinv.raw(), s.lapdiag.raw(), _i3(s.lo), _i3(s.hi), _d3(s.dx),
                               C.c_double(alpha), C.c_double(beta), rb, D)
        out_o[key] = p.a
        if ref:
            q = s.phi.copy()
            f = getattr(ref, "gsrbiter%ddortho" % D)
            f(q.fab(), s.rhs.fab(), *[s.jg[d].fab(d) for d in range(D)], s.jinv.fab(), s.lapdiag.fab(), _box(s.lo, s.hi),
              _vec(s.dx), alpha, beta, rb)
            out_r[key] = q.a
    return out_o, out_r


def case_gsrb_full(D, ref):
    """GSRBITER{2,3}D <- BaseGSRB::fullStencilGSRB, GSRB.cpp:447-487"""
    out_o, out_r = {}, ({} if ref else None)
    for (alpha, beta), rb in itertools.product([(0.0, 1.0), (0.3, 0.7)], [0, 1]):
        s = Setup(D, 12)
        key = "a%.1f_b%.1f_rb%d" % (alpha, beta, rb)
        p = s.phi.copy()
        _lib().orck_gsrb_full(p.raw(), s.ex.raw(), s.rhs.raw(), *_jg_raw(s), s.jinv.raw(), s.lapdiag.raw(), _i3(s.lo), _i3(s.hi),
                              _d3(s.dx), C.c_double(alpha), C.c_double(beta), rb, D)
        out_o[key] = p.a
        if ref:
            q = s.phi.copy()
            f = getattr(ref, "gsrbiter%dd" % D)
            f(q.fab(), s.ex.fab(), s.rhs.fab(), *[s.jg[d].fab() for d in range(D)], s.jinv.fab(), s.lapdiag.fab(),
              _box(s.lo, s.hi), _vec(s.dx), alpha, beta, rb)
            out_r[key] = q.a
    return out_o, out_r


# BCType codes per side (lo x, hi x, lo y, ...): -1 none (box interface), 0 Neumann, 1 Dirichlet, 2 periodic.  (All-Neumann is
# left out: every term of the diagonal is dropped and the Fortran divides by zero.)
STENCILS = {2: [(0, 1, 0, 1), (0, -1, 1, 0), (-1, 0, -1, 1), (2, 2, 0, 1), (1, 1, 1, 1), (-1, -1, -1, -1)],
            3: [(0, 1, 0, 1, 0, 1), (0, -1, 1, 0, -1, -1), (-1, 0, -1, 1, 0, 1), (2, 2, 0, 1, -1, 0), (1, 1, 1, 1, 1, 1),
                (-1, -1, -1, -1, -1, -1)]}


def _case_gsrb_boundary(D, ref, full):
    out_o, out_r = {}, ({} if ref else None)
    for st, rb, (alpha, beta) in itertools.product(STENCILS[D], [0, 1], [(0.0, 1.0), (0.3, 0.7)]):
        s = Setup(D, 13)
        key = "st%s_rb%d_a%.1f" % ("".join(str(x + 2) for x in st), rb, alpha)
        st6 = (C.c_int * 6)(*(list(st) + [0] * (6 - len(st))))
        p = s.phi.copy()
        if full:
            _lib().orck_gsrb_boundary_full(p.raw(), s.ex.raw(), s.rhs.raw(), *_jg_raw(s), s.jinv.raw(), _i3(s.lo), _i3(s.hi),
                                           _d3(s.dx), C.c_double(alpha), C.c_double(beta), st6, rb, D)
        else:
            _lib().orck_gsrb_boundary_ortho(p.raw(), s.rhs.raw(), *_jg_raw(s), s.jinv.raw(), _i3(s.lo), _i3(s.hi), _d3(s.dx),
                                            C.c_double(alpha), C.c_double(beta), st6, rb, D)
        out_o[key] = p.a
        if ref:
            q = s.phi.copy()
            if full:
                getattr(ref, "gsrbboundaryiter%dd" % D)(q.fab(), s.ex.fab(), s.rhs.fab(), *[s.jg[d].fab() for d in range(D)],
                                                        s.jinv.fab(), _box(s.lo, s.hi), _vec(s.dx), alpha, beta, *st, rb)
            else:
                getattr(ref, "gsrbboundaryiter%ddortho" % D)(q.fab(), s.rhs.fab(), *[s.jg[d].fab(d) for d in range(D)],
                                                             s.jinv.fab(), _box(s.lo, s.hi), _vec(s.dx), alpha, beta, *st, rb)
            out_r[key] = q.a
    return out_o, out_r


def case_gsrb_boundary_ortho(D, ref):
    """GSRBBOUNDARYITER{2,3}DORTHO <- BaseGSRB::boundaryGSRB, GSRB.cpp:489-653 (stencil codes = BCType of every side)"""
    return _case_gsrb_boundary(D, ref, False)


def case_gsrb_boundary_full(D, ref):
    """GSRBBOUNDARYITER{2,3}D <- BaseGSRB::boundaryGSRB, GSRB.cpp:489-653"""
    return _case_gsrb_boundary(D, ref, True)


def case_jacobi(D, ref):
    """JACOBIITER <- Jacobi::relax, Jacobi.cpp:50-90"""
    s = Setup(D, 14)
    p = s.phi.copy()
    _lib().orck_jacobi(p.raw(), s.rhs.raw(), s.lapdiag.raw(), _i3(s.lo), _i3(s.hi), C.c_double(0.3), C.c_double(0.7))
    out_r = None
    if ref:
        q = s.phi.copy()
        ref.jacobiiter(q.fab(), s.rhs.fab(), s.lapdiag.fab(), _box(s.lo, s.hi), 0.3, 0.7)
        out_r = {"phi": q.a}
    return {"phi": p.a}, out_r


def case_lapdiag(D, ref):
    """FILLMAPPEDLAPDIAG{2,3}D <- MappedAMRPoissonOpFactory::fill_MGfields, MappedAMRPoissonOpFactory.cpp:1199-1230"""
    s = Setup(D, 15)
    l = s.lapdiag.copy()
    _lib().orck_fill_lapdiag(l.raw(), *_jg_raw(s), s.jinv.raw(), _i3(s.lo), _i3(s.hi), _d3(s.dx), D)
    out_r = None
    if ref:
        q = s.lapdiag.copy()
        getattr(ref, "fillmappedlapdiag%dd" % D)(q.fab(), *[s.jg[d].fab() for d in range(D)], s.jinv.fab(), _box(s.lo, s.hi),
                                                 _vec(s.dx))
        out_r = {"lapdiag": q.a}
    return {"lapdiag": l.a}, out_r


def case_diagprecond(D, ref):
    """DIAGPRECOND <- MappedAMRPoissonOp::preCond, MappedAMRPoissonOp.cpp:684-734"""
    s = Setup(D, 16)
    p = s.phi.copy()
    _lib().orck_diagprecond(p.raw(), s.rhs.raw(), s.lapdiag.raw(), _i3(s.lo), _i3(s.hi), C.c_double(0.3), C.c_double(0.7))
    out_r = None
    if ref:
        q = s.phi.copy()
        ref.diagprecond(q.fab(), s.rhs.fab(), s.lapdiag.fab(), _box(s.lo, s.hi), 0.3, 0.7)
        out_r = {"phi": q.a}
    return {"phi": p.a}, out_r


def case_getflux(D, ref):
    """MAPPEDGETFLUX / MAPPEDGETFLUXORTHO <- MappedAMRPoissonOp::getFluxComplete, MappedAMRPoissonOp.cpp:2070-2160"""
    out_o, out_r = {}, ({} if ref else None)
    beta = 0.7
    for adir in range(D):
        s = Setup(D, 17 + adir)
        flo, fhi = _faces(s.lo, s.hi, adir)
        fo, ff = Arr(flo, fhi, 1, fill=0.0), Arr(flo, fhi, 1, fill=0.0)
        _lib().orck_getflux_ortho(fo.raw(), s.phi.raw(), s.jg[adir].raw(), adir, _i3(flo), _i3(fhi), C.c_double(beta / s.dx[adir]))
        # interior faces in the transverse directions only where the cross stencil stays inside phi + 1 ghost
        _lib().orck_getflux_full(ff.raw(), s.phi.raw(), s.ex.raw(), s.jg[adir].raw(), adir, _i3(flo), _i3(fhi), C.c_double(beta),
                                 _d3(s.dx), D)
        out_o["ortho%d" % adir], out_o["full%d" % adir] = fo.a, ff.a
        if ref:
            ro, rf = Arr(flo, fhi, 1, fill=0.0), Arr(flo, fhi, 1, fill=0.0)
            ref.mappedgetfluxortho(ro.fab(), s.phi.fab(), s.jg[adir].fab(adir), _box(flo, fhi), beta / s.dx[adir], adir)
            ref.mappedgetflux(rf.fab(), s.phi.fab(), s.ex.fab(), s.jg[adir].fab(), _box(flo, fhi), beta, _vec(s.dx), adir)
            out_r["ortho%d" % adir], out_r["full%d" % adir] = ro.a, rf.a
    return out_o, out_r


def case_divergence(D, ref):
    """MAPPEDFLUXDIVERGENCE{2,3}D <- MappedAMRPoissonOp::applyOpI (:880-893), Divergence::levelDivergenceMAC (Divergence.cpp:103-123)"""
    s = Setup(D, 21)
    fl = [Arr(*_faces(s.lo, s.hi, d), 1, s.rng) for d in range(D)]
    d_o = Arr(s.lo, s.hi, 1, fill=0.0)
    _lib().orck_flux_divergence(d_o.raw(), *[fl[d].raw() if d < D else None for d in range(3)], s.jinv.raw(), _i3(s.lo), _i3(s.hi),
                                _d3(s.dx), D)
    out_r = None
    if ref:
        d_r = Arr(s.lo, s.hi, 1, fill=0.0)
        getattr(ref, "mappedfluxdivergence%dd" % D)(d_r.fab(), *[f.fab() for f in fl], s.jinv.fab(), _box(s.lo, s.hi), _vec(s.dx))
        out_r = {"div": d_r.a}
    return {"div": d_o.a}, out_r


def case_macgrad(D, ref):
    """MAPPEDMACGRAD / MAPPEDMACGRADORTHO, normal derivatives <- Gradient::singleBoxMacGrad, Gradient.cpp:1056-1078"""
    out_o, out_r = {}, ({} if ref else None)
    for adir in range(D):
        s = Setup(D, 23 + adir)
        flo, fhi = _faces(s.lo, s.hi, adir)
        go, gf = Arr(flo, fhi, 1, fill=0.0), Arr(flo, fhi, 1, fill=0.0)
        _lib().orck_macgrad_ortho(go.raw(), s.phi.raw(), s.jg[adir].raw(), _i3(flo), _i3(fhi), C.c_double(s.dx[adir]), adir)
        _lib().orck_macgrad_full(gf.raw(), s.phi.raw(), s.ex.raw(), s.jg[adir].raw(), _i3(flo), _i3(fhi), _d3(s.dx), adir, D)
        out_o["ortho%d" % adir], out_o["full%d" % adir] = go.a, gf.a
        if ref:
            ro, rf = Arr(flo, fhi, 1, fill=0.0), Arr(flo, fhi, 1, fill=0.0)
            ref.mappedmacgradortho(ro.fab(), s.phi.fab(), s.ex.fab(), s.jg[adir].fab(), _box(flo, fhi), s.dx[adir], adir, adir)
            ref.mappedmacgrad(rf.fab(), s.phi.fab(), s.ex.fab(), s.jg[adir].fab(), _box(flo, fhi), _vec(s.dx), adir, adir)
            out_r["ortho%d" % adir], out_r["full%d" % adir] = ro.a, rf.a
    return out_o, out_r


def _ghost_box(s, d, side):
    lo, hi = list(s.lo), list(s.hi)
    if side == 0:
        lo[d] = hi[d] = s.lo[d] - 1
    else:
        lo[d] = hi[d] = s.hi[d] + 1
    return lo, hi


def case_extrapolate(D, ref):
    """EXTRAPOLATEFACENOEV <- ExtrapolateFaceNoEV, ExtrapolationUtils.cpp:34-62"""
    out_o, out_r = {}, ({} if ref else None)
    for d, side, order in itertools.product(range(D), [0, 1], [0, 1, 2, 3, 4]):
        if order >= [5, 4, 3][d]:
            continue                                  # the stencil must fit in the box
        s = Setup(D, 31)
        glo, ghi = _ghost_box(s, d, side)
        sign = 1 if side else -1
        p = s.phi.copy()
        _lib().orck_extrapolate_face_noev(p.raw(), s.phi.raw(), _i3(glo), _i3(ghi), d, sign, order)
        key = "d%d_s%d_o%d" % (d, side, order)
        out_o[key] = p.a
        if ref:
            q = s.phi.copy()
            ref.extrapolatefacenoev(q.fab(), s.phi.fab(), _box(glo, ghi), d, sign, order)
            out_r[key] = q.a
    return out_o, out_r


def case_bc_ghosts(D, ref):
    """ELLIPTICCONSTDIRIBCGHOST, ELLIPTICCONSTNEUMBCGHOST[ORTHO] <- setSideDiriBC / setSideNeumBC, EllipticBCUtils.cpp:41-215"""
    out_o, out_r = {}, ({} if ref else None)
    for d, side in itertools.product(range(D), [0, 1]):
        sign = 1 if side else -1
        for order, bcval in itertools.product([0, 1, 2], [0.0, 0.4]):
            s = Setup(D, 33)
            glo, ghi = _ghost_box(s, d, side)
            p = s.phi.copy()
            _lib().orck_diri_ghost(p.raw(), _i3(glo), _i3(ghi), C.c_double(bcval), d, sign, order)
            key = "diri_d%d_s%d_o%d_v%.1f" % (d, side, order, bcval)
            out_o[key] = p.a
            if ref:
                q = s.phi.copy()
                ref.ellipticconstdiribcghost(q.fab(), _box(glo, ghi), bcval, d, sign, order)
                out_r[key] = q.a
        for bcval in (0.0, 0.4):
            s = Setup(D, 34)
            glo, ghi = _ghost_box(s, d, side)
            p, pf = s.phi.copy(), s.phi.copy()
            _lib().orck_neum_ghost_ortho(p.raw(), s.jg[d].raw(), d, _i3(glo), _i3(ghi), C.c_double(bcval), d, sign, C.c_double(s.dx[d]))
            # the full-metric ghost reads extrap one cell beyond the ghost box in the transverse directions: shrink it
            flo, fhi = list(glo), list(ghi)
            for e in range(D):
                if e != d:
                    flo[e] += 1
                    fhi[e] -= 1
            _lib().orck_neum_ghost_full(pf.raw(), s.ex.raw(), s.jg[d].raw(), _i3(flo), _i3(fhi), C.c_double(bcval), d, sign,
                                        _d3(s.dx), D)
            ko, kf = "neumortho_d%d_s%d_v%.1f" % (d, side, bcval), "neumfull_d%d_s%d_v%.1f" % (d, side, bcval)
            out_o[ko], out_o[kf] = p.a, pf.a
            if ref:
                q, qf = s.phi.copy(), s.phi.copy()
                ref.ellipticconstneumbcghostortho(q.fab(), s.jg[d].fab(d), _box(glo, ghi), bcval, d, sign, s.dx[d])
                ref.ellipticconstneumbcghost(qf.fab(), s.ex.fab(), s.jg[d].fab(), _box(flo, fhi), bcval, d, sign, _vec(s.dx))
                out_r[ko], out_r[kf] = q.a, qf.a
    return out_o, out_r


def case_transfer(D, ref):
    """MAPPEDAVERAGE2, UNMAPPEDAVERAGEHARMONIC, UNMAPPEDAVERAGEFACE, ConstInterpPS, ConstInterpWithAvgPS <-
    MappedCoarseAverage.cpp:170-215,447-470, ProlongationStrategy.cpp:49-165"""
    out_o, out_r = {}, ({} if ref else None)
    rng = np.random.default_rng(41)
    ratios = {2: [(2, 2), (1, 2), (2, 1)], 3: [(1, 1, 2), (2, 2, 2), (2, 1, 2)]}[D]
    for r in ratios:
        clo = [1, -2, 0][:D]
        chi = [clo[d] + [2, 3, 1][d] for d in range(D)]
        flo = [clo[d] * r[d] for d in range(D)]
        fhi = [(chi[d] + 1) * r[d] - 1 for d in range(D)]
        fine, jinv = Arr(flo, fhi, 1, rng), Arr(flo, fhi, 1, rng, positive=True)
        key = "r" + "".join(map(str, r))
        co, ch = Arr(clo, chi, 1, fill=0.0), Arr(clo, chi, 1, fill=0.0)
        _lib().orck_mappedaverage2(co.raw(), fine.raw(), jinv.raw(), _i3(clo), _i3(chi), _i3(r, 1))
        _lib().orck_average_harmonic(ch.raw(), jinv.raw(), _i3(clo), _i3(chi), _i3(r, 1))
        out_o["avg2_" + key], out_o["harm_" + key] = co.a, ch.a
        crse = Arr(clo, chi, 1, rng)
        f1, f2 = fine.copy(), fine.copy()
        _lib().orck_const_interp(f1.raw(), crse.raw(), _i3(flo), _i3(fhi), _i3(r, 1))
        vol, tot = C.c_double(0.0), C.c_double(0.0)
        _lib().orck_const_interp_with_avg(f2.raw(), crse.raw(), _i3(flo), _i3(fhi), _i3(r, 1), jinv.raw(), C.c_double(0.37),
                                          C.byref(vol), C.byref(tot))
        out_o["interp_" + key], out_o["interpavg_" + key] = f1.a, f2.a
        out_o["interpavg_sums_" + key] = np.array([vol.value, tot.value])
        if ref:
            bref = _box([0] * D, [r[d] - 1 for d in range(D)])
            rc, rh = Arr(clo, chi, 1, fill=0.0), Arr(clo, chi, 1, fill=0.0)
            ref.mappedaverage2(rc.fab(), fine.fab(), jinv.fab(), _box(clo, chi), _vec(r, integer=True), bref)
            ref.unmappedaverageharmonic(rh.fab(), jinv.fab(), _box(clo, chi), _vec(r, integer=True), bref)
            out_r["avg2_" + key], out_r["harm_" + key] = rc.a, rh.a
            # the C++ caller shifts both boxes so that the fine box starts at 0 (ProlongationStrategy.cpp:69-78)
            g1, g2 = fine.copy(), fine.copy()
            sh_f = [0] * D
            sh_fh = [fhi[d] - flo[d] for d in range(D)]
            from oracle.chf_runtime import Fab
            shift = lambda arr, lo: Fab(arr.a, lo)
            ref.constinterpps(shift(g1, sh_f), shift(crse, [0] * D), _box(sh_f, sh_fh), _vec(r, integer=True))
            vs = _vec([0.0])
            ss = _vec([0.0])
            res = ref.constinterpwithavgps(shift(g2, sh_f), shift(crse, [0] * D), _box(sh_f, sh_fh), _vec(r, integer=True),
                                           shift(jinv, sh_f), 0.37, 0.0, 0.0)
            out_r["interp_" + key], out_r["interpavg_" + key] = g1.a, g2.a
            out_r["interpavg_sums_" + key] = np.array(res) if res is not None else None
        # face average in every direction
        for d in range(D):
            cflo, cfhi = _faces(clo, chi, d)
            fflo = [cflo[e] * r[e] for e in range(D)]
            ffhi = [(cfhi[e] + 1) * r[e] - 1 if e != d else cfhi[e] * r[e] for e in range(D)]
            ffine = Arr(fflo, ffhi, 1, rng)
            cf = Arr(cflo, cfhi, 1, fill=0.0)
            _lib().orck_average_face(cf.raw(), ffine.raw(), _i3(cflo), _i3(cfhi), d, _i3(r, 1))
            out_o["face%d_%s" % (d, key)] = cf.a
            if ref:
                rf = Arr(cflo, cfhi, 1, fill=0.0)
                rb_hi = [r[e] - 1 if e != d else 0 for e in range(D)]
                ref.unmappedaverageface(rf.fab(), ffine.fab(), _box(cflo, cfhi), d, _vec(r, integer=True), _box([0] * D, rb_hi))
                out_r["face%d_%s" % (d, key)] = rf.a
    return out_o, out_r


CASES = {
    "gsrb_ortho": case_gsrb_ortho, "gsrb_full": case_gsrb_full, "gsrb_boundary_ortho": case_gsrb_boundary_ortho,
    "gsrb_boundary_full": case_gsrb_boundary_full, "jacobi": case_jacobi, "lapdiag": case_lapdiag,
    "diagprecond": case_diagprecond, "getflux": case_getflux, "divergence": case_divergence, "macgrad": case_macgrad,
    "extrapolate": case_extrapolate, "bc_ghosts": case_bc_ghosts, "transfer": case_transfer,
}


def reference_module(D):
    """the translated reference kernels for CH_SPACEDIM = D, or None where neither /root/reference nor oracle/_ref exists"""
    import importlib
    from oracle import chf2py
    path = os.path.join(chf2py.OUT, "chf_kernels_%dd.py" % D)
    if os.path.isdir(chf2py.REF):
        chf2py.build()
    if not os.path.exists(path):
        return None
    return importlib.import_module("oracle._ref.chf_kernels_%dd" % D)


# ---------------------------------------------------------------------------------------------------------------
# Advection predictor: the numpy oracle (oracle/advect_ref.py, periodic global arrays) against the reference's Fortran
# kernels called in the order of the reference's C++ (cited), on the same arrays continued periodically by four ghost
# layers (ADVECT_GROW).  With every direction periodic the lo / hi one-sided boxes of PeriodicLoHiCenter* are empty.
# ---------------------------------------------------------------------------------------------------------------
G = 4


def _ext(a, D):
    """(nz, ny, nx) or (ny, nx) periodic array -> Arr over [-G, n + G - 1] in every direction"""
    b = np.pad(a, [(G, G)] * D, mode="wrap")
    o = Arr.__new__(Arr)
    o.D = D
    o.lo = [-G] * D
    o.hi = [a.shape[D - 1 - d] + G - 1 for d in range(D)]
    o.a = np.ascontiguousarray(b[None])
    return o


def _like(x, fill=0.0):
    o = x.copy()
    o.a[...] = fill
    return o


def _inner(x, n, D):
    sl = (0,) + tuple(slice(G, G + n[D - 1 - q]) for q in range(D))
    return x.a[sl]


def _shrunk(x, d, g):
    lo, hi = list(x.lo), list(x.hi)
    lo[d] += g
    hi[d] -= g
    return lo, hi


def _ref_van_leer(ref, W, d, lim, ho):
    """MappedGodunovUtilities::vanLeerSlopes, MappedGodunovUtilities.cpp:49-102 (periodic: centre box only)"""
    D = W.D
    dW, dWm, dWp = _like(W), _like(W), _like(W)
    empty = _box([0] * D, [-1] * D)
    c1 = _shrunk(W, d, 1)
    ref.secondslopediffsf(dW.fab(), dWm.fab(), dWp.fab(), W.fab(), 1, d, empty, 0, empty, 0, _box(*c1))
    if lim and ho:
        ref.extpreservingvanleerlimiterf(dW.fab(), dWm.fab(), dWp.fab(), 1, d, _box(*_shrunk(W, d, 2)))
    elif lim:
        ref.vanleerlimiterf(dW.fab(), dWm.fab(), dWp.fab(), 1, _box(*c1))
    return dW, dWm, dWp


def _ref_faces(ref, X, d, lim, ho):
    """one interpolation of MappedAdvectionUtil::PPMFaceValues (AdvectUtil.cpp:1246-1418): face f = low face of cell f"""
    D = X.D
    empty = _box([0] * D, [-1] * D)
    F = _like(X)
    fl, fh = _shrunk(X, d, 3)
    if ho:
        ref.fourthinterpfaces(F.fab(), X.fab(), 1, d, empty, empty, 0, empty, empty, 0, _box(fl, fh))
    else:
        dX, _, _ = _ref_van_leer(ref, X, d, lim, False)
        ref.ppmfacevaluesf(F.fab(), X.fab(), dX.fab(), 1, d, empty, 0, empty, 0, _box(fl, fh))
    return F


def case_advect(D, ref):
    """SecondSlopeDiffsF, VanLeerLimiterF, ExtPreservingVanLeerLimiterF, FourthSlopeDiffsF, PPMFaceValuesF, FourthInterpFaces,
    PPMLimiterF, GETSECONDDIFF, CheckCubicLimiterF, PPMNormalPredF, PLMNormalPredF, RiemannSolver, AvgStates,
    QuasilinearUpdateF <- MappedAdvectionUtil::PPMNormalPred / PLMNormalPred / PPMFaceValues / RiemannSolver /
    quasilinearUpdate (AdvectUtil.cpp:944-1418) and MappedGodunovUtilities (MappedGodunovUtilities.cpp:49-500)"""
    from oracle import advect_ref as ar
    rng = np.random.default_rng(51)
    n = [9, 8, 7][:D]
    shape = tuple(reversed(n))
    z = np.meshgrid(*[np.arange(m) / m for m in shape], indexing="ij")
    W = np.cos(2 * np.pi * z[-1]) * np.sin(2 * np.pi * z[-2] + 0.3) + 0.35 * rng.uniform(-1, 1, shape)
    J = 1.0 + 0.2 * rng.uniform(-1, 1, shape)
    vel = [rng.uniform(-1, 1, shape) for _ in range(D)]
    adv = [rng.uniform(-1, 1, shape) for _ in range(D)]
    for a in adv:
        a[..., ::4] = 0.0
    dtbydx = 0.37
    out_o, out_r = {}, ({} if ref else None)
    W3 = W if D == 3 else W[None]
    J3 = J if D == 3 else J[None]
    as3 = lambda a: a if D == 3 else a[None]
    un3 = lambda a: a if D == 3 else a[0]
    for d in range(D):
        for lim, ho in itertools.product([False, True], repeat=2):
            tag = "d%d_l%d_h%d" % (d, lim, ho)
            out_o["vl_" + tag] = un3(ar.van_leer_slopes(W3, d, lim, ho))
            out_o["face_" + tag] = un3(ar.ppm_face_values(W3, J3, d, lim, ho))
            m, p = ar.ppm_normal_pred(W3, J3, as3(vel[d]), d, dtbydx, lim, ho)
            out_o["ppm_m_" + tag], out_o["ppm_p_" + tag] = un3(m), un3(p)
            for fourth in (False, True):
                m, p = ar.plm_normal_pred(W3, as3(vel[d]), d, dtbydx, lim, fourth, ho)
                out_o["plm%d_m_%s" % (fourth, tag)], out_o["plm%d_p_%s" % (fourth, tag)] = un3(m), un3(p)
        wp, wm = rng.uniform(-1, 1, shape), rng.uniform(-1, 1, shape)
        for up in (True, False):
            out_o["riemann_d%d_u%d" % (d, up)] = un3(ar.riemann(as3(wp), as3(wm), as3(adv[d]), d, up))
        out_o["quasi_d%d" % d] = un3(ar.quasilinear_update(as3(wp), as3(vel[d]), -0.21, d))
        if not ref:
            continue
        eW, eJ, eV = _ext(W, D), _ext(J, D), _ext(vel[d], D)
        eJW = _ext(W * J, D)          # JW.copy(a_W); JW *= ccJ (AdvectUtil.cpp:1338-1341)
        empty = _box([0] * D, [-1] * D)
        for lim, ho in itertools.product([False, True], repeat=2):
            tag = "d%d_l%d_h%d" % (d, lim, ho)
            dW, _, _ = _ref_van_leer(ref, eW, d, lim, ho)
            out_r["vl_" + tag] = _inner(dW, n, D)
            # PPMFaceValues: interpolate J and J W with the same weights, divide (:1253-1417)
            FJ, FW = _ref_faces(ref, eJ, d, lim, ho), _ref_faces(ref, eJW, d, lim, ho)
            face = _like(eW)
            with np.errstate(divide="ignore", invalid="ignore"):
                face.a[...] = np.where(FJ.a != 0.0, FW.a / np.where(FJ.a != 0.0, FJ.a, 1.0), 0.0)
            out_r["face_" + tag] = _inner(face, n, D)
            # PPMNormalPred (:1158-1240): deltas from the two faces of the cell, limiter, predictor, "+= W"
            dm, dp = _like(eW), _like(eW)
            sl_hi = [slice(None)] * (D + 1)
            sl_lo = [slice(None)] * (D + 1)
            ax = D - d                                   # numpy axis of direction d in (1, [nz,] ny, nx)
            sl_hi[ax] = slice(1, None)
            sl_lo[ax] = slice(0, -1)
            dm.a[...] = (0.0 - eW.a) + face.a
            dp.a[tuple(sl_lo)] = (0.0 - eW.a[tuple(sl_lo)]) + face.a[tuple(sl_hi)]
            box = _box(*_shrunk(eW, d, 3))
            if lim and ho:                               # MappedGodunovUtilities::PPMLimiter, :386-461
                d2fcf = _like(eW)
                d2fcf.a[...] = (dm.a + dp.a) * 6.0
                d2W = _like(eW)
                ref.getseconddiff(d2W.fab(), eW.fab(), 1, d, empty, 0, empty, 0, _box(*_shrunk(eW, d, 1)))
                ref.checkcubiclimiterf(dm.fab(), dp.fab(), eW.fab(), d2W.fab(), d2fcf.fab(), 1, d, empty, empty, 0, empty, empty, 0,
                                       box, 1.25, 0.1, 1.0e-12)
            elif lim:
                ref.ppmlimiterf(dm.fab(), dp.fab(), 1, box)
            ref.ppmnormalpredf(dm.fab(), dp.fab(), eV.fab(), dtbydx, 1, box)
            out_r["ppm_m_" + tag] = _inner(dm, n, D) + W
            out_r["ppm_p_" + tag] = _inner(dp, n, D) + W
            # PLMNormalPred (:1066-1152)
            for fourth in (False, True):
                if fourth:
                    dvl, _, _ = _ref_van_leer(ref, eW, d, lim, ho)
                    dW4 = _like(eW)
                    ref.fourthslopediffsf(dW4.fab(), eW.fab(), dvl.fab(), 1, d, empty, 0, empty, 0, box)
                    _, dWm, dWp = _ref_van_leer(ref, eW, d, False, False)      # oneSidedDifferences (:143-171)
                    slope = dW4
                else:
                    slope, _, _ = _ref_van_leer(ref, eW, d, lim, ho)
                    dWm, dWp = _like(eW), _like(eW)                            # a_WMinus / a_WPlus .setVal(0.0)
                if lim or fourth:
                    ref.vanleerlimiterf(slope.fab(), dWm.fab(), dWp.fab(), 1, box)
                pm, pp = _like(eW), _like(eW)
                ref.plmnormalpredf(pm.fab(), pp.fab(), slope.fab(), eV.fab(), dtbydx, 1, box)
                out_r["plm%d_m_%s" % (fourth, tag)] = _inner(pm, n, D) + W
                out_r["plm%d_p_%s" % (fourth, tag)] = _inner(pp, n, D) + W
        # RiemannSolver (:944-998): the left state is the W+ of the cell below the face (shiftHalf)
        ewp, ewm, eadv = _ext(wp, D), _ext(wm, D), _ext(adv[d], D)
        from oracle.chf_runtime import Fab
        lo_shift = list(ewp.lo)
        lo_shift[d] += 1
        left = Fab(ewp.a, lo_shift)
        fbox = _box(*_shrunk(eW, d, 2))
        for up in (True, False):
            g = _like(eW)
            if up:
                ref.riemannsolver(g.fab(), left, ewm.fab(), eadv.fab(), d, fbox)
            else:
                ref.avgstates(g.fab(), left, ewm.fab(), d, fbox)
            out_r["riemann_d%d_u%d" % (d, up)] = _inner(g, n, D)
        q = _like(eW)
        ref.quasilinearupdatef(q.fab(), ewp.fab(), eV.fab(), -0.21, d, fbox)
        out_r["quasi_d%d" % d] = _inner(q, n, D)
    return out_o, out_r


CASES["advect"] = case_advect


def case_leptic(D, ref):
    """TRIDIAGPOISSONNN1DFAB (src/utils/TridiagUtilsF.ChF:85-164) <- LevelLepticSolver::verticalLineSolver (:1392-1404);
    UNMAPPEDVERTINTEGRAL, ADDEXTRUSION (src/utils/SubspaceF.ChF) and LEPTICACCUMDIV (LevelLepticSolverF.ChF:291-331) <-
    LevelLepticSolver::computeHorizRHS / addHorizontalCorrection.  Three-dimensional build only (the vertical is z)."""
    if D != 3:
        return {}, ({} if ref else None)
    from oracle import pyoracle as po
    from oracle.chf_runtime import Fab
    rng = np.random.default_rng(61)
    nx, ny, nv = 5, 4, 7
    lo = [2, -1, 0]
    dz = 0.37
    rhs = rng.uniform(-1, 1, (nv, ny, nx))
    jgzz = rng.uniform(0.5, 1.5, (nv + 1, ny, nx))
    out_o, out_r = {}, ({} if ref else None)
    phi, _, info = po.leptic_vertical_solve(rhs, jgzz, dz, 0, 0)
    assert info == 0
    out_o["tridiag_nn"] = phi
    flat0 = rng.uniform(-1, 1, (ny, nx))
    out_o["vert_integral"] = po.leptic_vert_integral(flat0, rhs, dz)
    out_o["add_extrusion"] = po.leptic_add_extrusion(rhs, flat0)
    fl = [rng.uniform(-1, 1, (ny, nx + 1)), rng.uniform(-1, 1, (ny + 1, nx))]
    for d in range(2):
        out_o["accum_div%d" % d] = po.leptic_accum_div(flat0, fl[d], d, -1.0 / 0.21)
    if ref:
        hi = [lo[0] + nx - 1, lo[1] + ny - 1, lo[2] + nv - 1]
        p = np.zeros((1, nv, ny, nx))
        ref.tridiagpoissonnn1dfab(Fab(p, lo), Fab(np.ascontiguousarray(rhs[None]), lo), Fab(np.ascontiguousarray(jgzz[None]), lo),
                                  _box(lo, [hi[0], hi[1], lo[2]]), nv, dz, 2)
        out_r["tridiag_nn"] = p[0]
        f = flat0[None, None].copy()
        ref.unmappedvertintegral(Fab(f, [lo[0], lo[1], 0]), Fab(np.ascontiguousarray(rhs[None]), lo), _box(lo, hi), dz)
        out_r["vert_integral"] = f[0, 0]
        dst = np.ascontiguousarray(rhs[None].copy())
        ref.addextrusion(Fab(dst, lo), Fab(flat0[None, None].copy(), [lo[0], lo[1], 0]), _box(lo, hi))
        out_r["add_extrusion"] = dst[0]
        for d in range(2):
            r2 = np.ascontiguousarray(flat0[None, None].copy())
            ref.lepticaccumdiv(Fab(r2, [lo[0], lo[1], 0]), Fab(np.ascontiguousarray(fl[d][None, None]), [lo[0], lo[1], 0]),
                               _box([lo[0], lo[1], 0], [hi[0], hi[1], 0]), -1.0 / 0.21, d)
            out_r["accum_div%d" % d] = r2[0, 0]
    return out_o, out_r


CASES["leptic"] = case_leptic


def case_line_gsrb(D, ref):
    """LineGSRBIter2D (GSRBF.ChF:1529-1729) <- LineGSRB::relax, GSRB.cpp:148-330 (work FABs D, B over [1, Nz], DU over
    [2, Nz], DL over [1, Nz - 1]; the box is shifted so that the column starts at 0; dgtsv is LAPACK's).
    LineGSRBIter3D (:1730-2042) compares `j` with jmin / jmax, which it never assigns (:1764,1831,1835): undefined in
    Fortran.  The tests sit in `(loYBCType .ne. BCType_Neum) .or. (j .ne. jmin)`, so with non-Neumann y sides the kernel's
    result does not depend on them and it IS pinned here (every vertical side type, both colours, diagonal and full
    metric).  Its x test compares `i` with the parity-shifted imin (:1795,1821) -- the first cell of the colour, not the
    first cell of the box -- so with a Neumann x side the rows whose colour starts one cell in are treated differently from
    the 2-D kernel; on the other rows it agrees with the box bound, and those rows are pinned too (the others are blanked
    on both sides).  The oracle restates the box bounds (its header says so)."""
    if D == 3:
        return _case_line_gsrb_3d(ref)
    from oracle.chf_runtime import Fab
    out_o, out_r = {}, ({} if ref else None)
    for st, rb, full, (alpha, beta) in itertools.product([(0, 1, 0, 0), (-1, -1, 0, 1), (1, 0, 1, 0), (2, 2, 0, 0)], [0, 1],
                                                         [False, True], [(0.0, 1.0), (0.3, 0.7)]):
        s = Setup(2, 71)
        # the caller's shift: the vertical index of the box starts at 0
        sh = s.lo[1]
        for a in [s.phi, s.ex, s.rhs, s.jinv, s.lapdiag] + s.jg:
            a.lo[1] -= sh
            a.hi[1] -= sh
        lo, hi = [s.lo[0], 0], [s.hi[0], s.hi[1] - sh]
        if not full:
            for d in range(2):
                s.jg[d].a[1 - d] = 0.0
        key = "st%s_rb%d_f%d_a%.1f" % ("".join(str(x + 2) for x in st), rb, full, alpha)
        st6 = (C.c_int * 6)(*(list(st) + [0, 0]))
        p = s.phi.copy()
        _lib().orck_line_gsrb(p.raw(), s.ex.raw(), s.rhs.raw(), s.jg[0].raw(), s.jg[1].raw(), None, s.jinv.raw(), _i3(lo), _i3(hi),
                              _d3(s.dx), C.c_double(2.0 * s.dx[1]), C.c_double(alpha), C.c_double(beta), rb, st6, 2, int(not full))
        out_o[key] = p.a
        if ref:
            q = s.phi.copy()
            nz = hi[1] + 1
            Dw, Bw = Fab(np.zeros((1, 1, nz)), [1, 1]), Fab(np.zeros((1, 1, nz)), [1, 1])
            DUw, DLw = Fab(np.zeros((1, 1, nz - 1)), [2, 1]), Fab(np.zeros((1, 1, nz - 1)), [1, 1])
            ref.linegsrbiter2d(q.fab(), s.ex.fab(), s.rhs.fab(), s.jg[0].fab(), s.jg[1].fab(), s.jinv.fab(), _box(lo, hi), _vec(s.dx),
                               2.0 * s.dx[1], alpha, beta, rb, DUw, Dw, DLw, Bw, *st)
            out_r[key] = q.a
    return out_o, out_r


def _case_line_gsrb_3d(ref):
    from oracle.chf_runtime import Fab
    out_o, out_r = {}, ({} if ref else None)
    # (loX, hiX, loY, hiY, loZ, hiZ); 0 Neumann, 1 Dirichlet, 3 coarse-fine, -1 / 2 anything else
    sts = [(1, 1, 1, -1, 0, 1), (-1, 2, 1, 1, 1, 0), (1, -1, 2, 1, 3, 3), (2, 1, -1, -1, -1, 0), (0, 0, 1, 1, 0, 0),
           (0, 1, -1, 1, 1, 3)]
    for st, rb, full, (alpha, beta) in itertools.product(sts, [0, 1], [False, True], [(0.0, 1.0), (0.3, 0.7)]):
        s = Setup(3, 73)
        sh = s.lo[2]
        for a in [s.phi, s.ex, s.rhs, s.jinv, s.lapdiag] + s.jg:
            a.lo[2] -= sh
            a.hi[2] -= sh
        lo, hi = [s.lo[0], s.lo[1], 0], [s.hi[0], s.hi[1], s.hi[2] - sh]
        if not full:
            for d in range(3):
                for e in range(3):
                    if e != d:
                        s.jg[d].a[e] = 0.0
        key = "st%s_rb%d_f%d_a%.1f" % ("".join(str(x + 2) for x in st), rb, full, alpha)
        # rows the reference's x test reads as intended: all of them without a Neumann x side, else those whose colour
        # starts on the first cell of the box
        rows = [j for j in range(lo[1], hi[1] + 1) if (st[0] != 0) or ((lo[0] + j + rb) % 2 == 0)]

        def keep(arr):                                            # arr: (1, nz + 2, ny + 2, nx + 2) with one ghost layer
            out = np.zeros_like(arr)
            for j in rows:
                out[:, :, j - (lo[1] - 1), :] = arr[:, :, j - (lo[1] - 1), :]
            return out
        st6 = (C.c_int * 6)(*st)
        p = s.phi.copy()
        _lib().orck_line_gsrb(p.raw(), s.ex.raw(), s.rhs.raw(), s.jg[0].raw(), s.jg[1].raw(), s.jg[2].raw(), s.jinv.raw(),
                              _i3(lo), _i3(hi), _d3(s.dx), C.c_double(2.0 * s.dx[2]), C.c_double(alpha), C.c_double(beta), rb,
                              st6, 3, int(not full))
        out_o[key] = keep(p.a)
        if ref:
            q = s.phi.copy()
            nz = hi[2] + 1
            Dw, Bw = Fab(np.zeros((1, 1, 1, nz)), [1, 1, 1]), Fab(np.zeros((1, 1, 1, nz)), [1, 1, 1])
            DUw, DLw = Fab(np.zeros((1, 1, 1, nz - 1)), [2, 1, 1]), Fab(np.zeros((1, 1, 1, nz - 1)), [1, 1, 1])
            ref.linegsrbiter3d(q.fab(), s.ex.fab(), s.rhs.fab(), s.jg[0].fab(), s.jg[1].fab(), s.jg[2].fab(), s.jinv.fab(),
                               _box(lo, hi), _vec(s.dx), 2.0 * s.dx[2], alpha, beta, rb, DUw, Dw, DLw, Bw, *st)
            out_r[key] = keep(q.a)
    return out_o, out_r


CASES["line_gsrb"] = case_line_gsrb


def _adj(lo, hi, d, side):
    lo, hi = list(lo), list(hi)
    if side == 0:
        lo[d] = hi[d] = lo[d] - 1
    else:
        lo[d] = hi[d] = hi[d] + 1
    return lo, hi


def _isect(a, b):
    lo = [max(x, y) for x, y in zip(a[0], b[0])]
    hi = [min(x, y) for x, y in zip(a[1], b[1])]
    return lo, hi


def case_leptic_bdry_gradient(D, ref):
    """LEPTICVERTHORIZGRAD (LevelLepticSolverF.ChF:74-122) <- LevelLepticSolver::levelVertHorizGradient (:1107-1176), with its
    in-place ExtrapolateFaceAndCopy (ExtrapolationUtils.cpp:109-163, EXTRAPOLATEFACENOEV) restated box by box in this
    harness, against the oracle's Op::lepticVertHorizGradient on a single-box terrain-following level"""
    if D != 3:
        return {}, ({} if ref else None)
    from common import Problem
    from oracle import pyoracle as po
    from oracle.chf_runtime import Fab
    pb = Problem(n=(10, 8, 6), map_name="terrain", aniso=4.0)
    orc = pb.oracle(po)
    phi = np.random.default_rng(81).uniform(-1, 1, pb.shape)
    lo_o, hi_o = orc.leptic_vert_horiz_grad(phi, -1.0)
    out_o, out_r = {"lo": lo_o, "hi": hi_o}, None
    if ref:
        nx, ny, nz = pb.n
        dom = ([0, 0, 0], [nx - 1, ny - 1, nz - 1])
        fab = ([-1, -1, -1], [nx, ny, nz])
        out_r = {}
        for side, name in ((0, "lo"), (1, "hi")):
            p = np.zeros((1, nz + 2, ny + 2, nx + 2))
            p[0, 1:-1, 1:-1, 1:-1] = phi
            P = Fab(p, fab[0])
            sign = 1 if side else -1

            def noev(box, d, s):                                   # ExtrapolateFaceNoEV, :34-62
                to = _isect(_adj(box[0], box[1], d, s), fab)
                if all(a <= b for a, b in zip(*to)):
                    ref.extrapolatefacenoev(P, P, _box(*to), d, 1 if s else -1, 2)
            valid = _isect(fab, dom)                               # domValid = phiFAB.box() & domain
            noev(valid, 2, side)
            ghost = _isect(_adj(valid[0], valid[1], 2, side), fab)
            near = ([ghost[0][0], ghost[0][1], ghost[0][2] - sign], [ghost[1][0], ghost[1][1], ghost[1][2] - sign])
            for edir in (0, 1):
                for es in (0, 1):
                    noev(ghost, edir, es)
                    noev(near, edir, es)
                for b in (ghost, near):
                    b[0][edir] -= 1
                    b[1][edir] += 1
                ghost, near = _isect(ghost, fab), _isect(near, fab)
            kf = nz if side else 0
            bc = np.zeros((1, 1, ny, nx))
            jgz = np.ascontiguousarray(np.asarray(pb.jgup[2]))       # (3, nz + 1, ny, nx)
            ref.lepticverthorizgrad(Fab(bc, [0, 0, kf]), P, Fab(jgz, [0, 0, 0]), _box([0, 0, kf], [nx - 1, ny - 1, kf]), sign,
                                    _vec(pb.spec.dx), -1.0)
            out_r[name] = bc[0, 0]
    return out_o, out_r


CASES["leptic_bdry_gradient"] = case_leptic_bdry_gradient

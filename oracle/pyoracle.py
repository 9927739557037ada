"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
The product package (somar_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

BC_NEUM, BC_DIRI, BC_PERIODIC = 0, 1, 2
RELAX_JACOBI, RELAX_LEVEL_GSRB, RELAX_LOOSE_GSRB, RELAX_LINE_GSRB = 0, 1, 2, 3


class OrcDesc(C.Structure):
    _fields_ = [("dim", C.c_int), ("n", C.c_int * 3), ("periodic", C.c_int * 3), ("nboxes", C.c_int),
                ("box_lo", C.POINTER(C.c_int)), ("box_hi", C.POINTER(C.c_int)), ("dx", C.c_double * 3),
                ("alpha", C.c_double), ("beta", C.c_double), ("is_diagonal", C.c_int), ("bc_type", C.c_int * 6),
                ("bc_value", C.c_double * 6), ("jgup_ncomp", C.c_int), ("jgup", C.c_void_p * 3), ("jinv", C.c_void_p)]


class OrcParams(C.Structure):
    _fields_ = [("eps", C.c_double), ("num_smooth_down", C.c_int), ("num_smooth_up", C.c_int),
                ("num_smooth_bottom", C.c_int), ("num_smooth_precond", C.c_int), ("numMG", C.c_int), ("imin", C.c_int),
                ("imax", C.c_int), ("hang", C.c_double), ("normThresh", C.c_double), ("maxDepth", C.c_int),
                ("relaxMode", C.c_int), ("precondMode", C.c_int), ("bottom_eps", C.c_double), ("bottom_reps", C.c_double),
                ("bottom_hang", C.c_double), ("bottom_small", C.c_double), ("bottom_imax", C.c_int),
                ("bottom_numRestarts", C.c_int), ("bottom_normType", C.c_int), ("verbosity", C.c_int)]


class OrcStats(C.Structure):
    _fields_ = [("iters", C.c_int), ("exitStatus", C.c_int), ("initial_rnorm", C.c_double), ("final_rnorm", C.c_double),
                ("bottom_iters_total", C.c_int), ("history_len", C.c_int), ("history", C.c_double * 128)]


def build(debug: bool = False) -> str:
    target = "liboracle_dbg.so" if debug else "liboracle.so"
    subprocess.run(["make", "-C", _HERE, target], check=True, stdout=subprocess.DEVNULL)
    return os.path.join(_HERE, target)


def lib(debug: bool = False):
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle_dbg.so" if debug else "liboracle.so")
        if not os.path.exists(path):
            build(debug)
        L = C.CDLL(path)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.POINTER(OrcDesc), C.POINTER(OrcParams)]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_num_depths.argtypes = [C.c_void_p]
        L.orc_norm.restype = C.c_double
        L.orc_timed_solve.restype = C.c_double
        _LIB = L
    return _LIB


def default_params(**kw) -> OrcParams:
    p = OrcParams()
    lib().orc_default_params(C.byref(p))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise KeyError(k)
        setattr(p, k, v)
    return p


def _f64(a) -> np.ndarray:
    if hasattr(a, "detach"):
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(a, dtype=np.float64)


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """One single-level mapped-Poisson problem with its multigrid hierarchy, on the CPU oracle."""

    def __init__(self, n: Sequence[int], dx: Sequence[float], jgup: Sequence, jinv, *, dim=3, periodic=(0, 0, 0),
                 boxes=None, alpha=0.0, beta=1.0, is_diagonal=True, bc_type=None, bc_value=None, jgup_ncomp=None,
                 params: Optional[OrcParams] = None):
        self.L = lib()
        self.n = tuple(int(x) for x in n)
        self.dim = dim
        if boxes is None:
            boxes = [((0, 0, 0), tuple(x - 1 for x in self.n))]
        self.boxes = boxes
        lo = np.ascontiguousarray([b[0] for b in boxes], dtype=np.int32)
        hi = np.ascontiguousarray([b[1] for b in boxes], dtype=np.int32)
        self._keep = [lo, hi]
        d = OrcDesc()
        d.dim = dim
        d.n[:] = self.n
        d.periodic[:] = [int(bool(x)) for x in periodic]
        d.nboxes = len(boxes)
        d.box_lo = lo.ctypes.data_as(C.POINTER(C.c_int))
        d.box_hi = hi.ctypes.data_as(C.POINTER(C.c_int))
        d.dx[:] = [float(x) for x in dx]
        d.alpha, d.beta = float(alpha), float(beta)
        d.is_diagonal = int(bool(is_diagonal))
        bt = bc_type if bc_type is not None else [BC_NEUM] * 6
        bv = bc_value if bc_value is not None else [0.0] * 6
        d.bc_type[:] = [int(x) for x in bt]
        d.bc_value[:] = [float(x) for x in bv]
        jg = [_f64(j) for j in jgup]
        ji = _f64(jinv)
        self._keep += jg + [ji]
        if jgup_ncomp is None:
            jgup_ncomp = jg[0].shape[0] if jg[0].ndim == 4 else 1
        d.jgup_ncomp = int(jgup_ncomp)
        for i in range(dim):
            d.jgup[i] = jg[i].ctypes.data
        d.jinv = ji.ctypes.data
        self.params = params if params is not None else default_params()
        self.h = C.c_void_p(self.L.orc_create(C.byref(d), C.byref(self.params)))
        self.ndepths = self.L.orc_num_depths(self.h)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.L.orc_destroy(self.h)
                self.h = None
        except Exception:
            pass

    # -- hierarchy inspection ---------------------------------------------------------------
    def depth_info(self, depth: int):
        n = (C.c_int * 3)()
        r = (C.c_int * 3)()
        dx = (C.c_double * 3)()
        z = C.c_int()
        nb = C.c_int()
        self.L.orc_depth_info(self.h, depth, n, r, dx, C.byref(z), C.byref(nb))
        return {"n": tuple(n), "ratio": tuple(r), "dx": tuple(dx), "zero_avg": bool(z.value), "nboundary": nb.value}

    def shape(self, depth: int):
        n = self.depth_info(depth)["n"]
        return (n[2], n[1], n[0])

    def metric(self, depth: int):
        n = self.depth_info(depth)["n"]
        nd = self.dim
        jg = []
        for d in range(3):
            if d < nd:
                jg.append(np.zeros((nd, n[2] + (d == 2), n[1] + (d == 1), n[0] + (d == 0))))
            else:
                jg.append(None)
        jinv = np.zeros(self.shape(depth))
        lap = np.zeros(self.shape(depth))
        self.L.orc_get_metric(self.h, depth, *[(_ptr(j) if j is not None else None) for j in jg], _ptr(jinv), _ptr(lap))
        return jg, jinv, lap

    def boundary_boxes(self, depth: int, simple=False):
        cap = 4096
        out = np.zeros((cap, 13), dtype=np.int32)
        cnt = self.L.orc_boundary_boxes(self.h, depth, int(simple), out.ctypes.data_as(C.POINTER(C.c_int)), cap)
        return out[:min(cnt, cap)]

    # -- operator ---------------------------------------------------------------------------
    def apply_op(self, phi, depth=0, homogeneous=True):
        phi = _f64(phi)
        out = np.zeros_like(phi)
        self.L.orc_apply_op(self.h, depth, _ptr(out), _ptr(phi), int(homogeneous))
        return out

    def residual(self, phi, rhs, depth=0, homogeneous=True):
        phi, rhs = _f64(phi), _f64(rhs)
        out = np.zeros_like(phi)
        self.L.orc_residual(self.h, depth, _ptr(out), _ptr(phi), _ptr(rhs), int(homogeneous))
        return out

    def relax(self, phi, rhs, iters=1, depth=0):
        phi = _f64(phi).copy()
        rhs = _f64(rhs)
        self.L.orc_relax(self.h, depth, _ptr(phi), _ptr(rhs), int(iters))
        return phi

    def precond(self, rhs, depth=0):
        rhs = _f64(rhs)
        out = np.zeros_like(rhs)
        self.L.orc_precond(self.h, depth, _ptr(out), _ptr(rhs))
        return out

    def restrict_residual(self, phi, rhs, depth=0):
        phi, rhs = _f64(phi), _f64(rhs)
        out = np.zeros(self.shape(depth + 1))
        self.L.orc_restrict_residual(self.h, depth, _ptr(out), _ptr(phi), _ptr(rhs))
        return out

    def prolong_increment(self, phi, crse, depth=0):
        phi = _f64(phi).copy()
        crse = _f64(crse)
        self.L.orc_prolong_increment(self.h, depth, _ptr(phi), _ptr(crse))
        return phi

    def norm(self, x, ord=0, depth=0):
        x = _f64(x)
        return float(self.L.orc_norm(self.h, depth, _ptr(x), int(ord)))

    def cycle(self, e, r, depth=0):
        e = _f64(e).copy()
        r = _f64(r)
        self.L.orc_cycle(self.h, depth, _ptr(e), _ptr(r))
        return e

    def bottom_solve(self, e, r, metric=0.0, eps=1e-6):
        e = _f64(e).copy()
        r = _f64(r)
        self.L.orc_bottom_solve.restype = C.c_int
        self.L.orc_bottom_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double]
        it = self.L.orc_bottom_solve(self.h, _ptr(e), _ptr(r), float(metric), float(eps))
        return e, it

    def set_convergence_metric(self, metric: float):
        self.L.orc_set_convergence_metric.argtypes = [C.c_void_p, C.c_double]
        self.L.orc_set_convergence_metric(self.h, float(metric))

    def blew_up(self) -> bool:
        """did the last solve end in the reference's "solver blew up" abort (MappedAMRMultiGrid.H:1140-1144)?"""
        self.L.orc_blew_up.argtypes = [C.c_void_p]
        return bool(self.L.orc_blew_up(self.h))

    def solve(self, rhs, phi0=None, zero_phi=True, force_homogeneous=True):
        rhs = _f64(rhs)
        phi = np.zeros_like(rhs) if phi0 is None else _f64(phi0).copy()
        st = OrcStats()
        self.L.orc_solve(self.h, _ptr(phi), _ptr(rhs), int(zero_phi), int(force_homogeneous), C.byref(st))
        return phi, _stats(st)

    # -- projector calculus (face arrays: +1 point in their own direction) ---------------------
    def face_shapes(self, depth=0):
        n = self.depth_info(depth)["n"]
        return [(n[2] + (d == 2), n[1] + (d == 1), n[0] + (d == 0)) for d in range(self.dim)]

    def mac_gradient(self, phi, depth=0, homogeneous=False):
        phi = _f64(phi)
        g = [np.zeros(sh) for sh in self.face_shapes(depth)] + [None] * (3 - self.dim)
        self.L.orc_mac_gradient(self.h, depth, *[(_ptr(x) if x is not None else None) for x in g], _ptr(phi),
                                int(homogeneous))
        return g[:self.dim]

    def mac_divergence(self, flux, depth=0):
        f = [_f64(x) for x in flux] + [None] * (3 - self.dim)
        out = np.zeros(self.shape(depth))
        self.L.orc_mac_divergence(self.h, depth, _ptr(out), *[(_ptr(x) if x is not None else None) for x in f])
        return out

    def project(self, vel, dt, pressure0=None, zero_pressure=True, force_homogeneous=False):
        v = [_f64(x).copy() for x in vel] + [None] * (3 - self.dim)
        p = np.zeros(self.shape(0)) if pressure0 is None else _f64(pressure0).copy()
        st = OrcStats()
        self.L.orc_project.argtypes = [C.c_void_p] * 5 + [C.c_double, C.c_int, C.c_int, C.POINTER(OrcStats)]
        self.L.orc_project(self.h, *[(_ptr(x) if x is not None else None) for x in v], _ptr(p), float(dt),
                           int(zero_pressure), int(force_homogeneous), C.byref(st))
        return v[:self.dim], p, _stats(st)

    def mac_gradient_bc(self, phi, grad_bc, depth=0, homogeneous=False):
        """grad_bc 0: the operator's BC holder; 1: order-2 extrapolated ghosts (the reference projectors' m_gradBC)"""
        phi = _f64(phi)
        g = [np.zeros(sh) for sh in self.face_shapes(depth)] + [None] * (3 - self.dim)
        self.L.orc_mac_gradient_bc.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_int, C.c_int]
        self.L.orc_mac_gradient_bc(self.h, depth, *[(_ptr(x) if x is not None else None) for x in g], _ptr(phi),
                                   int(homogeneous), int(grad_bc))
        return g[:self.dim]

    # -- cell-centred projector (LevelCCProjector) ------------------------------------------------
    def cc_divergence(self, vel):
        u = [_f64(x) for x in vel] + [None] * (3 - self.dim)
        out = np.zeros(self.shape(0))
        self.L.orc_cc_divergence.argtypes = [C.c_void_p] * 5
        self.L.orc_cc_divergence(self.h, _ptr(out), *[(_ptr(x) if x is not None else None) for x in u])
        return out

    def cc_gradient(self, phi):
        phi = _f64(phi)
        g = [np.zeros(self.shape(0)) for _ in range(self.dim)] + [None] * (3 - self.dim)
        self.L.orc_cc_gradient.argtypes = [C.c_void_p] * 5
        self.L.orc_cc_gradient(self.h, *[(_ptr(x) if x is not None else None) for x in g], _ptr(phi))
        return g[:self.dim]

    def project_cc(self, vel, dt, J=None, pressure0=None, zero_pressure=True, force_homogeneous=False):
        v = [_f64(x).copy() for x in vel] + [None] * (3 - self.dim)
        p = np.zeros(self.shape(0)) if pressure0 is None else _f64(pressure0).copy()
        Jp = None if J is None else _f64(J)
        st = OrcStats()
        self.L.orc_project_cc.argtypes = [C.c_void_p] * 6 + [C.c_double, C.c_int, C.c_int, C.POINTER(OrcStats)]
        self.L.orc_project_cc(self.h, *[(_ptr(x) if x is not None else None) for x in v],
                              None if Jp is None else _ptr(Jp), _ptr(p), float(dt), int(zero_pressure),
                              int(force_homogeneous), C.byref(st))
        return v[:self.dim], p, _stats(st)

    # -- leptic solver pieces with per-box ghosts (either metric) ---------------------------------
    def _flat_shape(self):
        return (self.n[1], self.n[0]) if self.dim == 3 else (self.n[0],)

    def leptic_horiz_rhs(self, phi):
        """LevelLepticSolver::computeHorizRHS: -d_m avg_z(J g^{mj} d_j phi), boundary faces zero"""
        phi = _f64(phi)
        out = np.zeros(self._flat_shape())
        self.L.orc_leptic_horiz_rhs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self.L.orc_leptic_horiz_rhs(self.h, _ptr(out), _ptr(phi))
        return out

    def leptic_vert_horiz_grad(self, phi, scale):
        """LevelLepticSolver::levelVertHorizGradient: scale * J g^{zm} d_m phi on the lower / upper boundary faces"""
        phi = _f64(phi)
        lo, hi = np.zeros(self._flat_shape()), np.zeros(self._flat_shape())
        self.L.orc_leptic_vert_horiz_grad.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double]
        self.L.orc_leptic_vert_horiz_grad(self.h, _ptr(lo), _ptr(hi), _ptr(phi), float(scale))
        return lo, hi

    def timed_solve(self, rhs, ncycles_cap=0):
        rhs = _f64(rhs)
        st = OrcStats()
        self.L.orc_timed_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(OrcStats)]
        secs = self.L.orc_timed_solve(self.h, _ptr(rhs), int(ncycles_cap), C.byref(st))
        return float(secs), _stats(st)


def _stats(st: OrcStats) -> dict:
    return {"iters": st.iters, "exit_status": st.exitStatus, "initial_rnorm": st.initial_rnorm,
            "final_rnorm": st.final_rnorm, "bottom_iters_total": st.bottom_iters_total,
            "history": [st.history[i] for i in range(st.history_len)]}


def set_threads(n: int):
    lib().orc_set_threads(int(n))


def max_threads() -> int:
    return int(lib().orc_max_threads())


# ---- leptic-solver sweeps (oracle/leptic.cpp): plain global arrays, no Oracle object needed --------------------
BCT_NEUM, BCT_DIRI, BCT_CF = 0, 1, 3


def leptic_vertical_solve(rhs, jgzz, dz, lo_bc, hi_bc, dz_crse=None, bdry_lo=None, bdry_hi=None):
    """returns (phi, rhs_after, info); rhs (nv, ..., nx), jgzz (nv+1, ..., nx), bdry_* flat (..., nx) or None"""
    L = lib()
    rhs = _f64(rhs).copy()
    jgzz = _f64(jgzz)
    nv = rhs.shape[0]
    nh = rhs.size // nv
    phi = np.zeros_like(rhs)
    bl = _f64(bdry_lo) if bdry_lo is not None else None
    bh = _f64(bdry_hi) if bdry_hi is not None else None
    L.orc_leptic_vertical_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_long, C.c_int, C.c_double, C.c_double,
                                            C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    info = L.orc_leptic_vertical_solve(_ptr(phi), _ptr(rhs), _ptr(jgzz), nh, nv, float(dz),
                                       float(dz if dz_crse is None else dz_crse), int(lo_bc), int(hi_bc),
                                       _ptr(bl) if bl is not None else None, _ptr(bh) if bh is not None else None)
    return phi, rhs, int(info)


def leptic_vert_integral(flat, func, dz):
    L = lib()
    flat = _f64(flat).copy()
    func = _f64(func)
    nv = func.shape[0]
    L.orc_leptic_vert_integral.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_int, C.c_double]
    L.orc_leptic_vert_integral(_ptr(flat), _ptr(func), func.size // nv, nv, float(dz))
    return flat


def leptic_add_extrusion(dest, flat):
    L = lib()
    dest = _f64(dest).copy()
    flat = _f64(flat)
    nv = dest.shape[0]
    L.orc_leptic_add_extrusion.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_int]
    L.orc_leptic_add_extrusion(_ptr(dest), _ptr(flat), dest.size // nv, nv)
    return dest


def leptic_accum_div(rhs, flux, dir, dx_scale):
    """flat (ny, nx) rhs; flux has +1 point in direction dir"""
    L = lib()
    rhs = _f64(rhs).copy()
    flux = _f64(flux)
    ny, nx = (rhs.shape if rhs.ndim == 2 else (1, rhs.shape[0]))
    L.orc_leptic_accum_div.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double]
    L.orc_leptic_accum_div(_ptr(rhs), _ptr(flux), nx, ny, int(dir), float(dx_scale))
    return rhs


def leptic_vertical_excess(rhs, bdry_lo, bdry_hi, dz):
    L = lib()
    rhs = _f64(rhs)
    nv = rhs.shape[0]
    out = np.zeros(rhs.shape[1:])
    L.orc_leptic_vertical_excess.argtypes = [C.c_void_p] * 4 + [C.c_long, C.c_int, C.c_double]
    L.orc_leptic_vertical_excess(_ptr(out), _ptr(rhs), _ptr(_f64(bdry_lo)), _ptr(_f64(bdry_hi)), rhs.size // nv, nv, float(dz))
    return out

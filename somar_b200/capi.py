"""ctypes binding of libsomar_b200.so (the C ABI in include/somar_b200.h) for the Python test and bench drivers.

This is plumbing: every call goes straight to the CUDA library; there is no Python or CPU implementation of any
operator here, and loading fails loudly when the library has not been built (`python -c "import __graft_entry__ as g;
g.build()"` or `make -C somar_b200/csrc`).
"""
from __future__ import annotations

import ctypes as C
import os
import weakref
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SOMAR_B200_LIB", os.path.join(_HERE, "libsomar_b200.so"))   # override: A/B builds only

BC_NEUM, BC_DIRI, BC_PERIODIC = 0, 1, 2
RELAX_JACOBI, RELAX_LEVEL_GSRB, RELAX_LOOSE_GSRB, RELAX_LINE_GSRB = 0, 1, 2, 3

# every symbol include/somar_b200.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "somar_last_error", "somar_version", "somar_default_params", "somar_device_count", "somar_comm_unique_id",
    "somar_level_create", "somar_level_destroy", "somar_level_set_metric_fab", "somar_level_set_metric_device",
    "somar_level_finalize", "somar_level_set_alpha_beta", "somar_level_set_convergence_metric",
    "somar_level_set_solver_params", "somar_level_num_depths", "somar_level_depth_info",
    "somar_level_sync", "somar_level_stream", "somar_level_launch_count", "somar_level_get_metric",
    "somar_level_profile_enable", "somar_level_profile_read", "somar_level_relax_kernel", "somar_level_halo_path",
    "somar_field_create", "somar_field_destroy", "somar_field_set_fab", "somar_field_get_fab", "somar_field_set",
    "somar_field_get", "somar_apply_op", "somar_residual", "somar_precond", "somar_set_to_zero", "somar_assign",
    "somar_incr", "somar_axby", "somar_scale", "somar_norm", "somar_local_max_norm", "somar_dot", "somar_relax",
    "somar_restrict_residual", "somar_prolong_increment", "somar_mg_cycle", "somar_solve",
    "somar_field_set_face", "somar_field_get_face", "somar_mac_gradient", "somar_mac_divergence",
    "somar_mac_apply_correction", "somar_project", "somar_leptic_vertical_solve", "somar_leptic_vert_integral",
    "somar_leptic_vertical_excess", "somar_leptic_add_extrusion", "somar_leptic_accum_div",
    "somar_advect_predict_scalar", "somar_advect_update",
    "somar_leptic_default_params", "somar_leptic_create", "somar_leptic_destroy", "somar_leptic_solve",
    "somar_leptic_level", "somar_norm_noalloc",
    "somar_level_set_projector_bc", "somar_mac_gradient_bc", "somar_mac_wall_flux_bc", "somar_cc_divergence",
    "somar_cc_gradient", "somar_project_cc",
]


class LevelDesc(C.Structure):
    _fields_ = [("dim", C.c_int), ("n", C.c_int * 3), ("periodic", C.c_int * 3), ("nboxes", C.c_int),
                ("box_lo", C.POINTER(C.c_int)), ("box_hi", C.POINTER(C.c_int)), ("box_rank", C.POINTER(C.c_int)),
                ("dx", C.c_double * 3), ("alpha", C.c_double), ("beta", C.c_double), ("is_diagonal", C.c_int),
                ("bc_type", C.c_int * 6), ("bc_value", C.c_double * 6), ("rank", C.c_int), ("nranks", C.c_int)]


class SolverParams(C.Structure):
    _fields_ = [("eps", C.c_double), ("num_smooth_down", C.c_int), ("num_smooth_up", C.c_int),
                ("num_smooth_bottom", C.c_int), ("num_smooth_precond", C.c_int), ("numMG", C.c_int), ("imin", C.c_int),
                ("imax", C.c_int), ("hang", C.c_double), ("normThresh", C.c_double), ("maxDepth", C.c_int),
                ("relaxMode", C.c_int), ("precondMode", C.c_int), ("bottom_eps", C.c_double), ("bottom_reps", C.c_double),
                ("bottom_hang", C.c_double), ("bottom_small", C.c_double), ("bottom_imax", C.c_int),
                ("bottom_numRestarts", C.c_int), ("bottom_normType", C.c_int), ("verbosity", C.c_int)]


class SolveStats(C.Structure):
    _fields_ = [("iters", C.c_int), ("exitStatus", C.c_int), ("initial_rnorm", C.c_double), ("final_rnorm", C.c_double),
                ("bottom_iters_total", C.c_int), ("history_len", C.c_int), ("history", C.c_double * 128),
                ("solve_ms", C.c_double), ("launches", C.c_longlong)]


class AdvectParams(C.Structure):
    _fields_ = [("normal_pred_order", C.c_int), ("use_limiting", C.c_int), ("use_high_order_limiter", C.c_int),
                ("use_upwinding", C.c_int), ("return_flux", C.c_int), ("use_fourth_order_slopes", C.c_int)]


class LepticParams(C.Structure):
    _fields_ = [("max_order", C.c_int), ("eps", C.c_double), ("hang", C.c_double), ("norm_type", C.c_int),
                ("horiz_rhs_tol", C.c_double), ("horiz", SolverParams), ("full", SolverParams)]


class LepticStats(C.Structure):
    _fields_ = [("orders", C.c_int), ("exit_status", C.c_int), ("horiz_solves", C.c_int), ("krylov_used", C.c_int),
                ("horiz_iters_total", C.c_int), ("num_res_norms", C.c_int), ("res_norms", C.c_double * 16),
                ("solve_ms", C.c_double), ("launches", C.c_longlong)]


class SomarError(RuntimeError):
    pass


_LIB = None


def _preload_bundled_nccl():
    """One NCCL per process.  libsomar_b200.so needs libnccl.so.2; a process that also imports PyTorch gets PyTorch's
    bundled copy (newer than the system one), and whichever is loaded first serves both.  Loading the system copy
    first breaks `import torch` later (missing newer symbols), so take the bundled one when it exists."""
    import sys
    for d in sys.path:
        cand = os.path.join(d, "nvidia", "nccl", "lib", "libnccl.so.2")
        if os.path.exists(cand):
            try:
                C.CDLL(cand, mode=C.RTLD_GLOBAL)
            except OSError:
                pass
            return


def lib():
    """Loads libsomar_b200.so; raises (never falls back) when it is missing."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise SomarError(f"{LIB_PATH} is not built: run __graft_entry__.build() (there is no CPU fallback)")
        _preload_bundled_nccl()
        L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        L.somar_last_error.restype = C.c_char_p
        L.somar_version.restype = C.c_char_p
        _LIB = L
    return _LIB


def _ck(rc: int):
    if rc != 0:
        raise SomarError(f"somar_b200 error {rc}: {lib().somar_last_error().decode()}")


def default_params(**kw) -> SolverParams:
    p = SolverParams()
    _ck(lib().somar_default_params(C.byref(p)))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise KeyError(k)
        setattr(p, k, v)
    return p


def leptic_default_params(horiz=None, full=None, **kw) -> LepticParams:
    """LevelLepticSolver::setDefaultParameters; `horiz` / `full` are dicts of SolverParams overrides"""
    p = LepticParams()
    _ck(lib().somar_leptic_default_params(C.byref(p)))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise KeyError(k)
        setattr(p, k, v)
    for sub, d in ((p.horiz, horiz), (p.full, full)):
        for k, v in (d or {}).items():
            if not hasattr(sub, k):
                raise KeyError(k)
            setattr(sub, k, v)
    return p


class LepticSolver:
    """LevelLepticSolver (define + solve) on a finalized Level: somar_leptic_create / _solve / _destroy"""

    def __init__(self, level: "Level", params: Optional[LepticParams] = None):
        self.level = level
        self.params = params if params is not None else leptic_default_params()
        self.h = None
        h = C.c_void_p()
        _ck(lib().somar_leptic_create(level.h, C.byref(self.params), C.byref(h)))
        self.h = h
        level._leptics.add(self)       # destroyed before the level: the derived levels borrow its communicator

    def solve(self, phi: "Field", rhs: "Field", homogeneous=False) -> dict:
        st = LepticStats()
        _ck(lib().somar_leptic_solve(self.h, phi.h, rhs.h, int(homogeneous), C.byref(st)))
        return {"orders": st.orders, "exit_status": st.exit_status, "horiz_solves": st.horiz_solves,
                "krylov_used": bool(st.krylov_used), "horiz_iters_total": st.horiz_iters_total,
                "res_norms": [st.res_norms[i] for i in range(st.num_res_norms)], "solve_ms": st.solve_ms,
                "launches": st.launches}

    def horiz_metric(self, full=False):
        """(J g^00, J g^11) of the flattened level, as (ny, nx+1) and (ny+1, nx) arrays (None without one); full=True:
        both components per face direction"""
        hv = C.c_void_p()
        _ck(lib().somar_leptic_level(self.h, 1, C.byref(hv)))
        if not hv.value:
            return None
        nx, ny = self.level.n[0], self.level.n[1]
        j0 = np.zeros((2, ny, nx + 1))
        j1 = np.zeros((2, ny + 1, nx))
        _ck(lib().somar_level_get_metric(hv, 0, j0.ctypes.data_as(C.c_void_p), j1.ctypes.data_as(C.c_void_p), None, None,
                                         None))
        if full:
            return j0, j1           # j0[b] = J g^{0b} on x faces, j1[b] = J g^{1b} on y faces
        return j0[0], j1[1]

    def destroy(self):
        if getattr(self, "h", None):
            lib().somar_leptic_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


NOALLOC_MAX_N = 67108864


def norm_noalloc(x, ord: int, out, stream: int = 0):
    """somar_norm_noalloc: x, out are contiguous float64 CUDA tensors (n values, 1 value); nothing is allocated and
    nothing synchronised -- the result is in `out` once `stream` (a cudaStream_t as int, 0 = default) has run."""
    assert x.is_cuda and out.is_cuda and x.is_contiguous() and str(x.dtype) == "torch.float64"
    _ck(lib().somar_norm_noalloc(C.c_void_p(x.data_ptr()), C.c_longlong(x.numel()), int(ord),
                                 C.c_void_p(out.data_ptr()), C.c_void_p(stream)))


def comm_unique_id() -> bytes:
    buf = (C.c_ubyte * 128)()
    _ck(lib().somar_comm_unique_id(buf))
    return bytes(buf)


def _i3(v):
    return (C.c_int * 3)(*[int(x) for x in v])


def _ptr(a):
    """numpy array, torch tensor (cpu or cuda) or None -> (void*, on_device)"""
    if a is None:
        return None, 0
    if hasattr(a, "data_ptr"):
        assert a.is_contiguous() and str(a.dtype) == "torch.float64"
        if a.is_cuda:
            # the library works on its own (non-blocking) stream: whatever produced this tensor on torch's current
            # stream must have finished before the library reads it (and before it overwrites it)
            import torch
            torch.cuda.current_stream(a.device).synchronize()
        return C.c_void_p(a.data_ptr()), int(a.is_cuda)
    assert isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return C.c_void_p(a.ctypes.data), 0


class Field:
    def __init__(self, level: "Level", depth: int):
        self.level, self.depth = level, depth
        h = C.c_void_p()
        _ck(lib().somar_field_create(level.h, depth, C.byref(h)))
        self.h = h
        level._fields.add(self)

    def shape(self):
        return self.level.local_shape(self.depth)

    def set(self, data):
        p, dev = _ptr(data)
        assert tuple(data.shape) == self.shape(), (tuple(data.shape), self.shape())
        _ck(lib().somar_field_set(self.h, p, dev))
        if hasattr(data, "data_ptr"):
            # the copy is only ENQUEUED on the library's stream; torch's caching allocator does not know that stream,
            # so a torch source (device, or pinned host) must not be recycled before the copy has run
            self.level.sync()
        return self

    def get(self, out=None):
        if out is None:
            out = np.empty(self.shape(), dtype=np.float64)
        p, dev = _ptr(out)
        _ck(lib().somar_field_get(self.h, p, dev))
        if dev:
            self.level.sync()
        return out

    def set_face(self, data, dir: int):
        """dense rank-local face array of direction dir: shape = cell shape with +1 point in that direction"""
        p, dev = _ptr(data)
        _ck(lib().somar_field_set_face(self.h, p, int(dir), dev))
        if hasattr(data, "data_ptr"):
            self.level.sync()
        return self

    def get_face(self, dir: int, out=None):
        if out is None:
            sh = list(self.shape())
            sh[2 - dir] += 1
            out = np.empty(sh, dtype=np.float64)
        p, dev = _ptr(out)
        _ck(lib().somar_field_get_face(self.h, p, int(dir), dev))
        if dev:
            self.level.sync()
        return out

    def set_fab(self, fab: np.ndarray, lo, hi, vlo, vhi):
        p, _ = _ptr(fab)
        _ck(lib().somar_field_set_fab(self.h, p, _i3(lo), _i3(hi), _i3(vlo), _i3(vhi)))

    def get_fab(self, fab: np.ndarray, lo, hi, vlo, vhi):
        p, _ = _ptr(fab)
        _ck(lib().somar_field_get_fab(self.h, p, _i3(lo), _i3(hi), _i3(vlo), _i3(vhi)))

    def destroy(self):
        # a field must die before its level (the C object points back at it); Level.destroy() sees to that
        if self.h and getattr(self.level, "h", None):
            lib().somar_field_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


class Level:
    """One AMR level on one GPU (rank): metric + multigrid hierarchy resident in HBM."""

    def __init__(self, n: Sequence[int], dx: Sequence[float], *, dim=3, periodic=(0, 0, 0), boxes=None, box_rank=None,
                 alpha=0.0, beta=1.0, is_diagonal=True, bc_type=None, bc_value=None,
                 params: Optional[SolverParams] = None, device=0, rank=0, nranks=1, comm_id: Optional[bytes] = None):
        L = lib()
        self.h = None
        self._fields = weakref.WeakSet()
        self._leptics = weakref.WeakSet()
        self.n = tuple(int(x) for x in n)
        self.dim = dim
        if boxes is None:
            boxes = [((0, 0, 0), tuple(x - 1 for x in self.n))]
        self.boxes = boxes
        lo = np.ascontiguousarray([b[0] for b in boxes], dtype=np.int32)
        hi = np.ascontiguousarray([b[1] for b in boxes], dtype=np.int32)
        rk = np.ascontiguousarray(box_rank if box_rank is not None else [0] * len(boxes), dtype=np.int32)
        self.box_rank = [int(x) for x in rk]
        d = LevelDesc()
        d.dim = dim
        d.n[:] = self.n
        d.periodic[:] = [int(bool(x)) for x in periodic]
        d.nboxes = len(boxes)
        d.box_lo = lo.ctypes.data_as(C.POINTER(C.c_int))
        d.box_hi = hi.ctypes.data_as(C.POINTER(C.c_int))
        d.box_rank = rk.ctypes.data_as(C.POINTER(C.c_int))
        d.dx[:] = [float(x) for x in dx]
        d.alpha, d.beta = float(alpha), float(beta)
        d.is_diagonal = int(bool(is_diagonal))
        d.bc_type[:] = [int(x) for x in (bc_type if bc_type is not None else [BC_NEUM] * 6)]
        d.bc_value[:] = [float(x) for x in (bc_value if bc_value is not None else [0.0] * 6)]
        d.rank, d.nranks = int(rank), int(nranks)
        self.rank, self.nranks = int(rank), int(nranks)
        self.params = params if params is not None else default_params()
        h = C.c_void_p()
        cid = (C.c_ubyte * 128).from_buffer_copy(comm_id) if comm_id is not None else None
        _ck(L.somar_level_create(C.byref(d), C.byref(self.params), int(device), cid, C.byref(h)))
        self.h = h
        nd_ = C.c_int()
        _ck(L.somar_level_num_depths(self.h, C.byref(nd_)))
        self.ndepths = nd_.value

    # -- set-up ---------------------------------------------------------------------------------------------
    def set_metric_fab(self, lo, hi, jgup: Sequence[Optional[np.ndarray]], jinv: Optional[np.ndarray], ncomp: int):
        ps = [_ptr(j)[0] for j in list(jgup) + [None] * (3 - len(jgup))]
        _ck(lib().somar_level_set_metric_fab(self.h, _i3(lo), _i3(hi), int(ncomp), ps[0], ps[1], ps[2], _ptr(jinv)[0]))

    def set_metric_global(self, jgup: Sequence[np.ndarray], jinv: np.ndarray):
        """Global (whole-domain) numpy metric arrays -> this rank's boxes, box by box (the CHF_FRA path)."""
        for b, ((lo, hi), r) in enumerate(zip(self.boxes, self.box_rank)):
            if r != self.rank:
                continue
            sl = tuple(slice(lo[d], hi[d] + 1) for d in (2, 1, 0))
            jg = []
            ncomp = 1
            for d in range(self.dim):
                fs = list(sl)
                fs[2 - d] = slice(lo[d], hi[d] + 2)
                a = jgup[d]
                if a.ndim == 4:
                    ncomp = a.shape[0]
                    jg.append(np.ascontiguousarray(a[(slice(None),) + tuple(fs)]))
                else:
                    jg.append(np.ascontiguousarray(a[tuple(fs)]))
            self.set_metric_fab(lo, hi, jg, np.ascontiguousarray(jinv[sl]), ncomp)

    def set_metric_device(self, jgup, jinv, ncomp: int):
        """torch CUDA tensors over the rank-local region (faces: +1 point in their direction)."""
        ps = [_ptr(j)[0] for j in list(jgup) + [None] * (3 - len(jgup))]
        _ck(lib().somar_level_set_metric_device(self.h, int(ncomp), ps[0], ps[1], ps[2], _ptr(jinv)[0]))
        self.sync()   # the sources are torch tensors the caller may free right away (see Field.set)

    def finalize(self):
        _ck(lib().somar_level_finalize(self.h))
        return self

    def set_alpha_beta(self, alpha, beta):
        _ck(lib().somar_level_set_alpha_beta(self.h, C.c_double(alpha), C.c_double(beta)))

    def set_convergence_metric(self, metric: float):
        _ck(lib().somar_level_set_convergence_metric(self.h, C.c_double(metric)))

    def set_params(self, params: "SolverParams"):
        """AMRPressureSolver::setAMRMGParameters / setBottomParameters on an existing level (not maxDepth)."""
        _ck(lib().somar_level_set_solver_params(self.h, C.byref(params)))
        self.params = params

    def depth_info(self, depth: int):
        n, r, lo, hi = (C.c_int * 3)(), (C.c_int * 3)(), (C.c_int * 3)(), (C.c_int * 3)()
        dx = (C.c_double * 3)()
        z = C.c_int()
        _ck(lib().somar_level_depth_info(self.h, depth, n, r, dx, C.byref(z), lo, hi))
        return {"n": tuple(n), "ratio": tuple(r), "dx": tuple(dx), "zero_avg": bool(z.value), "local_lo": tuple(lo),
                "local_hi": tuple(hi)}

    def local_shape(self, depth: int):
        i = self.depth_info(depth)
        return tuple(i["local_hi"][d] - i["local_lo"][d] + 1 for d in (2, 1, 0))

    def metric(self, depth: int):
        nz, ny, nx = self.local_shape(depth)
        nd = self.dim
        jg = [np.zeros((nd, nz + (d == 2), ny + (d == 1), nx + (d == 0))) if d < nd else None for d in range(3)]
        jinv, lap = np.zeros((nz, ny, nx)), np.zeros((nz, ny, nx))
        _ck(lib().somar_level_get_metric(self.h, depth, _ptr(jg[0])[0], _ptr(jg[1])[0], _ptr(jg[2])[0], _ptr(jinv)[0],
                                         _ptr(lap)[0]))
        return jg, jinv, lap

    def sync(self):
        _ck(lib().somar_level_sync(self.h))

    def stream(self) -> int:
        v = C.c_void_p()
        _ck(lib().somar_level_stream(self.h, C.byref(v)))
        return v.value or 0

    def profile(self, on: bool):
        _ck(lib().somar_level_profile_enable(self.h, int(on)))

    def profile_read(self):
        cnt, ms = (C.c_longlong * 6)(), (C.c_double * 6)()
        _ck(lib().somar_level_profile_read(self.h, cnt, ms))
        names = ["smooth0_kernel", "smooth_coarse", "restrict", "prolong", "bottom", "outer"]
        return {n: {"count": cnt[i], "ms": ms[i]} for i, n in enumerate(names)}

    RELAX_KERNELS = {0: "k_gsrb_ortho", 1: "k_gsrb_fused", 2: "k_gsrb_full", 3: "k_line_gsrb", 4: "k_jacobi", 5: "k_gsrb_async", 6: "k_gsrb_loose", 7: "k_gsrb_tma", 8: "k_line_pass_c", -1: "none"}

    def relax_kernel(self, depth=0, mode=RELAX_LEVEL_GSRB) -> str:
        v = C.c_int()
        _ck(lib().somar_level_relax_kernel(self.h, depth, mode, C.byref(v)))
        return self.RELAX_KERNELS[v.value]

    def halo_path(self) -> str:
        v = C.c_int()
        _ck(lib().somar_level_halo_path(self.h, C.byref(v)))
        return {0: "none", 1: "nccl", 2: "p2p"}[v.value]

    def launches(self) -> int:
        v = C.c_longlong()
        _ck(lib().somar_level_launch_count(self.h, C.byref(v)))
        return v.value

    def field(self, depth=0, data=None) -> Field:
        f = Field(self, depth)
        if data is not None:
            f.set(data)
        return f

    # -- operator -------------------------------------------------------------------------------------------
    def apply_op(self, lhs: Field, phi: Field, homogeneous=True):
        _ck(lib().somar_apply_op(self.h, phi.depth, lhs.h, phi.h, int(homogeneous)))

    def residual(self, res: Field, phi: Field, rhs: Field, homogeneous=True):
        _ck(lib().somar_residual(self.h, phi.depth, res.h, phi.h, rhs.h, int(homogeneous)))

    def relax(self, phi: Field, rhs: Field, iters=1):
        _ck(lib().somar_relax(self.h, phi.depth, phi.h, rhs.h, int(iters)))

    def precond(self, cor: Field, res: Field):
        _ck(lib().somar_precond(self.h, res.depth, cor.h, res.h))

    def restrict_residual(self, crse: Field, phi: Field, rhs: Field):
        _ck(lib().somar_restrict_residual(self.h, phi.depth, crse.h, phi.h, rhs.h))

    def prolong_increment(self, phi: Field, crse: Field):
        _ck(lib().somar_prolong_increment(self.h, phi.depth, phi.h, crse.h))

    def set_to_zero(self, x: Field):
        _ck(lib().somar_set_to_zero(self.h, x.h))

    def assign(self, y: Field, x: Field):
        _ck(lib().somar_assign(self.h, y.h, x.h))

    def incr(self, y: Field, x: Field, scale: float):
        _ck(lib().somar_incr(self.h, y.h, x.h, C.c_double(scale)))

    def axby(self, lhs: Field, x: Field, y: Field, a: float, b: float):
        _ck(lib().somar_axby(self.h, lhs.h, x.h, y.h, C.c_double(a), C.c_double(b)))

    def scale(self, x: Field, s: float):
        _ck(lib().somar_scale(self.h, x.h, C.c_double(s)))

    def norm(self, x: Field, ord=0) -> float:
        v = C.c_double()
        _ck(lib().somar_norm(self.h, x.h, int(ord), C.byref(v)))
        return v.value

    def dot(self, a: Field, b: Field) -> float:
        v = C.c_double()
        _ck(lib().somar_dot(self.h, a.h, b.h, C.byref(v)))
        return v.value

    def mg_cycle(self, e: Field, r: Field):
        _ck(lib().somar_mg_cycle(self.h, e.depth, e.h, r.h))

    def solve(self, phi: Field, rhs: Field, zero_phi=True, force_homogeneous=True) -> dict:
        st = SolveStats()
        _ck(lib().somar_solve(self.h, phi.h, rhs.h, int(zero_phi), int(force_homogeneous), C.byref(st)))
        return {"iters": st.iters, "exit_status": st.exitStatus, "initial_rnorm": st.initial_rnorm,
                "final_rnorm": st.final_rnorm, "bottom_iters_total": st.bottom_iters_total,
                "history": [st.history[i] for i in range(st.history_len)], "solve_ms": st.solve_ms,
                "launches": st.launches}

    # -- projector calculus -----------------------------------------------------------------
    @staticmethod
    def _h3(fs):
        fs = list(fs) + [None] * (3 - len(fs))
        return [f.h if f is not None else None for f in fs]

    def mac_gradient(self, grad, phi: Field, homogeneous=False):
        _ck(lib().somar_mac_gradient(self.h, *self._h3(grad), phi.h, int(homogeneous)))

    def mac_divergence(self, div: Field, flux, divide_by=0.0):
        L = lib()
        L.somar_mac_divergence.argtypes = [C.c_void_p] * 5 + [C.c_double]
        _ck(L.somar_mac_divergence(self.h, div.h, *self._h3(flux), float(divide_by)))

    def mac_apply_correction(self, vel, phi: Field, dt: float):
        L = lib()
        L.somar_mac_apply_correction.argtypes = [C.c_void_p] * 5 + [C.c_double]
        _ck(L.somar_mac_apply_correction(self.h, *self._h3(vel), phi.h, float(dt)))

    def project(self, vel, pressure: Field, dt: float, zero_pressure=True, force_homogeneous=False) -> dict:
        st = SolveStats()
        L = lib()
        L.somar_project.argtypes = [C.c_void_p] * 5 + [C.c_double, C.c_int, C.c_int, C.POINTER(SolveStats)]
        _ck(L.somar_project(self.h, *self._h3(vel), pressure.h, float(dt), int(zero_pressure), int(force_homogeneous),
                            C.byref(st)))
        return {"iters": st.iters, "exit_status": st.exitStatus, "initial_rnorm": st.initial_rnorm,
                "final_rnorm": st.final_rnorm, "bottom_iters_total": st.bottom_iters_total,
                "history": [st.history[i] for i in range(st.history_len)], "solve_ms": st.solve_ms,
                "launches": st.launches}

    @staticmethod
    def _stats(st):
        return {"iters": st.iters, "exit_status": st.exitStatus, "initial_rnorm": st.initial_rnorm,
                "final_rnorm": st.final_rnorm, "bottom_iters_total": st.bottom_iters_total,
                "history": [st.history[i] for i in range(st.history_len)], "solve_ms": st.solve_ms,
                "launches": st.launches}

    def set_projector_bc(self, mode: int):
        """0: the fluxes carry their wall values, gradient with the operator's BCs; 1: the reference projectors' own
        m_divBC / m_gradBC inside project()"""
        _ck(lib().somar_level_set_projector_bc(self.h, int(mode)))

    def mac_gradient_bc(self, grad, phi: Field, grad_bc: int, homogeneous=False):
        L = lib()
        L.somar_mac_gradient_bc.argtypes = [C.c_void_p] * 5 + [C.c_int, C.c_int]
        _ck(L.somar_mac_gradient_bc(self.h, *self._h3(grad), phi.h, int(homogeneous), int(grad_bc)))

    def mac_wall_flux_bc(self, flux):
        L = lib()
        L.somar_mac_wall_flux_bc.argtypes = [C.c_void_p] * 4
        _ck(L.somar_mac_wall_flux_bc(self.h, *self._h3(flux)))

    # -- cell-centred projector (LevelCCProjector) ---------------------------------------------------------------
    def cc_divergence(self, div: Field, vel, J: Field = None, divide_by=0.0):
        L = lib()
        L.somar_cc_divergence.argtypes = [C.c_void_p] * 6 + [C.c_double]
        _ck(L.somar_cc_divergence(self.h, div.h, *self._h3(vel), J.h if J is not None else None, float(divide_by)))

    def cc_gradient(self, grad, phi: Field):
        L = lib()
        L.somar_cc_gradient.argtypes = [C.c_void_p] * 5
        _ck(L.somar_cc_gradient(self.h, *self._h3(grad), phi.h))

    def project_cc(self, vel, pressure: Field, dt: float, J: Field = None, zero_pressure=True,
                   force_homogeneous=False) -> dict:
        st = SolveStats()
        L = lib()
        L.somar_project_cc.argtypes = [C.c_void_p] * 6 + [C.c_double, C.c_int, C.c_int, C.POINTER(SolveStats)]
        _ck(L.somar_project_cc(self.h, *self._h3(vel), J.h if J is not None else None, pressure.h, float(dt),
                               int(zero_pressure), int(force_homogeneous), C.byref(st)))
        return self._stats(st)

    # -- leptic-solver sweeps ---------------------------------------------------------------------------------
    def leptic_vertical_solve(self, phi: Field, rhs: Field, lo_bc: int, hi_bc: int, dz_crse=None, bdry_lo=None,
                              bdry_hi=None):
        L = lib()
        L.somar_leptic_vertical_solve.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_void_p]
        vd = self.dim - 1
        dzc = self.depth_info(0)["dx"][vd] if dz_crse is None else dz_crse
        _ck(L.somar_leptic_vertical_solve(self.h, phi.h, rhs.h, int(lo_bc), int(hi_bc), float(dzc),
                                          bdry_lo.h if bdry_lo is not None else None,
                                          bdry_hi.h if bdry_hi is not None else None))

    def leptic_vert_integral(self, flat: Field, func: Field, dz_scale: float):
        L = lib()
        L.somar_leptic_vert_integral.argtypes = [C.c_void_p] * 3 + [C.c_double]
        _ck(L.somar_leptic_vert_integral(self.h, flat.h, func.h, float(dz_scale)))

    def leptic_vertical_excess(self, excess: Field, rhs: Field, bdry_lo: Field, bdry_hi: Field):
        _ck(lib().somar_leptic_vertical_excess(self.h, excess.h, rhs.h, bdry_lo.h, bdry_hi.h))

    def leptic_add_extrusion(self, dest: Field, flat: Field):
        _ck(lib().somar_leptic_add_extrusion(self.h, dest.h, flat.h))

    def leptic_accum_div(self, flat_rhs: Field, flat_flux: Field, fc_dir: int, dx_scale: float):
        L = lib()
        L.somar_leptic_accum_div.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_double]
        _ck(L.somar_leptic_accum_div(self.h, flat_rhs.h, flat_flux.h, int(fc_dir), float(dx_scale)))

    # -- advection predictor --------------------------------------------------------------------------------
    def advect_predict_scalar(self, whalf, w_old: Field, old_vel, adv_vel, cc_j: Field, dt: float, source=None,
                              normal_pred_order=2, use_limiting=False, use_high_order_limiter=False, use_upwinding=True,
                              return_flux=True, use_fourth_order_slopes=False):
        prm = AdvectParams(int(normal_pred_order), int(use_limiting), int(use_high_order_limiter), int(use_upwinding),
                           int(return_flux), int(use_fourth_order_slopes))
        L = lib()
        L.somar_advect_predict_scalar.argtypes = [C.c_void_p] * 13 + [C.c_double, C.POINTER(AdvectParams)]
        _ck(L.somar_advect_predict_scalar(self.h, *self._h3(whalf), w_old.h, source.h if source is not None else None,
                                          *self._h3(old_vel), *self._h3(adv_vel), cc_j.h, float(dt), C.byref(prm)))

    def advect_update(self, w_new: Field, w_old: Field, flux, dt: float):
        L = lib()
        L.somar_advect_update.argtypes = [C.c_void_p] * 6 + [C.c_double]
        _ck(L.somar_advect_update(self.h, w_new.h, w_old.h, *self._h3(flux), float(dt)))

    def destroy(self):
        if getattr(self, "h", None):
            for s in list(self._leptics):
                s.destroy()
            for f in list(self._fields):
                f.destroy()
            lib().somar_level_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

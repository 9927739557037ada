"""oracle/leptic_ref.py -- TEST INFRASTRUCTURE ONLY (CPU oracle; never the product path).

The leptic ITERATION of SOMAR (SURVEY.md 8(a) "LevelLepticSolver" row, 8(f)-2) restated over the CPU oracle's operator
(oracle/somar_oracle.cpp through pyoracle.Oracle) and the sweeps of oracle/leptic.cpp, statement by statement after
  LevelLepticSolver::define             src/calculus/LepticSolver/LevelLepticSolver.cpp:147-437
      (vertical grids with Jinv = 1: :197-205; the 3-D multigrid of the fall-back: :251-300; the flattened horizontal
       level with vertically averaged J g^{ab}: :355-358 -> LevelGeometry::createVertAvgFCJgupPtr,
       src/geometry/LevelGeometryBasics.cpp:499-571; zero-Neumann horizontal BCs :363-368; m_horizRemoveAvg :410-433)
  LevelLepticSolver::setDefaultParameters   :463-514
  LevelLepticSolver::solve              :646-956
  ::computeHorizRHS                     :981-1097   (diagonal metric: MAPPEDMACGRADORTHO, DivCurlGradF.ChF:221-262)
  ::levelVertHorizGradient              :1107-1176  (diagonal metric: the boundary data stay zero, :1125)
  ::computeVerticalExcess / ::verticalLineSolver / ::horizontalSolver / ::addHorizontalCorrection / ::setZeroAvg
                                        :1183-1240, :1247-1425, :1427-1462, :1489-1515, :1665-1709
  ::gatherVerticalBCTypes               :1522-1644

Scope: one level, no coarser level (m_crsePhiPtr == NULL); diagonal metrics (the device driver's scope) and, through
Op::lepticHorizRHS / Op::lepticVertHorizGradient of oracle/somar_oracle.cpp, full metrics in 3-D (oracle only so far);
non-periodic vertical direction with Neumann or Dirichlet sides; when a horizontal solve is needed the problem must be
3-D with non-periodic horizontal directions (the reference leaves the domain-boundary faces of a periodic direction
unset in computeHorizRHS, :997-1001 + :1064-1069, BoundaryData.cpp:103-104).  The horizontal grids are the flattened
boxes of the level (the reference re-balances them with LepticBoxUtils::createHorizontalSolverGrids; the box layout
only enters through the multigrid schedule).

PARITY UNPINNED against reference output (the reference cannot be built here and ships no vectors); pinned by
tests/test_leptic_solver.py to what the algorithm promises: a separable problem is solved by the O(1) step, and the
residual norms fall order by order on a thin (leptic) domain.

Arrays are (nz, ny, nx) C-ordered; flat arrays (ny, nx)."""
from dataclasses import dataclass, field

import numpy as np

NEUM, DIRI = 0, 1
EXIT = {"NONE": -1, "CONVERGE": 0, "ITER": 1, "HANG": 2, "DIVERGE": 3, "KABOOM": 4}   # LevelLepticSolver.H ExitStatus


def horiz_mg_defaults():
    """setHorizMGParameters / setHorizBottomParameters defaults, :473-492, as pyoracle.default_params keywords"""
    return dict(imin=5, imax=20, num_smooth_down=4, num_smooth_bottom=4, num_smooth_up=4, num_smooth_precond=2,
                relaxMode=1, precondMode=1, maxDepth=-1, eps=1.0e-12, hang=1.0e-15, normThresh=1.0e-30, numMG=1,
                bottom_imax=80, bottom_eps=1.0e-12, bottom_numRestarts=5, bottom_hang=1.0e-15)


def full_mg_defaults():
    """setFullMGParameters / setFullBottomParameters defaults, :494-513 (the bottom solver's eps and hang are never
    handed over, :269-274: BiCGStab keeps its own 1e-6 / 1e-7)"""
    return dict(imin=5, imax=20, num_smooth_down=4, num_smooth_bottom=4, num_smooth_up=4, num_smooth_precond=4,
                relaxMode=3, precondMode=2, maxDepth=-1, eps=1.0e-6, hang=1.0e-15, normThresh=1.0e-30, numMG=1,
                bottom_imax=80, bottom_numRestarts=5)


@dataclass
class LepticParams:
    max_order: int = 4
    hang: float = 1.0e-15
    norm_type: int = 0
    horiz_rhs_tol: float = 1.0e-14
    horiz: dict = field(default_factory=horiz_mg_defaults)
    full: dict = field(default_factory=full_mg_defaults)


def vert_avg_faces(jg, nz):
    """createVertAvgFCJgupPtr: flat = 0; flat += J g^{aa}(k) * (1 / Nz), k ascending (UNMAPPEDVERTINTEGRAL)"""
    scale = 1.0 / float(nz)
    flat = np.zeros(jg.shape[1:])
    for k in range(jg.shape[0]):
        flat = flat + jg[k] * scale
    return flat


def horiz_rhs(phi, jgup, dx, nz):
    """computeHorizRHS for a diagonal metric: -d_m avg_z(J g^{mm} d_m phi) on the flat grid, boundary faces zero"""
    ny, nx = phi.shape[1], phi.shape[2]
    rhs = np.zeros((ny, nx))
    dz_scale = 1.0 / float(nz)
    for d in range(2):
        dxinv = 1.0 / dx[d]
        fshape = (ny + (d == 1), nx + (d == 0))
        avg = np.zeros(fshape)
        for k in range(phi.shape[0]):
            g = np.zeros(fshape)
            if d == 0:
                g[:, 1:nx] = dxinv * jgup[0][k][:, 1:nx] * (phi[k][:, 1:] - phi[k][:, :-1])
            else:
                g[1:ny, :] = dxinv * jgup[1][k][1:ny, :] * (phi[k][1:, :] - phi[k][:-1, :])
            avg = avg + g * dz_scale
        dx_scale = -1.0 / dx[d]
        if d == 0:
            rhs = rhs + (avg[:, 1:] - avg[:, :-1]) * dx_scale
        else:
            rhs = rhs + (avg[1:, :] - avg[:-1, :]) * dx_scale
    return rhs


class LevelLepticSolver:
    """define(op) + solve(phi, rhs) of the reference class, over the CPU oracle"""

    def __init__(self, po, n, dx, jgup, jinv, *, dim=3, periodic=(0, 0, 0), boxes=None, bc_type=None, bc_value=None,
                 params: LepticParams = None):
        self.po = po
        self.dim = dim                                          # 2: arrays (1, ny, nx), y is the vertical
        self.n = tuple(int(x) for x in n)
        self.dx = tuple(float(x) for x in dx)
        self.prm = params or LepticParams()
        self.jgup = [np.ascontiguousarray(j, dtype=np.float64) for j in jgup]
        self.jinv = np.ascontiguousarray(jinv, dtype=np.float64)
        nx, ny, nz = self.n
        bt = list(bc_type) if bc_type is not None else [NEUM] * 6
        bv = list(bc_value) if bc_value is not None else [0.0] * 6
        vd = dim - 1
        if periodic[vd]:
            raise ValueError("periodic vertical direction: unforeseen BC type in verticalLineSolver (:1316)")
        self.vert_bc = (bt[2 * vd], bt[2 * vd + 1])
        if any(t not in (NEUM, DIRI) for t in self.vert_bc):
            raise ValueError("vertical BCs must be Neumann or Dirichlet")
        self.do_horiz = self.vert_bc == (NEUM, NEUM)            # gatherVerticalBCTypes :1615-1628
        if self.do_horiz and dim != 3:
            raise ValueError("horizontal solves need a 3-D level")
        self.diag = all(j.ndim == 3 for j in self.jgup)         # full metric: jgup[d] is (dim, nz, ny, nx) with +1 in d
        if not self.diag and dim != 3:
            raise ValueError("the full-metric iteration is restated for 3-D levels")
        kw = dict(dim=dim, periodic=periodic, boxes=boxes, alpha=0.0, beta=1.0, is_diagonal=self.diag, bc_type=bt, bc_value=bv)
        # the operator handed to define() (m_origOpPtr)
        self.orig = po.Oracle(self.n, self.dx, self.jgup, self.jinv, params=po.default_params(), **kw)
        # m_opPtr and m_mgSolverPtr share the vertical grids' metric: Jgup copied, Jinv = 1 (:197-205)
        # both bottom solvers take the leptic solver's norm type (:274, :387)
        full = dict(self.prm.full, bottom_normType=self.prm.norm_type)
        hor = dict(self.prm.horiz, bottom_normType=self.prm.norm_type)
        self.vert = po.Oracle(self.n, self.dx, self.jgup, np.ones_like(self.jinv), params=po.default_params(**full), **kw)
        self.horiz = None
        if self.do_horiz:
            if periodic[0] or periodic[1]:
                raise ValueError("horizontal solves need non-periodic horizontal directions")
            flat_boxes = None
            if boxes is not None:
                seen = []
                for lo, hi in boxes:
                    b = ((lo[0], lo[1], 0), (hi[0], hi[1], 0))
                    if b not in seen:
                        seen.append(b)
                flat_boxes = seen
            if self.diag:
                self.hjg = [vert_avg_faces(self.jgup[0], nz), vert_avg_faces(self.jgup[1], nz)]
                hmetric = [h[None] for h in self.hjg]
            else:                                               # J g^{ab}, a, b horizontal (LevelGeometryBasics.cpp:545-568)
                self.hjg = [np.stack([vert_avg_faces(self.jgup[a][b], nz) for b in range(2)]) for a in range(2)]
                hmetric = [h[:, None] for h in self.hjg]
            self.horiz = po.Oracle((nx, ny, 1), (self.dx[0], self.dx[1], 1.0), hmetric,
                                   np.ones((1, ny, nx)), dim=2, periodic=(periodic[0], periodic[1], 0),
                                   boxes=flat_boxes, alpha=0.0, beta=1.0, is_diagonal=self.diag,
                                   bc_type=[NEUM] * 6, bc_value=[0.0] * 6, params=po.default_params(**hor))
        self.horiz_remove_avg = True                            # the flattened boxes cover the domain (:410-433)
        self.res_norms = []
        self.exit_status = EXIT["NONE"]
        self.horiz_solves = 0
        self.krylov_used = False

    def solve(self, phi, rhs, homogeneous=False):
        po, prm = self.po, self.prm
        nx, ny, nz = self.n
        if self.dim == 2:                                       # columns run along y: nv = ny, flat arrays are (1, nx)
            nz, ny = ny, 1
        dz = self.dx[self.dim - 1]
        H = dz * nz
        max_order = prm.max_order
        # J * residual to the vertical grids (:698-716)
        res = self.orig.residual(phi, rhs, homogeneous=homogeneous)
        rhs_v = res / self.jinv
        self.res_norms = [self.vert.norm(rhs_v, prm.norm_type)]
        res0 = self.res_norms[0]
        bdry_lo = np.zeros((ny, nx))
        bdry_hi = np.zeros((ny, nx))
        phi_total = np.zeros_like(rhs_v)
        vert_phi = np.zeros_like(rhs_v)
        excess = None
        use_excess = use_horiz = self.do_horiz
        self.horiz_solves = 0
        self.krylov_used = False
        jzz = self.jgup[self.dim - 1] if self.diag else self.jgup[self.dim - 1][self.dim - 1]
        jgzz = np.ascontiguousarray(jzz).reshape(nz + 1, -1)
        for order in range(max_order + 1):
            if self.do_horiz:
                if order >= 1 and self.diag:
                    bdry_lo = np.zeros((ny, nx))     # levelVertHorizGradient: diagonal metric -> zero (:1122-1125)
                    bdry_hi = np.zeros((ny, nx))
                elif order >= 1:                     # -Grad[phi_{order-1}]^z on the boundary faces (:760-762)
                    bdry_lo, bdry_hi = self.vert.leptic_vert_horiz_grad(vert_phi, -1.0)
                if order >= 1 and use_excess:
                    bdry_hi = bdry_hi + excess * 1.0                      # vertPlus (:766)
                if use_excess:
                    excess = po.leptic_vertical_excess(rhs_v.reshape(nz, -1), bdry_lo.ravel(), bdry_hi.ravel(),
                                                       dz).reshape(ny, nx)
                    if order == 1:
                        use_excess = False
                if order == 0 and use_excess:
                    bdry_hi = bdry_hi + excess * -1.0
            sol, rhs_after, info = po.leptic_vertical_solve(rhs_v.reshape(nz, -1), jgzz, dz, self.vert_bc[0],
                                                            self.vert_bc[1], dz_crse=dz, bdry_lo=bdry_lo.ravel(),
                                                            bdry_hi=bdry_hi.ravel())
            if info != 0:
                raise RuntimeError("LepticLapackVerticalSolver: INFO = %d" % info)
            vert_phi = sol.reshape(phi_total.shape)
            rhs_v = rhs_after.reshape(phi_total.shape)
            if use_horiz:
                flat_rhs = horiz_rhs(vert_phi, self.jgup, self.dx, nz) if self.diag else self.vert.leptic_horiz_rhs(vert_phi)
                if use_excess:
                    flat_rhs = flat_rhs + excess * (-1.0 / H)
                horiz_rhs_f = np.zeros((ny, nx)) + flat_rhs                 # setVal(0) + addTo (:811-812)
                hnorm = self.horiz.norm(horiz_rhs_f[None], prm.norm_type)
                if prm.horiz_rhs_tol * res0 > hnorm:
                    use_horiz = False
            if use_horiz:
                hphi, _ = self.horiz.solve(horiz_rhs_f[None], zero_phi=True, force_homogeneous=True)
                hphi = hphi[0]
                if self.horiz_remove_avg:                                   # setZeroAvg (:1665-1709)
                    hphi = hphi - float(np.sum(hphi)) / float(nx * ny)
                vert_phi = vert_phi + hphi[None]                            # ADDEXTRUSION
                self.horiz_solves += 1
            tmp = self.vert.residual(vert_phi, rhs_v, homogeneous=True)
            res_norm = self.vert.norm(tmp, prm.norm_type)
            rel = res_norm / res0
            prev = self.res_norms[-1] / res0
            redu = prev - rel
            if redu <= prm.hang and order == max_order:                    # Krylov (full multigrid) fall-back (:851-875)
                vert_phi, _ = self.vert.solve(rhs_v, zero_phi=True, force_homogeneous=True)
                tmp = self.vert.residual(vert_phi, rhs_v, homogeneous=True)
                res_norm = self.vert.norm(tmp, prm.norm_type)
                rel = res_norm / res0
                self.krylov_used = True
            rhs_v = tmp                                                     # std::swap(rhsPtr, tmpRhsPtr)
            self.res_norms.append(res_norm)
            redu = prev - rel
            if redu > prm.hang or order < max_order:
                phi_total = phi_total + vert_phi
                self.exit_status = EXIT["CONVERGE"] if order < max_order - 1 else EXIT["ITER"]
            elif -redu > prm.hang:
                self.exit_status = EXIT["KABOOM"] if order == 0 else EXIT["DIVERGE"]
                break
            else:
                self.exit_status = EXIT["KABOOM"] if order == 0 else EXIT["HANG"]
                break
            if self.diag:
                use_horiz = False                                           # LevelGeometry::isDiagonal() (:924-926)
        if self.exit_status != EXIT["KABOOM"]:
            return phi + phi_total * 1.0
        return phi.copy()

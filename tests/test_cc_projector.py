"""The cell-centred level projection, LevelCCProjector (src/projection/LevelCCProjector.cpp:156-255): the second of the two
projections of the reference's time step (AMRNavierStokes::doCCProjection, AMRNavierStokesAdvancePPM.cpp:1513-1573), and
the reference projectors' own BC holders (m_divBC = uStarFuncBC on solid walls, m_gradBC = order-2 extrapolated ghosts).

CPU part: the oracle's restatement against things that are not the oracle -- the order-2 ghost differentiates a
quadratic exactly at a wall, the composed projection is the composition of its parts, and on a periodic Cartesian grid
the discrete CC projection removes exactly the modes the wide (2h) Laplacian sees.  GPU part: the CUDA path through the C
ABI against the oracle on identical seeded inputs -- bit-exact for diagonal metrics, <= 1e-13 for full metrics."""
import numpy as np
import pytest

from common import Problem, rel_err

TOL = 1e-13

CASES = {
    "stretched3d": dict(n=(32, 24, 16), map_name="stretched", aniso=4.0),
    "stretched3d_boxes": dict(n=(32, 32, 16), map_name="stretched", aniso=4.0, parts=(1, 4, 1)),
    "periodic_y": dict(n=(32, 16, 16), map_name="stretched", aniso=2.0, periodic=(0, 1, 0)),
    "periodic_xy": dict(n=(16, 16, 8), map_name="stretched", aniso=2.0, periodic=(1, 1, 0)),
    "stretched2d": dict(n=(32, 16, 1), dim=2, map_name="stretched", aniso=4.0),
    "terrain3d": dict(n=(32, 24, 16), map_name="terrain", aniso=4.0),
    "terrain3d_boxes": dict(n=(32, 32, 16), map_name="terrain", aniso=4.0, parts=(2, 2, 1)),
    "terrain2d": dict(n=(32, 16, 1), dim=2, map_name="terrain", aniso=4.0),
    "terrain_periodic_y": dict(n=(16, 16, 16), map_name="terrain", aniso=4.0, periodic=(0, 1, 0)),
}
DIAG = [k for k, v in CASES.items() if v["map_name"] == "stretched"]
SOLVE = dict(eps=1e-11, imin=0, imax=40, num_smooth_down=4, num_smooth_up=4)


def _cells(pb, seed=5):
    rng = np.random.default_rng(seed)
    nx, ny, nz = pb.n
    return [np.ascontiguousarray(rng.uniform(-1.0, 1.0, size=(nz, ny, nx))) for _ in range(pb.dim)]


def _J(pb):
    """a cell-centred Jacobian field (the reference caches J and 1/J separately: LevelGeometry::getCCJ / getCCJinv)"""
    return np.ascontiguousarray(1.0 / pb.jinv)


# ---------------------------------------------------------------------------------------------------------------
# CPU: the oracle
# ---------------------------------------------------------------------------------------------------------------
def test_oracle_extrap_ghost_is_exact_for_quadratics(oracle_lib):
    """ghost = 3 (p1 - p2) + p3 reproduces a quadratic, so the wall-face gradient of phi = x^2 is J g^{xx} 2 x_wall."""
    n = 8
    pb = Problem(n=(n, n, n), map_name="cartesian", aniso=1.0, L=(1.0, 1.0, 1.0))
    orc = pb.oracle(oracle_lib)
    h = pb.spec.dx[0]
    x = (np.arange(n) + 0.5) * h
    phi = np.broadcast_to(x * x, (n, n, n)).copy()
    g = orc.mac_gradient_bc(phi, 1)
    xf = np.arange(n + 1) * h
    want = pb.jgup[0] * (2.0 * xf)
    assert np.abs(g[0] - want).max() <= 1e-12 * np.abs(want).max()
    # the operator's (Neumann) holder puts zero on the wall faces instead
    g0 = orc.mac_gradient_bc(phi, 0)
    assert np.all(g0[0][:, :, 0] == 0.0) and np.all(g0[0][:, :, -1] == 0.0)
    assert np.array_equal(g0[0][:, :, 1:-1], g[0][:, :, 1:-1])


@pytest.mark.parametrize("name", ["stretched3d_boxes", "terrain3d", "periodic_y"])
def test_oracle_cc_projection_is_its_parts(name, oracle_lib):
    """project_cc == [multByJ] -> levelDivergenceCC / dt -> solve -> levelGradientCC -> vel += corr * (-dt) -> [divByJ],
    bit for bit, and levelDivergenceCC == levelDivergenceMAC(wall BC(CellToEdge(u)))"""
    pb = Problem(**CASES[name])
    orc = pb.oracle(oracle_lib, **SOLVE)
    u, J, dt = _cells(pb), _J(pb), 0.25
    new, p, st = orc.project_cc(u, dt, J=J)
    assert st["final_rnorm"] <= 1e-10 * st["initial_rnorm"]
    uJ = [a * J for a in u]
    rhs = orc.cc_divergence(uJ) / dt
    p2, st2 = orc.solve(rhs)
    assert np.array_equal(p, p2)
    gc = orc.cc_gradient(p)
    for d in range(pb.dim):
        assert np.array_equal(new[d], (uJ[d] + gc[d] * (-dt)) * pb.jinv)
    # levelDivergenceCC from numpy faces
    nx, ny, nz = pb.n
    faces = []
    for d in range(pb.dim):
        ax = 2 - d
        a = uJ[d]
        if pb.periodic[d]:
            lo = np.roll(a, 1, axis=ax)
            f = 0.5 * (a + lo)
            f = np.concatenate([f, np.take(f, [0], axis=ax)], axis=ax)
        else:
            sl_lo = [slice(None)] * 3
            sl_hi = [slice(None)] * 3
            sl_lo[ax] = slice(0, -1)
            sl_hi[ax] = slice(1, None)
            inner = 0.5 * (a[tuple(sl_hi)] + a[tuple(sl_lo)])
            z = np.zeros_like(np.take(a, [0], axis=ax))
            f = np.concatenate([z, inner, z], axis=ax)
        faces.append(np.ascontiguousarray(f))
    assert np.array_equal(orc.mac_divergence(faces), orc.cc_divergence(uJ))
    # and levelGradientCC is EdgeToCell of the MAC gradient taken with the extrapolated ghosts
    gm = orc.mac_gradient_bc(p, 1)
    for d in range(pb.dim):
        ax = 2 - d
        sl_lo = [slice(None)] * 3
        sl_hi = [slice(None)] * 3
        sl_lo[ax] = slice(0, -1)
        sl_hi[ax] = slice(1, None)
        assert np.array_equal(gc[d], 0.5 * (gm[d][tuple(sl_lo)] + gm[d][tuple(sl_hi)]))


def test_oracle_cc_projection_periodic_cartesian(oracle_lib):
    """Periodic Cartesian box: Div_cc Grad_cc is the 2h-wide Laplacian, the solver inverts the h-wide one, so the
    projection is approximate -- a smooth mode k loses its divergence up to the factor 1 - cos^2(k h / 2) = O(h^2)."""
    res = []
    for n in (16, 32):
        pb = Problem(n=(n, n, 1), dim=2, map_name="cartesian", L=(1.0, 1.0), periodic=(1, 1, 0))
        orc = pb.oracle(oracle_lib, eps=1e-12, imin=0, imax=40, num_smooth_down=4, num_smooth_up=4)
        c = (np.arange(n) + 0.5) / n
        Y, X = np.meshgrid(c, c, indexing="ij")
        k = 2.0 * np.pi
        # a pure gradient field: u = grad(cos kx cos ky)
        u = [(-k * np.sin(k * X) * np.cos(k * Y)).reshape(1, n, n), (-k * np.cos(k * X) * np.sin(k * Y)).reshape(1, n, n)]
        new, p, st = orc.project_cc(u, 0.5)
        d0 = np.abs(orc.cc_divergence(u)).max()
        d1 = np.abs(orc.cc_divergence(new)).max()
        res.append(d1 / d0)
        assert np.abs(new[0]).max() <= 0.2 * np.abs(u[0]).max()
    assert res[0] < 0.05 and res[1] < res[0] / 3.5, res


# ---------------------------------------------------------------------------------------------------------------
# GPU: CUDA path through the C ABI vs the oracle
# ---------------------------------------------------------------------------------------------------------------
def _check(got, want, exact, what):
    if exact:
        assert np.array_equal(got, want), f"{what}: not bit-exact, rel err {rel_err(got, want):.3e}"
    else:
        assert rel_err(got, want) <= TOL, f"{what}: rel err {rel_err(got, want):.3e}"


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_cc_divergence_gradient_parity(name, oracle_lib, somar):
    pb = Problem(**CASES[name])
    orc, lev = pb.oracle(oracle_lib), pb.level(somar)
    exact = name in DIAG
    u_h, J_h = _cells(pb), _J(pb)
    u = [lev.field(0, a) for a in u_h]
    J = lev.field(0, J_h)
    div = lev.field(0)
    # pure streaming arithmetic: bit-exact for every metric
    lev.cc_divergence(div, u)
    assert np.array_equal(div.get(), orc.cc_divergence(u_h))
    lev.cc_divergence(div, u, J=J, divide_by=0.3)
    assert np.array_equal(div.get(), orc.cc_divergence([a * J_h for a in u_h]) / 0.3)
    for a, f in zip(u_h, u):
        assert np.array_equal(f.get(), a)            # multByJ is folded in: the caller's velocity is not touched
    phi_h = pb.rand()
    phi = lev.field(0, phi_h)
    g = [lev.field(0) for _ in range(pb.dim)]
    lev.cc_gradient(g, phi)
    want = orc.cc_gradient(phi_h)
    for d in range(pb.dim):
        _check(g[d].get(), want[d], exact, f"ccGradient dir {d}")
    # the MAC gradient with the projectors' BC holder, and the wall BC of the divergence
    gf = [lev.field(0) for _ in range(pb.dim)]
    lev.mac_gradient_bc(gf, phi, 1)
    wantf = orc.mac_gradient_bc(phi_h, 1)
    for d in range(pb.dim):
        _check(gf[d].get_face(d), wantf[d], exact, f"macGradient(extrap BC) dir {d}")
    lev.mac_wall_flux_bc(gf)
    for d in range(pb.dim):
        got = gf[d].get_face(d)
        ax = 2 - d
        inner = [slice(None)] * 3
        inner[ax] = slice(1, -1)
        _check(got[tuple(inner)], wantf[d][tuple(inner)], exact, f"wall BC keeps interior faces, dir {d}")
        if not pb.periodic[d]:
            assert np.all(np.take(got, [0, -1], axis=ax) == 0.0)
        else:
            _check(got, wantf[d], exact, f"wall BC leaves periodic direction {d} alone")


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["stretched3d_boxes", "terrain3d", "stretched2d", "periodic_xy", "terrain_periodic_y"])
@pytest.mark.parametrize("with_j", [True, False])
def test_project_cc_parity(name, with_j, oracle_lib, somar):
    pb = Problem(**CASES[name])
    orc, lev = pb.oracle(oracle_lib, **SOLVE), pb.level(somar, **SOLVE)
    u_h, J_h, dt = _cells(pb), _J(pb), 0.25
    want_u, want_p, st_o = orc.project_cc(u_h, dt, J=J_h if with_j else None)
    u = [lev.field(0, a) for a in u_h]
    J = lev.field(0, J_h) if with_j else None
    p = lev.field(0)
    st = lev.project_cc(u, p, dt, J=J)
    assert abs(st["iters"] - st_o["iters"]) <= 1, (st["iters"], st_o["iters"])
    assert st["final_rnorm"] <= 1e-10 * st["initial_rnorm"]
    assert st["initial_rnorm"] == st_o["initial_rnorm"]          # rhs = Div[u J]/dt is bit-exact
    a, b = p.get(), want_p
    assert np.abs((a - a.mean()) - (b - b.mean())).max() <= 1e-7 * np.abs(b).max()
    for d in range(pb.dim):
        assert np.abs(u[d].get() - want_u[d]).max() <= 1e-7 * max(np.abs(want_u[d]).max(), 1.0)
    # with the oracle's pressure the correction pass itself is exact: u <- ((u J) + grad_cc(p) (-dt)) Jinv
    u2 = [lev.field(0, a) for a in u_h]
    p2 = lev.field(0, want_p)
    g = [lev.field(0) for _ in range(pb.dim)]
    lev.cc_gradient(g, p2)
    gh = orc.cc_gradient(want_p)
    for d in range(pb.dim):
        _check(g[d].get(), gh[d], name in DIAG, f"ccGradient of the converged pressure, dir {d}")


@pytest.mark.gpu
def test_project_mac_with_reference_bc_holders(oracle_lib, somar):
    """somar_level_set_projector_bc(1): somar_project zeroes the wall fluxes itself and corrects with the extrapolated-ghost
    gradient, as LevelMACProjector::computeDiv / computeGrad do."""
    pb = Problem(**CASES["stretched3d_boxes"])
    orc, lev = pb.oracle(oracle_lib, **SOLVE), pb.level(somar, **SOLVE)
    rng = np.random.default_rng(9)
    nx, ny, nz = pb.n
    vel_h = [np.ascontiguousarray(rng.uniform(-1, 1, size=(nz + (d == 2), ny + (d == 1), nx + (d == 0)))) for d in range(3)]
    dt = 0.25
    # reference sequence on the oracle: wall faces <- 0, rhs, solve, v += grad_extrap * (-dt)
    w = [v.copy() for v in vel_h]
    for d in range(3):
        ax = 2 - d
        idx = [slice(None)] * 3
        for e in (0, -1):
            idx[ax] = e
            w[d][tuple(idx)] = 0.0
    rhs = orc.mac_divergence(w) / dt
    p_o, st_o = orc.solve(rhs)
    g = orc.mac_gradient_bc(p_o, 1)
    want = [w[d] + g[d] * (-dt) for d in range(3)]
    lev.set_projector_bc(1)
    vel = [lev.field(0).set_face(vel_h[d], d) for d in range(3)]
    p = lev.field(0)
    st = lev.project(vel, p, dt)
    assert st["initial_rnorm"] == st_o["initial_rnorm"]
    assert abs(st["iters"] - st_o["iters"]) <= 1
    for d in range(3):
        assert np.abs(vel[d].get_face(d) - want[d]).max() <= 1e-7 * max(np.abs(want[d]).max(), 1.0)
    # the wall faces did receive the one-sided correction
    assert np.abs(vel[0].get_face(0)[:, :, 0]).max() > 0.0
    lev.set_projector_bc(0)


@pytest.mark.gpu
def test_project_cc_rejects_bad_arguments(somar):
    pb = Problem(n=(16, 16, 8))
    lev = pb.level(somar)
    f = [lev.field(0) for _ in range(3)]
    p, J = lev.field(0), lev.field(0)
    with pytest.raises(somar.SomarError):
        lev.project_cc([f[0], f[0], f[2]], p, 0.1)           # aliased components
    with pytest.raises(somar.SomarError):
        lev.project_cc([f[0], f[1], p], p, 0.1)              # pressure aliases a component
    with pytest.raises(somar.SomarError):
        lev.project_cc(f, p, 0.1, J=p)                       # J aliases the pressure
    with pytest.raises(somar.SomarError):
        lev.set_projector_bc(7)
    thin = Problem(n=(16, 16, 2))
    lev2 = thin.level(somar)
    g = [lev2.field(0) for _ in range(3)]
    with pytest.raises(somar.SomarError):
        lev2.cc_gradient(g, lev2.field(0))                   # the order-2 ghost needs three cells

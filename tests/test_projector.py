"""Projector calculus either side of the pressure solve: MAC divergence, MAC gradient,
velocity correction and the whole BaseProjector::project (BaseProjectorI.H:176-300, a_velIsFlux = true).

CPU part: the oracle's restatement is pinned to identities the reference relies on -- div(grad phi) is the
operator (the projector and the solver must agree on L, or the projection is not idempotent), and a projected
velocity is divergence-free to the solver tolerance.  GPU part: the CUDA path through the C ABI against the oracle
on identical seeded inputs; bit-exact for diagonal metrics, <= 1e-13 for full metrics (BASELINE.json)."""
import numpy as np
import pytest

from common import Problem, rel_err

TOL = 1e-13

CASES = {
    "stretched3d": dict(n=(32, 24, 16), map_name="stretched", aniso=4.0),
    "stretched3d_boxes": dict(n=(32, 32, 16), map_name="stretched", aniso=4.0, parts=(1, 4, 1)),
    "periodic_y": dict(n=(32, 16, 16), map_name="stretched", aniso=2.0, periodic=(0, 1, 0)),
    "dirichlet_top": dict(n=(16, 16, 16), map_name="stretched", aniso=4.0, bc_type=[0, 0, 0, 0, 0, 1],
                          bc_value=[0, 0, 0, 0, 0, 0.5]),
    "stretched2d": dict(n=(32, 16, 1), dim=2, map_name="stretched", aniso=4.0),
    "terrain3d": dict(n=(32, 24, 16), map_name="terrain", aniso=4.0),
    "terrain3d_boxes": dict(n=(32, 32, 16), map_name="terrain", aniso=4.0, parts=(2, 2, 1)),
    "terrain2d": dict(n=(32, 16, 1), dim=2, map_name="terrain", aniso=4.0),
    "terrain_periodic_y": dict(n=(16, 16, 16), map_name="terrain", aniso=4.0, periodic=(0, 1, 0)),
}
DIAG = [k for k, v in CASES.items() if v["map_name"] == "stretched"]


def _faces(pb, seed=3):
    """seeded face-centred 'J u' with zero normal component on non-periodic domain boundaries"""
    rng = np.random.default_rng(seed)
    nx, ny, nz = pb.n
    out = []
    for d in range(pb.dim):
        v = rng.uniform(-1.0, 1.0, size=(nz + (d == 2), ny + (d == 1), nx + (d == 0)))
        ax = 2 - d
        if pb.periodic[d]:
            v[(slice(None),) * ax + (-1,)] = v[(slice(None),) * ax + (0,)]
        else:
            v[(slice(None),) * ax + (0,)] = 0.0
            v[(slice(None),) * ax + (-1,)] = 0.0
        out.append(np.ascontiguousarray(v))
    return out


# ---------------------------------------------------------------------------------------------------------------
# CPU: the oracle
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", list(CASES))
def test_oracle_div_grad_is_the_operator(name, oracle_lib):
    """div(J g grad phi) == L phi bit for bit: MAPPEDMACGRAD and MAPPEDGETFLUX are the same statements (beta = 1)."""
    pb = Problem(**CASES[name])
    orc = pb.oracle(oracle_lib)
    phi = pb.rand()
    for homog in (True, False):
        g = orc.mac_gradient(phi, homogeneous=homog)
        assert np.array_equal(orc.mac_divergence(g), orc.apply_op(phi, homogeneous=homog))


def test_oracle_gradient_of_linear_field(oracle_lib):
    """Cartesian metric (J g^{dd} = dx dy dz / dx_d^2 ... = const): the face gradient of phi = x_d is J g^{dd}."""
    pb = Problem(n=(8, 8, 8), map_name="cartesian", aniso=1.0, periodic=(0, 0, 0), bc_type=[1] * 6)
    orc = pb.oracle(oracle_lib)
    z, y, x = np.meshgrid(*[np.arange(n) + 0.5 for n in (8, 8, 8)], indexing="ij")
    g = orc.mac_gradient(x * pb.spec.dx[0], homogeneous=False)
    assert np.abs(g[0][:, :, 1:-1] - pb.jgup[0][:, :, 1:-1]).max() <= 1e-13 * np.abs(pb.jgup[0]).max()
    assert np.abs(g[1][:, 1:-1, :]).max() <= 1e-13 and np.abs(g[2][1:-1]).max() <= 1e-13


@pytest.mark.parametrize("name", ["stretched3d_boxes", "terrain3d"])
def test_oracle_projection_is_divergence_free(name, oracle_lib):
    pb = Problem(**CASES[name])
    orc = pb.oracle(oracle_lib, eps=1e-11, imin=0, imax=40, num_smooth_down=4, num_smooth_up=4)
    vel = _faces(pb)
    dt = 0.25
    d0 = np.abs(orc.mac_divergence(vel)).max()
    new, p, st = orc.project(vel, dt)
    d1 = np.abs(orc.mac_divergence(new)).max()
    assert st["final_rnorm"] <= 1e-10 * st["initial_rnorm"]
    assert d1 <= 1e-9 * d0, (d0, d1)
    # the correction is -dt * grad(p), nothing else
    g = orc.mac_gradient(p)
    for d in range(pb.dim):
        assert np.array_equal(new[d], vel[d] + g[d] * (-dt))


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
def test_mac_gradient_divergence_parity(name, oracle_lib, somar):
    pb = Problem(**CASES[name])
    orc, lev = pb.oracle(oracle_lib), pb.level(somar)
    exact = name in DIAG
    phi_h = pb.rand()
    phi = lev.field(0, phi_h)
    grads = [lev.field(0) for _ in range(pb.dim)]
    for homog in (False, True):
        lev.mac_gradient(grads, phi, homogeneous=homog)
        want = orc.mac_gradient(phi_h, homogeneous=homog)
        for d in range(pb.dim):
            _check(grads[d].get_face(d), want[d], exact, f"macGradient dir {d} homog {homog}")
    # divergence of seeded fluxes (pure streaming arithmetic: bit-exact for every metric), plain and divided by dt
    flux_h = _faces(pb, seed=11)
    flux = [lev.field(0).set_face(flux_h[d], d) for d in range(pb.dim)]
    div = lev.field(0)
    lev.mac_divergence(div, flux)
    assert np.array_equal(div.get(), orc.mac_divergence(flux_h))
    lev.mac_divergence(div, flux, divide_by=0.3)
    assert np.array_equal(div.get(), orc.mac_divergence(flux_h) / 0.3)
    # div(grad phi) == L phi on the device as well (the projector and the solver share one L)
    lev.mac_gradient(grads, phi, homogeneous=True)
    lev.mac_divergence(div, grads)
    lhs = lev.field(0)
    lev.apply_op(lhs, phi, homogeneous=True)
    assert np.array_equal(div.get(), lhs.get())
    # the correction: v -= dt * grad, fused (the gradient is never stored)
    vel = [lev.field(0).set_face(flux_h[d], d) for d in range(pb.dim)]
    lev.mac_apply_correction(vel, phi, 0.25)
    g = orc.mac_gradient(phi_h, homogeneous=False)
    for d in range(pb.dim):
        _check(vel[d].get_face(d), flux_h[d] + g[d] * (-0.25), exact, f"applyCorrection dir {d}")


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["stretched3d_boxes", "terrain3d", "stretched2d"])
def test_project_parity(name, oracle_lib, somar):
    pb = Problem(**CASES[name])
    kw = dict(eps=1e-11, imin=0, imax=40, num_smooth_down=4, num_smooth_up=4)
    orc, lev = pb.oracle(oracle_lib, **kw), pb.level(somar, **kw)
    vel_h = _faces(pb)
    dt = 0.25
    want_v, want_p, st_o = orc.project(vel_h, dt)
    vel = [lev.field(0).set_face(vel_h[d], d) for d in range(pb.dim)]
    p = lev.field(0)
    st = lev.project(vel, p, dt)
    assert abs(st["iters"] - st_o["iters"]) <= 1, (st["iters"], st_o["iters"])
    assert st["final_rnorm"] <= 1e-10 * st["initial_rnorm"]
    assert st["initial_rnorm"] == st_o["initial_rnorm"]          # rhs = Div[v]/dt is bit-exact
    # converged fields agree to the solver tolerance (the pressure up to its null-space constant)
    a, b = p.get(), want_p
    assert np.abs((a - a.mean()) - (b - b.mean())).max() <= 1e-7 * np.abs(b).max()
    div = lev.field(0)
    lev.mac_divergence(div, vel)
    d0 = np.abs(orc.mac_divergence(vel_h)).max()
    assert np.abs(div.get()).max() <= 1e-9 * d0
    for d in range(pb.dim):
        assert np.abs(vel[d].get_face(d) - want_v[d]).max() <= 1e-7 * max(np.abs(want_v[d]).max(), 1.0)


@pytest.mark.gpu
def test_project_rejects_bad_arguments(somar):
    pb = Problem(n=(16, 16, 8))
    lev = pb.level(somar)
    f = [lev.field(0) for _ in range(3)]
    p = lev.field(0)
    with pytest.raises(somar.SomarError):
        lev.project([f[0], f[0], f[2]], p, 0.1)      # aliased components
    with pytest.raises(somar.SomarError):
        lev.project([f[0], f[1], p], p, 0.1)         # pressure aliases a component
    c = lev.field(1)
    with pytest.raises(somar.SomarError):
        lev.mac_gradient([f[0], f[1], c], p)         # wrong depth

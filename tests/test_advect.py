"""PPM / corner-transport-upwind scalar predictor (SURVEY.md 8(a) MappedAdvectionUtil::predictScalar row, 8(f)-3) on
fully periodic levels (BASELINE.json config "TaylorGreen 2D/3D periodic").

CPU: the numpy restatement oracle/advect_ref.py is pinned to properties of the scheme (the reference ships no
vectors): constants are preserved, a smooth profile carried by a uniform velocity is predicted to second order at
t + dt/2, stagnant faces take the average state, the limiter keeps the face states inside the local data range; the
high-order (extremum-preserving) branch gives 4th-order faces and keeps a smooth extremum that PPMLimiterF flattens.
GPU: the CUDA kernels through the C ABI against the oracle on seeded inputs -- BIT-EXACT (same statements, no FMA
contraction) for every option combination, in 2-D and 3-D."""
import itertools

import numpy as np
import pytest

from common import Problem
from oracle import advect_ref as ar


def _smooth(shape, rng, amp=1.0):
    nz, ny, nx = shape
    z, y, x = np.meshgrid(np.arange(nz) / max(nz, 1), np.arange(ny) / ny, np.arange(nx) / nx, indexing="ij")
    a = np.zeros(shape)
    for _ in range(4):
        kx, ky, kz = rng.integers(0, 3, size=3)
        ph = rng.uniform(0, 2 * np.pi, size=3)
        a += rng.uniform(-1, 1) * np.cos(2 * np.pi * kx * x + ph[0]) * np.cos(2 * np.pi * ky * y + ph[1]) * \
            np.cos(2 * np.pi * (kz if nz > 1 else 0) * z + ph[2])
    return amp * a


def _inputs(shape, dim, seed=0):
    rng = np.random.default_rng(seed)
    W = _smooth(shape, rng) + 0.2 * rng.uniform(-1, 1, shape)
    J = 1.0 + 0.3 * _smooth(shape, rng, 0.25)
    assert J.min() > 0.2
    old_vel = [_smooth(shape, rng) for _ in range(dim)]
    adv_vel = [_smooth(shape, rng) for _ in range(dim)]
    adv_vel[0][..., ::5] = 0.0          # some stagnant faces: the average-state branch of the Riemann solver
    src = _smooth(shape, rng)
    return W, J, old_vel, adv_vel, src


# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("lim,ho", [(False, False), (True, False), (False, True), (True, True)])
def test_oracle_preserves_constants(dim, lim, ho):
    shape = (6 if dim == 3 else 1, 10, 12)
    _, J, ov, av, _ = _inputs(shape, dim)
    out = ar.predict_scalar(np.full(shape, 2.5), J, ov, av, 0.01, [0.1] * 3, dim, use_limiting=lim, return_flux=False,
                            high_order=ho)
    for o in out:
        assert np.abs(o - 2.5).max() <= 1e-14


@pytest.mark.parametrize("dim", [2, 3])
def test_oracle_uniform_translation_second_order(dim):
    errs = []
    u = [0.7, -0.4, 0.3][:dim]
    for n in (16, 32, 64):
        nz = n if dim == 3 else 1
        dx = [1.0 / n] * 3
        z, y, x = np.meshgrid((np.arange(nz) + 0.5) / max(nz, 1), (np.arange(n) + 0.5) / n, (np.arange(n) + 0.5) / n,
                              indexing="ij")
        f = lambda x, y, z: np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y) * (np.cos(2 * np.pi * z) if dim == 3 else 1.0)
        W = f(x, y, z)
        vel = [np.full_like(W, c) for c in u]
        dt = 0.4 * dx[0]
        out = ar.predict_scalar(W, np.ones_like(W), vel, vel, dt, dx, dim, return_flux=False)
        exact = f(x - 0.5 * dx[0] - u[0] * dt / 2, y - u[1] * dt / 2, (z - u[2] * dt / 2) if dim == 3 else z)
        errs.append(np.abs(out[0] - exact).max())
    assert errs[1] < errs[0] / 3.5 and errs[2] < errs[1] / 3.5, errs


def test_oracle_stagnant_faces_and_limiter_bounds():
    shape = (1, 16, 16)
    W, J, ov, av, _ = _inputs(shape, 2, seed=3)
    zero = [np.zeros(shape), np.zeros(shape)]
    a = ar.predict_scalar(W, J, zero, zero, 0.01, [0.1] * 3, 2, use_upwinding=True, return_flux=False)
    b = ar.predict_scalar(W, J, zero, zero, 0.01, [0.1] * 3, 2, use_upwinding=False, return_flux=False)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)                       # |u| < SMALLNUMBER: RiemannSolver == AvgStates
    # with limiting and J = 1 the normal-predictor states stay within the range of the three cells around them
    m, p = ar.ppm_normal_pred(W, np.ones(shape), np.zeros(shape), 0, 0.1, True)
    lo = np.minimum(np.minimum(np.roll(W, 1, 2), W), np.roll(W, -1, 2)) - 1e-13
    hi = np.maximum(np.maximum(np.roll(W, 1, 2), W), np.roll(W, -1, 2)) + 1e-13
    assert (m >= lo).all() and (m <= hi).all() and (p >= lo).all() and (p <= hi).all()


def test_oracle_fourth_order_faces_and_extremum_preserving_limiter():
    """m_useHighOrderLimiter: FourthInterpFaces is exact for cell averages of cubics (4th order in general); the
    CheckCubicLimiterF limiter leaves a smooth extremum alone where PPMLimiterF flattens it, and still limits at a step."""
    errs = []
    for n in (16, 32):
        xf = np.arange(n + 1) / n
        F = -np.cos(2 * np.pi * xf) / (2 * np.pi)                     # antiderivative of sin(2 pi x)
        avg = ((F[1:] - F[:-1]) * n).reshape(1, 1, n)                 # exact cell averages
        face = ar.ppm_face_values(avg, np.ones_like(avg), 0, False, high_order=True)
        errs.append(np.abs(face[0, 0] - np.sin(2 * np.pi * xf[:-1])).max())
    assert errs[1] < errs[0] / 14.0, errs                              # 4th order
    n = 32
    x = (np.arange(n) + 0.5) / n
    W = np.sin(2 * np.pi * x).reshape(1, 1, n)
    one, zero = np.ones_like(W), np.zeros_like(W)
    i_max = int(np.argmax(W))
    m_un, p_un = ar.ppm_normal_pred(W, one, zero, 0, 0.0, False, high_order=True)
    m_ho, p_ho = ar.ppm_normal_pred(W, one, zero, 0, 0.0, True, high_order=True)
    m_lo, p_lo = ar.ppm_normal_pred(W, one, zero, 0, 0.0, True, high_order=False)
    assert m_ho[0, 0, i_max] == m_un[0, 0, i_max] and p_ho[0, 0, i_max] == p_un[0, 0, i_max]   # extremum kept
    assert m_lo[0, 0, i_max] == W[0, 0, i_max] and p_lo[0, 0, i_max] == W[0, 0, i_max]         # PPM flattens it
    step = np.where(x < 0.5, 0.0, 1.0).reshape(1, 1, n)
    m, p = ar.ppm_normal_pred(step, one, zero, 0, 0.0, True, high_order=True)
    m_u, p_u = ar.ppm_normal_pred(step, one, zero, 0, 0.0, False, high_order=True)
    # at the step the unlimited parabola of the last 0-cell reaches 1/2; the limiter pulls it back to 1/6 (rho = 0:
    # dWP = -2 dWM) and never moves a state further out than the 1/12 under/overshoot of the 4th-order face value
    assert p_u.max() > 1.0 + 1.0 / 13.0 and np.abs(p - p_u).max() > 0.3
    assert m.min() >= -1.0 / 12.0 - 1e-12 and p.max() <= 1.0 + 1.0 / 12.0 + 1e-12


# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dim", [2, 3])
def test_oracle_low_order_predictors(dim):
    """normalPredOrder 0 (CTU) and 1 (PLM): constants preserved; PLM with fourth-order slopes is second order on a
    uniform translation, CTU first order; and the reference's quirk -- PLM with useLimiting but WITHOUT fourth-order
    slopes hands zeroed one-sided differences to the limiter (AdvectUtil.cpp:1116-1128) and so equals CTU bit for bit."""
    shape = (6 if dim == 3 else 1, 10, 12)
    W, J, ov, av, _ = _inputs(shape, dim)
    for order, lim, fourth in itertools.product([0, 1], [False, True], [False, True]):
        out = ar.predict_scalar(np.full(shape, 2.5), J, ov, av, 0.01, [0.1] * 3, dim, use_limiting=lim, return_flux=False,
                                normal_pred_order=order, fourth_order_slopes=fourth)
        for o in out:
            assert np.abs(o - 2.5).max() <= 1e-14
    ctu = ar.predict_scalar(W, J, ov, av, 0.01, [0.1] * 3, dim, normal_pred_order=0)
    plm_lim = ar.predict_scalar(W, J, ov, av, 0.01, [0.1] * 3, dim, normal_pred_order=1, use_limiting=True)
    plm = ar.predict_scalar(W, J, ov, av, 0.01, [0.1] * 3, dim, normal_pred_order=1)
    for a, b, c in zip(ctu, plm_lim, plm):
        assert np.array_equal(a, b)
        assert not np.array_equal(a, c)
    u = [0.7, -0.4, 0.3][:dim]
    errs = {0: [], 1: []}
    for n in (16, 32, 64):
        nz = n if dim == 3 else 1
        dx = [1.0 / n] * 3
        z, y, x = np.meshgrid((np.arange(nz) + 0.5) / max(nz, 1), (np.arange(n) + 0.5) / n, (np.arange(n) + 0.5) / n,
                              indexing="ij")
        f = lambda x, y, z: np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y) * (np.cos(2 * np.pi * z) if dim == 3 else 1.0)
        Wn = f(x, y, z)
        vel = [np.full_like(Wn, c) for c in u]
        dt = 0.4 * dx[0]
        exact = f(x - 0.5 * dx[0] - u[0] * dt / 2, y - u[1] * dt / 2, (z - u[2] * dt / 2) if dim == 3 else z)
        for order in (0, 1):
            out = ar.predict_scalar(Wn, np.ones_like(Wn), vel, vel, dt, dx, dim, return_flux=False, normal_pred_order=order,
                                    fourth_order_slopes=True)
            errs[order].append(np.abs(out[0] - exact).max())
    assert errs[1][1] < errs[1][0] / 3.5 and errs[1][2] < errs[1][1] / 3.5, errs
    assert errs[0][2] > errs[0][1] / 2.5 and errs[0][2] > 10 * errs[1][2], errs      # CTU: first order only


OPTS = list(itertools.product([False, True], [False, True], [False, True], [False, True]))   # lim, upwind, src, flux


@pytest.mark.gpu
@pytest.mark.parametrize("dim,n", [(2, (24, 20, 1)), (3, (16, 12, 10))])
def test_predict_scalar_parity(dim, n, somar):
    pb = Problem(n=n, dim=dim, map_name="cartesian", aniso=1.0, periodic=(1, 1, 1 if dim == 3 else 0))
    lev = pb.level(somar)
    W, J, ov, av, src = _inputs(pb.shape, dim, seed=5)
    dt = 0.3 * min(pb.spec.dx[:dim])
    fW, fJ, fS = lev.field(0, W), lev.field(0, J), lev.field(0, src)
    fov = [lev.field(0, v) for v in ov]
    fav = [lev.field(0, v) for v in av]
    out = [lev.field(0) for _ in range(dim)]
    for (lim, up, with_src, flux), ho in itertools.product(OPTS, [False, True]):
        lev.advect_predict_scalar(out, fW, fov, fav, fJ, dt, source=fS if with_src else None, use_limiting=lim,
                                  use_upwinding=up, return_flux=flux, use_high_order_limiter=ho)
        want = ar.predict_scalar(W, J, ov, av, dt, pb.spec.dx, dim, source=src if with_src else None, use_limiting=lim,
                                 use_upwinding=up, return_flux=flux, high_order=ho)
        for d in range(dim):
            got = out[d].get()
            assert np.array_equal(got, want[d]), (dim, d, lim, up, with_src, flux, ho, np.abs(got - want[d]).max())
    # the other two normal predictors: CTU (order 0) and PLM (order 1, with and without fourth-order slopes)
    for order, lim, fourth, with_src, ho in itertools.product([0, 1], [False, True], [False, True], [False, True], [False, True]):
        lev.advect_predict_scalar(out, fW, fov, fav, fJ, dt, source=fS if with_src else None, use_limiting=lim,
                                  normal_pred_order=order, use_fourth_order_slopes=fourth, use_high_order_limiter=ho)
        want = ar.predict_scalar(W, J, ov, av, dt, pb.spec.dx, dim, source=src if with_src else None, use_limiting=lim,
                                 normal_pred_order=order, fourth_order_slopes=fourth, high_order=ho)
        for d in range(dim):
            got = out[d].get()
            assert np.array_equal(got, want[d]), (dim, d, order, lim, fourth, with_src, ho, np.abs(got - want[d]).max())
    lev.advect_predict_scalar(out, fW, fov, fav, fJ, dt, source=fS, use_limiting=True, use_high_order_limiter=True)
    # the flux's periodic image is in place: the MAC divergence of the device equals the wrapped numpy divergence
    div = lev.field(0)
    lev.mac_divergence(div, out)
    want = ar.predict_scalar(W, J, ov, av, dt, pb.spec.dx, dim, source=src, use_limiting=True, use_upwinding=True,
                             high_order=True)      # the last combination of the loop
    s = (np.roll(want[0], -1, 2) - want[0]) * (1.0 / pb.spec.dx[0])
    s = s + (np.roll(want[1], -1, 1) - want[1]) * (1.0 / pb.spec.dx[1])
    if dim == 3:
        s = s + (np.roll(want[2], -1, 0) - want[2]) * (1.0 / pb.spec.dx[2])
    assert np.array_equal(div.get(), pb.jinv * s)
    # the finite-volume update of getNewScalar: new = old + (-Div[flux]) * dt
    new = lev.field(0)
    lev.advect_update(new, fW, out, dt)
    assert np.array_equal(new.get(), W + ((pb.jinv * s) * -1.0) * dt)
    # conservation on the periodic domain: sum of dV-weighted change is zero (dV = prod(dx) / Jinv)
    assert abs(((new.get() - W) / pb.jinv).sum()) <= 1e-11 * np.abs(W / pb.jinv).sum()


@pytest.mark.gpu
def test_predict_scalar_rejects_what_is_not_built(somar):
    pb = Problem(n=(16, 16, 8), map_name="cartesian", aniso=1.0, periodic=(1, 1, 0))
    lev = pb.level(somar)
    f = [lev.field(0) for _ in range(9)]
    with pytest.raises(somar.SomarError):          # a non-periodic side
        lev.advect_predict_scalar(f[0:3], f[3], f[4:7], f[4:7], f[8], 0.1)
    pb = Problem(n=(16, 16, 8), map_name="cartesian", aniso=1.0, periodic=(1, 1, 1))
    lev = pb.level(somar)
    f = [lev.field(0) for _ in range(12)]
    with pytest.raises(somar.SomarError):
        lev.advect_predict_scalar(f[0:3], f[3], f[4:7], f[7:10], f[10], 0.1, normal_pred_order=3)
    with pytest.raises(somar.SomarError):          # output aliases an input
        lev.advect_predict_scalar([f[3], f[1], f[2]], f[3], f[4:7], f[7:10], f[10], 0.1)

"""BASELINE.json config 2, "TaylorGreen 2D periodic (projection + advection correctness vs analytic decay)".

The Taylor-Green vortex  u = sin(kx) cos(ky) F,  v = -cos(kx) sin(ky) F,  p = F^2 (cos 2kx + cos 2ky) / 4  (the reference's
analytic solution: src/BCutil/TaylorGreenBCUtilF.ChF:32-105, F(t) = exp(-2 nu k^2 t)) is, for nu = 0, a STEADY solution of
the Euler equations: -(u.grad)u = grad p.  One inviscid step of the reference's algorithm on a periodic level,

    u* = u - dt Div[u_half u_adv]        MappedAdvectionUtil::predictScalar per component (traced with the lagged pressure
                                         gradient as its source, as the reference's velocity predictor does -- without it
                                         the half-time states are O(dt) off and the step is first order) + the update
    u_f = average of u* to the faces;    u_f -= dt Grad[p],  Lap[p] = Div[u_f] / dt        BaseProjector::project

must therefore leave the velocity unchanged and return the analytic pressure, both to second order in h.  That pins the
advection predictor, the update, the MAC divergence / gradient and the pressure solve TOGETHER against something that
is neither the oracle nor a restatement: a wrong upwinding, a missing transverse term or a sign in the correction
shows up as an O(1) acceleration.  Run for the oracle stack on the CPU and for the CUDA path on the GPU.
"""
import numpy as np
import pytest

from common import Problem
from oracle import advect_ref as ar

K = 2.0 * np.pi
SOLVE = dict(eps=1e-12, num_smooth_down=4, num_smooth_up=4, num_smooth_bottom=2, imin=0, imax=40)


def _tg(x, y):
    return np.sin(K * x) * np.cos(K * y), -np.cos(K * x) * np.sin(K * y)


def _grids(n):
    h = 1.0 / n
    c = (np.arange(n) + 0.5) * h          # centres
    f = np.arange(n) * h                   # low faces
    return h, c, f


def _measure(n, uf_new, vf_new, p, dt):
    """max-norm acceleration of the face velocities and pressure error (mean removed) for the result of one step"""
    h, c, f = _grids(n)
    Y, X = np.meshgrid(c, c, indexing="ij")
    ucc, vcc = _tg(X, Y)
    uf_old = 0.5 * (np.roll(ucc, 1, 1) + ucc)      # the same face average the step applies to u*
    vf_old = 0.5 * (np.roll(vcc, 1, 0) + vcc)
    acc = max(np.abs(uf_new - uf_old).max(), np.abs(vf_new - vf_old).max()) / dt
    p_exact = 0.25 * (np.cos(2 * K * X) + np.cos(2 * K * Y))
    perr = np.abs((p - p.mean()) - (p_exact - p_exact.mean())).max()
    return acc, perr


def _inputs(n):
    h, c, f = _grids(n)
    Y, X = np.meshgrid(c, c, indexing="ij")
    ucc, vcc = _tg(X, Y)
    Yx, Xx = np.meshgrid(c, f, indexing="ij")      # x-faces: (x_f, y_c)
    Yy, Xy = np.meshgrid(f, c, indexing="ij")      # y-faces: (x_c, y_f)
    uadv = _tg(Xx, Yx)[0]
    vadv = _tg(Xy, Yy)[1]
    # the analytic face velocities are discretely divergence-free (k = l)
    div = (np.roll(uadv, -1, 1) - uadv) / h + (np.roll(vadv, -1, 0) - vadv) / h
    assert np.abs(div).max() < 1e-11
    dt = 0.4 * h
    sh = (1, n, n)
    src = [0.5 * K * np.sin(2 * K * X), 0.5 * K * np.sin(2 * K * Y)]      # -grad p of the analytic pressure
    return h, dt, [a.reshape(sh) for a in (ucc, vcc)], [a.reshape(sh) for a in (uadv, vadv)], [a.reshape(sh) for a in src]


def _step_oracle(n, po):
    h, dt, cc, adv, src = _inputs(n)
    pb = Problem(n=(n, n, 1), dim=2, map_name="cartesian", L=(1.0, 1.0), periodic=(1, 1, 0))
    one = np.ones((1, n, n))
    star = []
    for w, sc in zip(cc, src):
        flux = ar.predict_scalar(w, one, cc, adv, dt, pb.spec.dx, 2, source=sc, use_limiting=True, use_upwinding=True,
                                 return_flux=True, high_order=True)
        div = (np.roll(flux[0], -1, 2) - flux[0]) / h + (np.roll(flux[1], -1, 1) - flux[1]) / h
        star.append(w - dt * div)
    uf = 0.5 * (np.roll(star[0], 1, 2) + star[0])
    vf = 0.5 * (np.roll(star[1], 1, 1) + star[1])
    # oracle face arrays carry n + 1 points in their own direction (the periodic image)
    uf1 = np.concatenate([uf, uf[:, :, :1]], axis=2)
    vf1 = np.concatenate([vf, vf[:, :1, :]], axis=1)
    orc = pb.oracle(po, **SOLVE)
    vel, p, st = orc.project([uf1, vf1], dt)
    assert st["exit_status"] & 1 or st["final_rnorm"] <= 1e-10 * st["initial_rnorm"], st
    return _measure(n, vel[0][0, :, :n], vel[1][0, :n, :], p[0], dt)


def _step_cuda(n, somar):
    h, dt, cc, adv, src = _inputs(n)
    pb = Problem(n=(n, n, 1), dim=2, map_name="cartesian", L=(1.0, 1.0), periodic=(1, 1, 0))
    lev = pb.level(somar, **SOLVE)
    fcc = [lev.field(0, w) for w in cc]
    fadv = [lev.field(0, w) for w in adv]
    fsrc = [lev.field(0, w) for w in src]
    fJ = lev.field(0, np.ones((1, n, n)))
    flux = [lev.field(0), lev.field(0)]
    star = []
    for w, sc in zip(fcc, fsrc):
        lev.advect_predict_scalar(flux, w, fcc, fadv, fJ, dt, source=sc, use_limiting=True, use_upwinding=True,
                                  return_flux=True, use_high_order_limiter=True)
        new = lev.field(0)
        lev.advect_update(new, w, flux, dt)
        star.append(new.get())
    uf = 0.5 * (np.roll(star[0], 1, 2) + star[0])
    vf = 0.5 * (np.roll(star[1], 1, 1) + star[1])
    fu, fv, fp = lev.field(0), lev.field(0), lev.field(0)
    fu.set_face(np.ascontiguousarray(np.concatenate([uf, uf[:, :, :1]], axis=2)), 0)
    fv.set_face(np.ascontiguousarray(np.concatenate([vf, vf[:, :1, :]], axis=1)), 1)
    st = lev.project([fu, fv], fp, dt)
    assert st["exit_status"] & 1 or st["final_rnorm"] <= 1e-10 * st["initial_rnorm"], st
    out = _measure(n, fu.get_face(0)[0, :, :n], fv.get_face(1)[0, :n, :], fp.get()[0], dt)
    lev.destroy()
    return out


def _check_second_order(results):
    (a0, p0), (a1, p1), (a2, p2) = results
    # steady state: the acceleration is truncation error only, and it falls like h^2 (|u.grad u| is ~ k = 6.3)
    assert a0 < 0.5 and a1 < a0 / 3.3 and a2 < a1 / 3.3, results
    assert p1 < p0 / 3.3 and p2 < p1 / 3.3, results
    assert p2 < 2e-3, results


def test_taylor_green_step_oracle_stack(oracle_lib):
    _check_second_order([_step_oracle(n, oracle_lib) for n in (16, 32, 64)])


@pytest.mark.gpu
def test_taylor_green_step_cuda(oracle_lib, somar):
    res = [_step_cuda(n, somar) for n in (32, 64, 128)]
    _check_second_order(res)
    # and the two stacks agree with each other at the common size
    a_o, p_o = _step_oracle(64, oracle_lib)
    assert abs(res[1][0] - a_o) <= 1e-6 * a_o + 1e-12 and abs(res[1][1] - p_o) <= 1e-6 * p_o + 1e-12, (res[1], a_o, p_o)

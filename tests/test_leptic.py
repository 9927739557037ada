"""Sweeps of the leptic solver (SURVEY.md 8(a) LevelLepticSolver row): batched vertical tridiagonal solves
(TRIDIAGPOISSONNN1DFAB for Neum-Neum columns, LAPACK dptsv otherwise), vertical integral / excess, extrusion and
the flat horizontal divergence.

CPU: the oracle restatement (oracle/leptic.cpp) is pinned to the mathematical definition -- the exact solution of
each column's tridiagonal system (dense solve) -- because the reference ships no vectors and LAPACK is not vendored.
GPU: the CUDA kernels through the C ABI against the oracle, bit-exact (same statements, no FMA contraction)."""
import numpy as np
import pytest

from common import Problem

NEUM, DIRI, CF = 0, 1, 3


def _jgzz(pb):
    j = pb.jgup[pb.dim - 1]
    return np.ascontiguousarray(j[pb.dim - 1] if j.ndim == 4 else j)


def _cols(pb, a):
    """(nv, nh) view of a cell/face array of the problem: the vertical index first"""
    if pb.dim == 3:
        return a.reshape(a.shape[0], -1)
    return a.reshape(a.shape[1], a.shape[2])      # (1, ny, nx): y is the vertical


def _matrix(J, dz, lo, hi, dzc):
    nv = len(J) - 1
    A = np.zeros((nv, nv))
    for k in range(nv):
        A[k, k] = -(J[k] + J[k + 1])
        if k > 0:
            A[k, k - 1] = J[k]
        if k < nv - 1:
            A[k, k + 1] = J[k + 1]
    al = 1.0 - 2.0 * dz / (dzc + dz)
    A[0, 0] = {NEUM: -J[1], DIRI: -(2 * J[0] + J[1]), CF: -((1 - al) * J[0] + J[1])}[lo]
    A[-1, -1] = {NEUM: -J[nv - 1], DIRI: -(J[nv - 1] + 2 * J[nv]), CF: -(J[nv - 1] + (1 - al) * J[nv])}[hi]
    return A / dz ** 2


CASES = {
    "stretched3d": dict(n=(24, 12, 32), map_name="stretched", aniso=4.0),
    "terrain3d": dict(n=(16, 12, 24), map_name="terrain", aniso=4.0),
    "stretched2d": dict(n=(32, 24, 1), dim=2, map_name="stretched", aniso=4.0),
}
BCS = [(DIRI, NEUM), (NEUM, DIRI), (DIRI, DIRI), (CF, DIRI), (NEUM, CF)]


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_vertical_solve_is_exact(name, oracle_lib):
    pb = Problem(**CASES[name])
    J = _jgzz(pb)
    dz = pb.spec.dx[pb.dim - 1]
    rhs = pb.rand()
    Jc, rc = _cols(pb, J), _cols(pb, rhs)
    for lo, hi in BCS:
        phi, rhs_after, info = oracle_lib.leptic_vertical_solve(rc, Jc, dz, lo, hi, dz_crse=2 * dz)
        assert info == 0 and np.array_equal(rhs_after, rc)
        for c in range(0, rc.shape[1], 7):
            x = np.linalg.solve(_matrix(Jc[:, c], dz, lo, hi, 2 * dz), rc[:, c])
            assert np.abs(x - phi[:, c]).max() <= 1e-11 * np.abs(x).max()
    # Neum-Neum: solvable right-hand sides (zero column sums); the solver returns the zero-mean solution
    r0 = rc - rc.mean(axis=0, keepdims=True)
    phi, _, _ = oracle_lib.leptic_vertical_solve(r0, Jc, dz, NEUM, NEUM)
    assert np.abs(phi.mean(axis=0)).max() <= 1e-12 * np.abs(phi).max()
    for c in range(0, rc.shape[1], 7):
        res = _matrix(Jc[:, c], dz, NEUM, NEUM, dz) @ phi[:, c] - r0[:, c]
        assert np.abs(res).max() <= 1e-10 * np.abs(r0[:, c]).max()


def test_oracle_neumann_values_are_rolled_into_rhs(oracle_lib):
    """d_z(J d_z phi) = rhs with J d_z phi = b at the bottom equals the homogeneous problem for rhs + b/dz (:1343)."""
    pb = Problem(**CASES["stretched3d"])
    J, rhs = _cols(pb, _jgzz(pb)), _cols(pb, pb.rand())
    dz = pb.spec.dx[2]
    b = np.random.default_rng(5).uniform(-1, 1, rhs.shape[1])
    a, rhs_after, _ = oracle_lib.leptic_vertical_solve(rhs, J, dz, NEUM, DIRI, bdry_lo=b)
    r2 = rhs.copy()
    r2[0] += b / dz
    c, _, _ = oracle_lib.leptic_vertical_solve(r2, J, dz, NEUM, DIRI)
    assert np.abs(a - c).max() <= 1e-13 * np.abs(c).max()
    assert np.abs(rhs_after - rhs).max() <= 4e-16 * np.abs(rhs).max() * 8      # (rhs + r) - r, as the reference leaves it


def test_oracle_flat_sweeps(oracle_lib):
    rng = np.random.default_rng(2)
    f = rng.uniform(-1, 1, (10, 6, 8))
    flat = rng.uniform(-1, 1, (6, 8))
    assert np.abs(oracle_lib.leptic_vert_integral(flat, f, 0.3) - (flat + 0.3 * f.sum(axis=0))).max() <= 1e-14
    assert np.array_equal(oracle_lib.leptic_add_extrusion(f, flat), f + flat[None])
    fx = rng.uniform(-1, 1, (6, 9))
    assert np.array_equal(oracle_lib.leptic_accum_div(flat, fx, 0, 2.0), flat + (fx[:, 1:] - fx[:, :-1]) * 2.0)
    fy = rng.uniform(-1, 1, (7, 8))
    assert np.array_equal(oracle_lib.leptic_accum_div(flat, fy, 1, 2.0), flat + (fy[1:] - fy[:-1]) * 2.0)
    lo, hi = rng.uniform(-1, 1, (6, 8)), rng.uniform(-1, 1, (6, 8))
    ex = oracle_lib.leptic_vertical_excess(f, lo, hi, 0.3)
    assert np.abs(ex - (hi - lo - 0.3 * f.sum(axis=0))).max() <= 1e-14


# ---------------------------------------------------------------------------------------------------------------
def _flat_field(lev, pb, flat2d):
    """a depth-0 field whose vertical index 0 holds `flat2d` ((ny, nx) in 3-D, (nx,) in 2-D)"""
    a = np.zeros(pb.shape)
    if pb.dim == 3:
        a[0] = flat2d
    else:
        a[0, 0] = flat2d
    return lev.field(0, a)


def _flat_of(pb, f):
    a = f.get()
    return a[0] if pb.dim == 3 else a[0, 0]


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_vertical_solve_parity(name, oracle_lib, somar):
    pb = Problem(**CASES[name])
    lev = pb.level(somar)
    J = _cols(pb, _jgzz(pb))
    dz = pb.spec.dx[pb.dim - 1]
    rhs_h = pb.rand()
    rng = np.random.default_rng(9)
    nh = J.shape[1]
    blo, bhi = rng.uniform(-1, 1, nh), rng.uniform(-1, 1, nh)
    hshape = (pb.n[1], pb.n[0]) if pb.dim == 3 else (pb.n[0],)
    flo, fhi = _flat_field(lev, pb, blo.reshape(hshape)), _flat_field(lev, pb, bhi.reshape(hshape))
    phi = lev.field(0)
    for lo, hi in BCS + [(NEUM, NEUM)]:
        for with_b in (False, True):
            rhs = lev.field(0, rhs_h)
            kw = dict(bdry_lo=flo, bdry_hi=fhi) if with_b else {}
            lev.leptic_vertical_solve(phi, rhs, lo, hi, dz_crse=2 * dz, **kw)
            okw = dict(bdry_lo=blo, bdry_hi=bhi) if with_b else {}
            want, rhs_after, info = oracle_lib.leptic_vertical_solve(_cols(pb, rhs_h), J, dz, lo, hi, dz_crse=2 * dz, **okw)
            assert info == 0
            assert np.array_equal(_cols(pb, phi.get()), want), (lo, hi, with_b)
            assert np.array_equal(_cols(pb, rhs.get()), rhs_after), (lo, hi, with_b)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["stretched3d", "stretched2d"])
def test_flat_sweeps_parity(name, oracle_lib, somar):
    pb = Problem(**CASES[name])
    lev = pb.level(somar)
    rng = np.random.default_rng(4)
    f_h = pb.rand()
    hshape = (pb.n[1], pb.n[0]) if pb.dim == 3 else (pb.n[0],)
    flat_h, lo_h, hi_h = (rng.uniform(-1, 1, hshape) for _ in range(3))
    dz = pb.spec.dx[pb.dim - 1]
    f = lev.field(0, f_h)
    flat = _flat_field(lev, pb, flat_h)
    lev.leptic_vert_integral(flat, f, -dz)
    want = oracle_lib.leptic_vert_integral(flat_h.reshape(-1), _cols(pb, f_h), -dz)
    assert np.array_equal(_flat_of(pb, flat).reshape(-1), want)
    lo, hi = _flat_field(lev, pb, lo_h), _flat_field(lev, pb, hi_h)
    ex = lev.field(0)
    lev.leptic_vertical_excess(ex, f, lo, hi)
    want = oracle_lib.leptic_vertical_excess(_cols(pb, f_h), lo_h.reshape(-1), hi_h.reshape(-1), dz)
    assert np.array_equal(_flat_of(pb, ex).reshape(-1), want)
    flat2 = _flat_field(lev, pb, flat_h)
    lev.leptic_add_extrusion(f, flat2)
    want = oracle_lib.leptic_add_extrusion(_cols(pb, f_h), flat_h.reshape(-1))
    assert np.array_equal(_cols(pb, f.get()), want)
    # flat horizontal divergence of a face field (vertical index 0 of a face-centred field)
    for d in range(pb.dim - 1):
        sh = list(pb.shape)
        sh[2 - d] += 1
        fx_h = np.zeros(sh)
        fl = rng.uniform(-1, 1, [s + (i == len(hshape) - 1 - d) for i, s in enumerate(hshape)])
        if pb.dim == 3:
            fx_h[0] = fl
        else:
            fx_h[0, 0] = fl
        fx = lev.field(0).set_face(fx_h, d)
        acc = _flat_field(lev, pb, flat_h)
        lev.leptic_accum_div(acc, fx, d, 1.0 / pb.spec.dx[d])
        want = oracle_lib.leptic_accum_div(flat_h if pb.dim == 3 else flat_h[None], fl if pb.dim == 3 else fl[None], d,
                                           1.0 / pb.spec.dx[d])
        assert np.array_equal(_flat_of(pb, acc).reshape(-1), want.reshape(-1))


@pytest.mark.gpu
def test_vertical_solve_rejects_periodic_vertical(somar):
    pb = Problem(n=(16, 16, 8), map_name="cartesian", aniso=1.0, periodic=(0, 0, 1))
    lev = pb.level(somar)
    a, b = lev.field(0), lev.field(0)
    with pytest.raises(somar.SomarError):
        lev.leptic_vertical_solve(a, b, NEUM, DIRI)
    with pytest.raises(somar.SomarError):
        lev.leptic_vertical_solve(a, a, NEUM, DIRI)

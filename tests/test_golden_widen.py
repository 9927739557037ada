"""Committed fixtures tests/golden/widen_golden.npz (made by tests/golden/make_golden_widen.py from the oracle):
CPU -- the oracle still reproduces them bit for bit; GPU -- the CUDA path through the C ABI reproduces them
(bit-exact for diagonal metrics and for the advection / leptic kernels, 1e-13 for the full-metric gradient)."""
import os
import sys

import numpy as np
import pytest

from common import Problem, rel_err

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden_widen as mk  # noqa: E402
from oracle import advect_ref as ar  # noqa: E402

G = np.load(os.path.join(HERE, "golden", "widen_golden.npz"))


def test_oracle_reproduces_fixtures(oracle_lib):
    pb = Problem(**mk.PROJ)
    orc = pb.oracle(oracle_lib)
    for d, g in enumerate(orc.mac_gradient(G["proj/phi"], homogeneous=False)):
        assert np.array_equal(g, G[f"proj/grad{d}"])
    assert np.array_equal(orc.mac_divergence(orc.mac_gradient(G["proj/phi"], homogeneous=True)), G["proj/div"])
    pb = Problem(**mk.LOOSE)
    assert np.array_equal(pb.oracle(oracle_lib, relaxMode=2).relax(G["loose/phi"], G["loose/rhs"], iters=2), G["loose/relax2"])
    pb = Problem(**mk.LEPTIC)
    J = pb.jgup[2].reshape(pb.n[2] + 1, -1)
    dz = pb.spec.dx[2]
    for lo, hi in [(0, 1), (1, 0), (1, 1), (3, 1), (0, 0)]:
        got = oracle_lib.leptic_vertical_solve(G["leptic/rhs"], J, dz, lo, hi, dz_crse=2 * dz)[0]
        assert np.array_equal(got, G[f"leptic/phi_{lo}{hi}"])
    W, J, ov, av, src = mk.advect_inputs()
    dx = [1.0 / n for n in mk.ADV_N]
    for lim in (0, 1):
        for d, f in enumerate(ar.predict_scalar(W, J, ov, av, 0.02, dx, 3, source=src, use_limiting=bool(lim))):
            assert np.array_equal(f, G[f"advect/flux{d}_lim{lim}"])
    for d, f in enumerate(ar.predict_scalar(W, J, ov, av, 0.02, dx, 3, source=src, use_limiting=True, high_order=True)):
        assert np.array_equal(f, G[f"advect/flux{d}_ho"])
    for name in mk.LEPTIC_SOLVE:
        _, rhs, phi, norms = mk.leptic_solve(name)
        assert np.array_equal(rhs, G[f"leptic_solve/{name}_rhs"])
        tol = 0.0 if name == "diri" else 1e-9          # "neum" runs a multigrid solve (OpenMP reductions)
        assert np.abs(phi - G[f"leptic_solve/{name}_phi"]).max() <= tol * np.abs(phi).max()
        assert np.allclose(norms, G[f"leptic_solve/{name}_norms"], rtol=1e-9, atol=0)


@pytest.mark.gpu
def test_cuda_reproduces_fixtures(somar):
    pb = Problem(**mk.PROJ)
    lev = pb.level(somar)
    phi = lev.field(0, G["proj/phi"])
    gr = [lev.field(0) for _ in range(3)]
    lev.mac_gradient(gr, phi, homogeneous=False)
    for d in range(3):
        assert rel_err(gr[d].get_face(d), G[f"proj/grad{d}"]) <= 1e-13
    lev.mac_gradient(gr, phi, homogeneous=True)
    div = lev.field(0)
    lev.mac_divergence(div, gr)
    assert rel_err(div.get(), G["proj/div"]) <= 1e-12
    pb = Problem(**mk.LOOSE)
    lev = pb.level(somar, relaxMode=2)
    f, r = lev.field(0, G["loose/phi"]), lev.field(0, G["loose/rhs"])
    lev.relax(f, r, 2)
    assert np.array_equal(f.get(), G["loose/relax2"])
    pb = Problem(**mk.LEPTIC)
    lev = pb.level(somar)
    dz = pb.spec.dx[2]
    out = lev.field(0)
    for lo, hi in [(0, 1), (1, 0), (1, 1), (3, 1), (0, 0)]:
        rhs = lev.field(0, G["leptic/rhs"].reshape(pb.shape))
        lev.leptic_vertical_solve(out, rhs, lo, hi, dz_crse=2 * dz)
        assert np.array_equal(out.get().reshape(pb.n[2], -1), G[f"leptic/phi_{lo}{hi}"])
    W, J, ov, av, src = mk.advect_inputs()
    pb = Problem(n=mk.ADV_N, map_name="cartesian", aniso=1.0, periodic=(1, 1, 1))
    lev = somar.Level(mk.ADV_N, [1.0 / n for n in mk.ADV_N], dim=3, periodic=(1, 1, 1), is_diagonal=True,
                      params=somar.default_params())
    lev.set_metric_global(pb.jgup, pb.jinv)
    lev.finalize()
    fl = [lev.field(0) for _ in range(3)]
    for lim in (0, 1):
        lev.advect_predict_scalar(fl, lev.field(0, W), [lev.field(0, v) for v in ov], [lev.field(0, v) for v in av],
                                  lev.field(0, J), 0.02, source=lev.field(0, src), use_limiting=bool(lim))
        for d in range(3):
            assert np.array_equal(fl[d].get(), G[f"advect/flux{d}_lim{lim}"])
    lev.advect_predict_scalar(fl, lev.field(0, W), [lev.field(0, v) for v in ov], [lev.field(0, v) for v in av],
                              lev.field(0, J), 0.02, source=lev.field(0, src), use_limiting=True, use_high_order_limiter=True)
    for d in range(3):
        assert np.array_equal(fl[d].get(), G[f"advect/flux{d}_ho"])
    for name, (pk, lk) in mk.LEPTIC_SOLVE.items():
        pb = Problem(**pk)
        lev = pb.level(somar)
        ls = somar.LepticSolver(lev, somar.leptic_default_params(**lk))
        phi, rhs = lev.field(0), lev.field(0, G[f"leptic_solve/{name}_rhs"])
        st = ls.solve(phi, rhs, homogeneous=True)
        want = G[f"leptic_solve/{name}_phi"]
        tol = 1e-13 if name == "diri" else 1e-9
        assert np.abs(phi.get() - want).max() <= tol * np.abs(want).max()
        assert np.allclose(st["res_norms"], G[f"leptic_solve/{name}_norms"], rtol=1e-8, atol=0)

"""Solve parity on the grid shapes BASELINE.json's configs name (VERDICT r1 row +g): the oracle and the CUDA path on the
shipped decks' grids, box layouts, periodicity and AMRMG.* / bottom.* settings.

  1. LockExchange_Cartesian2D   exec/inputs.LockExchange_Cartesian2D.machine:12-15   128 x 64, L = (30, 2)
  2. TaylorGreen                exec/inputs.TaylorGreen.machine:8-17,104             64 x 64, L = (1, 1), periodic, eps 1e-12
                                (Cartesian as shipped, and the twisted map the deck offers: geometry.coordMap = 1)
  3. LockExchange_Cartesian3D   exec/inputs.LockExchange_Cartesian3D.machine:12-17   64 x 96 x 64, L = (15, 3, 2), periodic y,
                                32^3 boxes
  4. LockExchange_DEM_3D        exec/inputs.LockExchange_DEM_3D.machine:14-19        64 x 96 x 64, periodic y, 32 x 16 x 64
                                boxes, terrain-following map (a Gaussian seamount stands in for the HDF5 DEM)
  5. a terrain-following solve at 128 x 128 x 32
"""
import numpy as np
import pytest

from common import Problem

pytestmark = pytest.mark.gpu

# exec/inputs.LockExchange_Cartesian3D.machine:96-117 (the same block in every shipped deck)
DECK = dict(eps=1e-6, relaxMode=1, num_smooth_down=4, num_smooth_up=4, num_smooth_bottom=2, num_smooth_precond=2, numMG=1,
            imin=5, imax=20, hang=1e-15, normThresh=1e-30, maxDepth=-1, bottom_eps=1e-6, bottom_reps=1e-12, bottom_imax=80,
            bottom_numRestarts=5, bottom_hang=1e-15, bottom_small=1e-30, bottom_normType=2)

CONFIGS = {
    "lockexchange2d": (dict(n=(128, 64, 1), dim=2, map_name="cartesian", L=(30.0, 2.0)), DECK),
    "taylorgreen_cartesian": (dict(n=(64, 64, 1), dim=2, map_name="cartesian", L=(1.0, 1.0), periodic=(1, 1, 0)),
                              dict(DECK, eps=1e-12)),
    "taylorgreen_twisted": (dict(n=(64, 64, 1), dim=2, map_name="twisted", L=(1.0, 1.0), periodic=(1, 1, 0),
                                 map_params={"pert": 0.05}), dict(DECK, eps=1e-12)),
    "lockexchange3d": (dict(n=(64, 96, 64), map_name="cartesian", L=(15.0, 3.0, 2.0), periodic=(0, 1, 0), parts=(2, 3, 2)),
                       DECK),
    "lockexchange_dem3d": (dict(n=(64, 96, 64), map_name="terrain", L=(15.0, 3.0, 2.0), periodic=(0, 1, 0), parts=(2, 6, 1),
                                map_params={"amp": 0.3, "width": 0.2}), DECK),
    "terrain_128": (dict(n=(128, 128, 32), map_name="terrain", aniso=8.0, parts=(2, 2, 1), map_params={"amp": 0.2}),
                    dict(DECK, eps=1e-10, imin=0, imax=40)),
    "lockexchange3d_1e-10": (dict(n=(64, 96, 64), map_name="cartesian", L=(15.0, 3.0, 2.0), periodic=(0, 1, 0),
                                  parts=(2, 3, 2)), dict(DECK, eps=1e-10, imin=0, imax=40)),
}


STALLS = {"lockexchange_dem3d"}


@pytest.mark.parametrize("name", list(CONFIGS))
def test_deck_shape_solve_matches_oracle(name, oracle_lib, somar):
    case, knobs = CONFIGS[name]
    pb = Problem(**case)
    rhs = pb.rhs()
    orc = pb.oracle(oracle_lib, **knobs)
    lev = pb.level(somar, **knobs)
    # same hierarchy
    assert lev.ndepths == orc.ndepths
    for d in range(lev.ndepths):
        assert tuple(lev.depth_info(d)["n"]) == tuple(orc.depth_info(d)["n"])
    phi_o, st_o = orc.solve(rhs)
    fphi, frhs = lev.field(0), lev.field(0, rhs)
    st = lev.solve(fphi, frhs)
    assert st["initial_rnorm"] == st_o["initial_rnorm"]
    if name in STALLS:
        # The reference algorithm does not converge on this configuration (the oracle hangs at ~1e-4 with the deck knobs,
        # with bottom.imax 1000 / numRestarts 50 and with LineGSRB alike): once the outer iteration stalls, its history
        # is rounding noise amplified by the Krylov bottom solver.  What is defined: the cycles before the stall (they
        # agree to six digits) and the fact of the stall.
        for a, b in zip(st["history"][:4], st_o["history"][:4]):
            assert abs(a - b) <= 1e-4 * b, (st["history"], st_o["history"])
        assert not (st["exit_status"] & 1) and not (st_o["exit_status"] & 1)
        assert st["final_rnorm"] > knobs["eps"] * st["initial_rnorm"]
        lev.destroy()
        return
    assert abs(st["iters"] - st_o["iters"]) <= 1, (st["iters"], st_o["iters"], st["history"], st_o["history"])
    assert st["exit_status"] == st_o["exit_status"] or abs(st["iters"] - st_o["iters"]) == 1, (st, st_o)
    for a, b in zip(st["history"][:5], st_o["history"][:5]):
        assert abs(a - b) <= 0.02 * b + 1e-9 * st["initial_rnorm"], (st["history"], st_o["history"])
    if st_o["exit_status"] & 1:
        assert st["final_rnorm"] <= knobs["eps"] * st["initial_rnorm"]
    phi = fphi.get()
    w = 1.0 / pb.jinv
    d = (phi - (phi * w).sum() / w.sum()) - (phi_o - (phi_o * w).sum() / w.sum())
    # both are within eps of the discrete solution: a residual of eps * r0 is an error of ~ eps * |phi| * (a few)
    assert np.abs(d).max() <= max(100.0 * knobs["eps"], 1e-9) * np.abs(phi_o).max(), np.abs(d).max() / np.abs(phi_o).max()
    # the operator itself on this shape: bit-exact (diagonal) or 1e-13 (full metric)
    res = lev.field(0)
    lev.residual(res, fphi, frhs)
    want = orc.residual(phi, rhs)
    got = res.get()
    if pb.is_diagonal:
        assert np.array_equal(got, want)
    else:
        assert np.abs(got - want).max() <= 1e-13 * max(np.abs(rhs).max(), np.abs(want).max())
    lev.destroy()

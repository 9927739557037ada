"""Generates tests/golden/widen_golden.npz: seeded inputs and ORACLE outputs for the components added around the
solver (projector calculus, leptic sweeps, advection predictor, LooseGSRB).  Like oracle_golden.npz these pin the
oracle against regressions and give the GPU tests a travel-safe target; they are not reference outputs (the reference
cannot be built here and ships none).      Run:  python tests/golden/make_golden_widen.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from common import Problem  # noqa: E402
from oracle import advect_ref as ar  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

PROJ = dict(n=(12, 8, 6), map_name="terrain", aniso=4.0, parts=(1, 2, 1))
LOOSE = dict(n=(12, 8, 6), map_name="stretched", aniso=4.0, parts=(2, 2, 1))
LEPTIC = dict(n=(8, 6, 10), map_name="stretched", aniso=4.0)
ADV_N = (8, 6, 5)
LEPTIC_SOLVE = {"diri": (dict(n=(16, 12, 8), aniso=32.0, bc_type=[0, 0, 0, 0, 0, 1]), dict(max_order=3)),
                "neum": (dict(n=(16, 16, 8), aniso=32.0), dict(max_order=1))}


def leptic_solve(name):
    from oracle import leptic_ref as lr
    pk, lk = LEPTIC_SOLVE[name]
    pb = Problem(**pk)
    rhs = pb.rhs()
    s = lr.LevelLepticSolver(po, pb.n, pb.spec.dx, pb.jgup, pb.jinv, boxes=pb.boxes, bc_type=pb.bc_type,
                             params=lr.LepticParams(**lk))
    phi = s.solve(np.zeros_like(rhs), rhs, homogeneous=True)
    return pb, rhs, phi, np.array(s.res_norms)


def advect_inputs():
    rng = np.random.default_rng(17)
    shape = (ADV_N[2], ADV_N[1], ADV_N[0])
    W = rng.uniform(-1, 1, shape)
    J = 1.0 + 0.2 * rng.uniform(-1, 1, shape)
    ov = [rng.uniform(-1, 1, shape) for _ in range(3)]
    av = [rng.uniform(-1, 1, shape) for _ in range(3)]
    av[1][..., ::3] = 0.0
    return W, J, ov, av, rng.uniform(-1, 1, shape)


def main():
    po.build()
    out = {}
    pb = Problem(**PROJ)
    orc = pb.oracle(po)
    phi = pb.rand()
    out["proj/phi"] = phi
    for d, g in enumerate(orc.mac_gradient(phi, homogeneous=False)):
        out[f"proj/grad{d}"] = g
    out["proj/div"] = orc.mac_divergence(orc.mac_gradient(phi, homogeneous=True))
    pb = Problem(**LOOSE)
    orc = pb.oracle(po, relaxMode=2)
    phi, rhs = pb.rand(), pb.rand()
    out["loose/phi"], out["loose/rhs"] = phi, rhs
    out["loose/relax2"] = orc.relax(phi, rhs, iters=2)
    pb = Problem(**LEPTIC)
    J = pb.jgup[2].reshape(pb.n[2] + 1, -1)
    rhs = pb.rand().reshape(pb.n[2], -1)
    out["leptic/rhs"] = rhs
    dz = pb.spec.dx[2]
    for lo, hi in [(0, 1), (1, 0), (1, 1), (3, 1), (0, 0)]:
        out[f"leptic/phi_{lo}{hi}"] = po.leptic_vertical_solve(rhs, J, dz, lo, hi, dz_crse=2 * dz)[0]
    W, J, ov, av, src = advect_inputs()
    dx = [1.0 / n for n in ADV_N]
    for lim in (0, 1):
        for d, f in enumerate(ar.predict_scalar(W, J, ov, av, 0.02, dx, 3, source=src, use_limiting=bool(lim))):
            out[f"advect/flux{d}_lim{lim}"] = f
    for d, f in enumerate(ar.predict_scalar(W, J, ov, av, 0.02, dx, 3, source=src, use_limiting=True, high_order=True)):
        out[f"advect/flux{d}_ho"] = f
    for name in LEPTIC_SOLVE:
        _, rhs, phi, norms = leptic_solve(name)
        out[f"leptic_solve/{name}_rhs"], out[f"leptic_solve/{name}_phi"], out[f"leptic_solve/{name}_norms"] = rhs, phi, norms
    np.savez_compressed(os.path.join(HERE, "widen_golden.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()

"""Generates tests/golden/oracle_golden.{npz,json}: small seeded inputs and the oracle's outputs for them.

The reference (UNC-CFD/somar) ships no golden vectors and cannot be built in this container (needs Chombo 3.1,
Fortran, MPI, HDF5), so these fixtures pin the ORACLE against regressions; they are not reference outputs.
Both the CPU tests (oracle vs fixture) and the GPU tests (CUDA path vs fixture) read them.
Run:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from common import Problem  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

CASES = {
    "stretched3d": dict(problem=dict(n=(16, 16, 8), map_name="stretched", aniso=4.0),
                        params=dict(eps=1e-10, num_smooth_down=4, num_smooth_up=4, imin=0, imax=40)),
    "terrain3d_line": dict(problem=dict(n=(16, 16, 8), map_name="terrain", aniso=4.0, map_params={"amp": 0.1}),
                           params=dict(eps=1e-10, num_smooth_down=4, num_smooth_up=4, imin=0, imax=40, relaxMode=3)),
    "lockexchange2d": dict(problem=dict(n=(128, 64, 1), dim=2, map_name="cartesian", aniso=128 * 2.0 / (64 * 30.0)),
                           params=dict(eps=1e-6, num_smooth_down=4, num_smooth_up=4, imin=5, imax=20)),
}


def main():
    po.build()
    out, meta = {}, {"cases": {}}
    for name, case in CASES.items():
        pb = Problem(**case["problem"])
        orc = pb.oracle(po, **case["params"])
        phi, rhs = pb.rand(), pb.rhs()
        out[f"{name}/phi"], out[f"{name}/rhs"] = phi, rhs
        out[f"{name}/apply"] = orc.apply_op(phi)
        out[f"{name}/relax"] = orc.relax(phi, rhs, iters=1)
        out[f"{name}/restrict"] = orc.restrict_residual(phi, rhs)
        sol, st = orc.solve(rhs)
        out[f"{name}/solution"] = sol
        meta["cases"][name] = dict(problem=case["problem"], params=case["params"], iters=st["iters"],
                                   exit_status=st["exit_status"], history=st["history"])
        print(name, st["iters"], st["exit_status"], st["final_rnorm"] / st["initial_rnorm"])
    np.savez_compressed(os.path.join(HERE, "oracle_golden.npz"), **out)
    json.dump(meta, open(os.path.join(HERE, "oracle_golden.json"), "w"), indent=1)


if __name__ == "__main__":
    main()

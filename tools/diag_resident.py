"""mode 1 / 2 / 3 bottom solvers and the oracle on the small test problems: V-cycles, exit status, histories"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from common import Problem
from somar_b200 import capi
from oracle import pyoracle as po
SOLVE = dict(eps=1e-10, num_smooth_down=4, num_smooth_up=4, num_smooth_bottom=2, imin=0, imax=40)
cases = [dict(n=(64, 64, 16), map_name="stretched", aniso=16.0), dict(n=(48, 40, 16), map_name="stretched", aniso=16.0, parts=(1, 2, 1))]
extras = [{}, dict(bottom_imax=1000, bottom_numRestarts=50), dict(bottom_imax=1000, bottom_numRestarts=50, bottom_hang=0.9),
          dict(bottom_imax=1000, bottom_numRestarts=50, bottom_normType=0)]
for case in cases:
    pb = Problem(**case)
    rhs = pb.rhs()
    for extra in extras:
        prm = dict(SOLVE, **extra)
        print("==", case["n"], extra)
        for variant in ("1", "2", "3"):
            os.environ["SOMAR_B200_BOTTOM"] = variant
            lev = pb.level(capi, **prm)
            fphi, frhs = lev.field(0), lev.field(0, rhs)
            try:
                st = lev.solve(fphi, frhs)
                print(f" mode {variant}: iters {st['iters']} exit {st['exit_status']} bottom {st['bottom_iters_total']} launches {st['launches']}",
                      " ".join(f"{h / st['history'][0]:.2e}" for h in st["history"]))
            except capi.SomarError as e:
                print(f" mode {variant}: {e}")
            lev.destroy()
        _, so = pb.oracle(po, **prm).solve(rhs)
        print(f" oracle: iters {so['iters']} exit {so['exit_status']} bottom {so['bottom_iters_total']}",
              " ".join(f"{h / so['history'][0]:.2e}" for h in so["history"]))

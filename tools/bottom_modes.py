"""Development check: full solves with the three bottom-solver modes side by side (histories, bottom iterations)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from common import Problem  # noqa: E402
from somar_b200 import capi  # noqa: E402

SOLVE = dict(eps=1e-10, num_smooth_down=4, num_smooth_up=4, num_smooth_bottom=2, imin=0, imax=40, relaxMode=1)
cases = [dict(n=(64, 64, 16), map_name="stretched", aniso=16.0), dict(n=(64, 32, 1), dim=2, map_name="stretched", aniso=4.0),
         dict(n=(128, 128, 32), map_name="stretched", aniso=16.0)]
for case in cases:
    pb = Problem(**case)
    rhs = pb.rhs()
    ref = None
    for mode in ("0", "2"):
        os.environ["SOMAR_B200_BOTTOM"] = mode
        lev = pb.level(capi, **SOLVE)
        fphi, frhs = lev.field(0), lev.field(0, rhs)
        st = lev.solve(fphi, frhs)
        phi = fphi.get()
        print(case["n"], "mode", mode, "iters", st["iters"], "exit", st["exit_status"], "bottom", st["bottom_iters_total"],
              "launches", st["launches"])
        print("   ", " ".join(f"{h:.4e}" for h in st["history"]))
        if ref is None:
            ref = phi
        else:
            print("    max diff vs mode 0:", np.abs((phi - phi.mean()) - (ref - ref.mean())).max() / np.abs(ref).max())
        lev.destroy()

"""Which of the reference's own solver knobs make the headline-family solve reach 1e-10?  (CPU oracle only.)

The knobs are the reference's: AMRMG.num_smooth_precond, bottom.imax, bottom.numRestarts, AMRMG.num_smooth_bottom,
AMRMG.relax_mode (src/utils/ProblemContext.cpp:1147-1216, src/projection/AMRPressureSolver.cpp:253-262).
usage: python tools/converge_study.py NX NY NZ  key=value[,key=value...] ...

Round-2 findings (8 y-slab boxes, stretched, dx/dz = 16, pre/post/bottom 4/4/2, eps 1e-10), V-cycles / exit / residual:
  256x256x64  deck (bottom.imax 80, numRestarts 5)                 7 / hang / 1.1e-3     231 bottom iterations
              bottom.imax 1000, numRestarts 50                    15 / conv / 6.1e-11   1930
              num_smooth_precond 8, bottom.imax 400               15 / conv / 6.1e-11    594   (3x the preconditioner work)
              num_smooth_precond 4, bottom.imax 400, restarts 20  15 / conv / 6.1e-11   1310
              num_smooth_precond 16, bottom.imax 200              15 / conv / 6.1e-11    797
              bottom.imax 200, numRestarts 50                     16 / conv / 2.0e-11   3199
              bottom.imax 100, numRestarts 50                     17 / conv / 9.6e-11   1359
              bottom.imax 300, numRestarts 20                     18 / conv / 2.5e-11   1320
              num_smooth_precond 1, imax 1000, restarts 50        20 / conv / 8.3e-11   4485
              num_smooth_precond 0, imax 1000, restarts 50         8 / hang / 7.3e-4
  128x128x32  deck 7 / hang / 3.9e-4;  imax 1000 + restarts 50: 14 / conv;  precond 8 + imax 400: 14 / conv
=> the stall is the 256x256x4 bottom problem; once the bottom is solved the outer rate is ~0.2 per V-cycle whatever
   knob solves it.  bench.py uses bottom.imax 1000 + numRestarts 50 (the mildest change: the two bottom.* knobs only).
"""
import sys
import time
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from oracle import pyoracle as po
from somar_b200 import synthetic as syn


def main():
    grid = tuple(int(x) for x in sys.argv[1:4])
    runs = sys.argv[4:]
    aniso, map_name, nboxes = 16.0, "stretched", 8
    spec = syn.headline_grid(n=grid, aniso=aniso, map_name=map_name)
    m = syn.build_metric(spec)
    jg = [np.ascontiguousarray(j[0].numpy()) if m.ncomp == 1 else j.numpy() for j in m.jgup]
    rhs = syn.random_rhs(spec, m.jinv).numpy()
    boxes = syn.split_boxes(spec.n, (1, nboxes, 1))
    base = dict(eps=1e-10, num_smooth_down=4, num_smooth_up=4, num_smooth_bottom=2, imin=0, imax=50)
    for r in runs:
        kw = dict(base)
        for kv in r.split(","):
            if not kv:
                continue
            k, v = kv.split("=")
            kw[k] = float(v) if ("." in v or "e" in v) else int(v)
        orc = po.Oracle(spec.n, spec.dx, jg, m.jinv.numpy(), dim=3, is_diagonal=spec.is_diagonal, boxes=boxes,
                        params=po.default_params(**kw))
        t = time.time()
        phi, st = orc.solve(rhs)
        dt = time.time() - t
        h = st["history"]
        print(f"{r or 'deck'}: depths {orc.ndepths} bottom {orc.depth_info(orc.ndepths - 1)['n']} iters {st['iters']} "
              f"exit {st['exit_status']} rel {st['final_rnorm'] / st['initial_rnorm']:.3e} bottom_iters {st['bottom_iters_total']} "
              f"{dt:.1f}s", flush=True)
        print("   hist", " ".join(f"{x / h[0]:.2e}" for x in h), flush=True)


if __name__ == "__main__":
    main()

"""one terrain-following (full-metric) 512x512x128 solve, as bench.py's terrain block runs it; exits 0 whatever the solver's
exit status (the terrain family stalls at 2.4e-7 on both stacks), so that ncu can be pointed at it"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from somar_b200 import capi, synthetic as syn

n = (512, 512, 128)
spec = syn.headline_grid(n=n, aniso=16.0, map_name="terrain")
boxes = syn.split_boxes(n, (1, 8, 1))
lev = capi.Level(n, spec.dx, dim=3, is_diagonal=False, boxes=boxes, box_rank=[0] * 8, params=capi.default_params(), device=0)
m = syn.build_metric(spec, device="cuda")
lev.set_metric_device(m.jgup, m.jinv, m.ncomp)
lev.finalize()
phi, rhs = lev.field(0), lev.field(0)
r = torch.rand(n[2], n[1], n[0], dtype=torch.float64, device="cuda")
r -= r.mean()
rhs.set(r)
l0 = lev.launches()
t0 = time.time()
try:
    st = lev.solve(phi, rhs)
    print({k: st[k] for k in st if k != "history"})
except Exception as e:  # noqa: BLE001
    print("solve raised", e)
lev.sync()
print("solve %.1f ms, %d launches" % (1e3 * (time.time() - t0), lev.launches() - l0))

"""relaxes a terrain-following (full-metric) 512x512x128 level a few times and times the colour pass: the command the ncu
capture of k_gsrb_full runs (profiles/r2_ncu_k_gsrb_full.txt)"""
import sys, os
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
s = torch.cuda.ExternalStream(lev.stream())
phi, rhs = lev.field(0), lev.field(0)
phi.set(torch.rand(n[2], n[1], n[0], dtype=torch.float64, device="cuda"))
rhs.set(torch.rand(n[2], n[1], n[0], dtype=torch.float64, device="cuda"))
lev.relax(phi, rhs, iters=1)
lev.sync()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(s):
    a.record(s)
    lev.relax(phi, rhs, iters=4)
    b.record(s)
lev.sync()
t = a.elapsed_time(b) / 8
cells = n[0] * n[1] * n[2]
print("k_gsrb_full colour pass (incl. its exchange) %.3f ms  %.0f GB/s by the 104 B/cell model" % (t, 104 * cells / t / 1e6))

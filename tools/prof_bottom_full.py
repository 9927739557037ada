"""times the operations one BiCGStab iteration of the full-metric (terrain) bottom solve is made of, on the bottom depth of
the 512x512x128 terrain level: apply, one relax iteration (two colour passes), the preconditioner, a dot product.
Exits 0; prints one line per operation (CUDA events on the level's stream, 200 repetitions each)."""
import os
import sys

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
for d in range(lev.ndepths):
    print("depth", d, lev.depth_info(d))


REPS = int(os.environ.get("PROF_REPS", "0"))


def timed(what, fn, reps=200):
    reps = REPS or reps
    fn()
    lev.sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = lev.launches()
    with torch.cuda.stream(s):
        a.record(s)
        for _ in range(reps):
            fn()
        b.record(s)
    lev.sync()
    print("%-28s %8.1f us  (%d launches)" % (what, 1e3 * a.elapsed_time(b) / reps, (lev.launches() - l0) // reps))


for d in (lev.ndepths - 1, lev.ndepths - 2, 0):
    shp = lev.local_shape(d)
    print("depth", d, "shape", shp)
    x, y, z = lev.field(d), lev.field(d), lev.field(d)
    x.set(torch.rand(*shp, dtype=torch.float64, device="cuda"))
    y.set(torch.rand(*shp, dtype=torch.float64, device="cuda"))
    reps = 200 if d > 0 else 5
    timed("apply_op", lambda: lev.apply_op(z, x), reps)
    timed("relax x1 (2 colour passes)", lambda: lev.relax(x, y, iters=1), reps)
    timed("precond", lambda: lev.precond(z, y), reps)
    timed("incr (axpy)", lambda: lev.incr(z, y, 0.5), reps)

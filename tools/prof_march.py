"""times the depth-0 residual (k_march<1>), fused restrictResidual (k_march<2>) and prolongIncrement at the headline size"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from somar_b200 import capi, synthetic as syn

n = (1024, 1024, 256)
spec = syn.headline_grid(n=n, aniso=16.0)
lev = capi.Level(spec.n, spec.dx, dim=3, is_diagonal=True, params=capi.default_params(), device=0)
m = syn.build_metric(spec, device="cuda")
lev.set_metric_device([j[0].contiguous() for j in m.jgup], m.jinv, 1)
lev.finalize()
phi, rhs, res, crse = lev.field(0), lev.field(0), lev.field(0), lev.field(1)
rhs.set(torch.rand(n[2], n[1], n[0], dtype=torch.float64, device="cuda"))
phi.set(torch.rand(n[2], n[1], n[0], dtype=torch.float64, device="cuda"))
s = torch.cuda.ExternalStream(lev.stream())
def timeit(fn, reps=5):
    fn(); lev.sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        a.record(s)
        for _ in range(reps): fn()
        b.record(s)
    lev.sync()
    return a.elapsed_time(b) / reps
print("residual %.3f ms  restrict %.3f ms  prolong %.3f ms" % (
    timeit(lambda: lev.residual(res, phi, rhs)), timeit(lambda: lev.restrict_residual(crse, phi, rhs)),
    timeit(lambda: lev.prolong_increment(phi, crse))))

"""times the advection predictor on a fully periodic 512^3 level (device-resident synthetic fields)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from somar_b200 import capi, synthetic as syn

n = (512, 512, 512)
cells = n[0] * n[1] * n[2]
spec = syn.GridSpec(n=n, L=(1.0, 1.0, 1.0), dim=3, periodic=(True, True, True), map_name="cartesian", map_params={})
lev = capi.Level(spec.n, spec.dx, dim=3, periodic=(1, 1, 1), is_diagonal=True, params=capi.default_params(), device=0)
m = syn.build_metric(spec, device="cuda")
lev.set_metric_device([j[0].contiguous() for j in m.jgup], m.jinv, 1)
lev.finalize()
s = torch.cuda.ExternalStream(lev.stream())
rnd = lambda: torch.rand(n[2], n[1], n[0], dtype=torch.float64, device="cuda")
W, J = lev.field(0), lev.field(0)
W.set(rnd()); J.set(1.0 + 0.1 * rnd())
ov = [lev.field(0) for _ in range(3)]
av = [lev.field(0) for _ in range(3)]
for f in ov + av:
    f.set(rnd() - 0.5)
out = [lev.field(0) for _ in range(3)]
for lim, ho in ((False, False), (True, False), (True, True)):
    fn = lambda: lev.advect_predict_scalar(out, W, ov, av, J, 1e-3, use_limiting=lim, use_high_order_limiter=ho)
    fn(); lev.sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        a.record(s)
        for _ in range(3):
            fn()
        b.record(s)
    lev.sync()
    t = a.elapsed_time(b) / 3
    print("predict_scalar 512^3 limiting=%d high_order=%d  %7.3f ms  %7.1f Mcell/s" % (lim, ho, t, cells / t / 1e3))

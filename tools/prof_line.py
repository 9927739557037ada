"""ncu target: a few LineGSRB iterations at the headline size (relax_mode 3), nothing else."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from somar_b200 import capi, synthetic as syn

n = tuple(int(x) for x in sys.argv[1:4]) if len(sys.argv) > 3 else (1024, 1024, 256)
spec = syn.headline_grid(n=n, aniso=16.0)
lev = capi.Level(spec.n, spec.dx, dim=3, is_diagonal=True, params=capi.default_params(relaxMode=3), device=0)
m = syn.build_metric(spec, device="cuda")
lev.set_metric_device([j[0].contiguous() for j in m.jgup], m.jinv, 1)
lev.finalize()
phi, rhs = lev.field(0), lev.field(0)
rhs.set(torch.rand(n[2], n[1], n[0], dtype=torch.float64, device="cuda"))
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
lev.relax(phi, rhs, 2)
lev.sync()
s = torch.cuda.ExternalStream(lev.stream())
with torch.cuda.stream(s):
    ev[0].record(s)
    lev.relax(phi, rhs, 4)
    ev[1].record(s)
lev.sync()
print("line relax: %.3f ms per iteration (2 colour passes)" % (ev[0].elapsed_time(ev[1]) / 4))

"""Development probe: time ONE bottom visit (relaxations + BiCGStab) on a level whose only depth has the size of the
headline problem's bottom depth (256 x 256 x 4).  usage: python tools/bottom_probe.py [reps]
Honours SOMAR_B200_BOTTOM (0 host-sequenced, 1 device scalars, 2 persistent kernel)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from somar_b200 import capi, synthetic as syn  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
n = (256, 256, 4)
spec = syn.headline_grid(n=n, aniso=1.0)
m = syn.build_metric(spec, device="cuda")
prm = capi.default_params(relaxMode=1, maxDepth=0, eps=1e-10, num_smooth_down=4, num_smooth_up=4, num_smooth_bottom=2,
                          bottom_imax=80)
lev = capi.Level(n, spec.dx, dim=3, is_diagonal=True, params=prm)
lev.set_metric_device([j[0].contiguous() for j in m.jgup], m.jinv, 1)
lev.finalize()
assert lev.ndepths == 1
gen = torch.Generator(device="cuda")
gen.manual_seed(7)
rhs = torch.rand((n[2], n[1], n[0]), dtype=torch.float64, device="cuda", generator=gen) * 2 - 1
w = 1.0 / m.jinv
rhs -= (rhs * w).sum() / w.sum()
fr, fe = lev.field(0, rhs), lev.field(0)
stream = torch.cuda.ExternalStream(lev.stream())
zero = torch.zeros_like(rhs)
for it in range(reps + 2):
    fe.set(zero)
    torch.cuda.synchronize()
    l0 = lev.launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    lev.mg_cycle(fe, fr)
    e1.record(stream)
    e1.synchronize()
    if it >= 2:
        print(f"mode {os.environ.get('SOMAR_B200_BOTTOM', '2')}: bottom visit {e0.elapsed_time(e1):8.3f} ms, "
              f"{lev.launches() - l0} launches")

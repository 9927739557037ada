"""times the projector / leptic / advection kernels at a headline-class size (device-resident synthetic fields)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from somar_b200 import capi, synthetic as syn


def timeit(lev, s, fn, reps=3):
    fn(); lev.sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        a.record(s)
        for _ in range(reps):
            fn()
        b.record(s)
    lev.sync()
    return a.elapsed_time(b) / reps


def rnd(n):
    return torch.rand(n[2], n[1], n[0], dtype=torch.float64, device="cuda")


n = (1024, 1024, 256)
cells = n[0] * n[1] * n[2]
spec = syn.headline_grid(n=n, aniso=16.0)
lev = capi.Level(spec.n, spec.dx, dim=3, is_diagonal=True, params=capi.default_params(), device=0)
m = syn.build_metric(spec, device="cuda")
lev.set_metric_device([j[0].contiguous() for j in m.jgup], m.jinv, 1)
lev.finalize()
s = torch.cuda.ExternalStream(lev.stream())
phi, rhs, div = lev.field(0), lev.field(0), lev.field(0)
phi.set(rnd(n)); rhs.set(rnd(n))
g = [lev.field(0) for _ in range(3)]
t = timeit(lev, s, lambda: lev.mac_gradient(g, phi))
print("mac_gradient      %7.3f ms  %6.0f GB/s by 56 B/cell (phi + 3 Jg + 3 out)" % (t, 56 * cells / t / 1e6))
t = timeit(lev, s, lambda: lev.mac_divergence(div, g))
print("mac_divergence    %7.3f ms  %6.0f GB/s by 40 B/cell" % (t, 40 * cells / t / 1e6))
t = timeit(lev, s, lambda: lev.mac_apply_correction(g, phi, 0.1))
print("apply_correction  %7.3f ms  %6.0f GB/s by 80 B/cell" % (t, 80 * cells / t / 1e6))
t = timeit(lev, s, lambda: lev.leptic_vertical_solve(phi, rhs, 0, 1))
print("leptic LDL^T      %7.3f ms  %6.0f GB/s by 72 B/cell" % (t, 72 * cells / t / 1e6))
t = timeit(lev, s, lambda: lev.leptic_vertical_solve(phi, rhs, 0, 0))
print("leptic NN         %7.3f ms  %6.0f GB/s by 80 B/cell" % (t, 80 * cells / t / 1e6))
t = timeit(lev, s, lambda: lev.leptic_vert_integral(div, rhs, 0.1))
print("vert_integral     %7.3f ms  %6.0f GB/s by 8 B/cell" % (t, 8 * cells / t / 1e6))
t = timeit(lev, s, lambda: lev.leptic_add_extrusion(phi, div))
print("add_extrusion     %7.3f ms  %6.0f GB/s by 16 B/cell" % (t, 16 * cells / t / 1e6))
lev.destroy()
del lev, phi, rhs, div, g
torch.cuda.empty_cache()

# advection: fully periodic cartesian level, 512^3 (15 scratch fields + 12 user fields of 1.1 GB)
n = (512, 512, 512)
cells = n[0] * n[1] * n[2]
spec = syn.GridSpec(n=n, L=(1.0, 1.0, 1.0), dim=3, periodic=(True, True, True), map_name="cartesian", map_params={})
lev = capi.Level(spec.n, spec.dx, dim=3, periodic=(1, 1, 1), is_diagonal=True, params=capi.default_params(), device=0)
m = syn.build_metric(spec, device="cuda")
lev.set_metric_device([j[0].contiguous() for j in m.jgup], m.jinv, 1)
lev.finalize()
s = torch.cuda.ExternalStream(lev.stream())
W, J = lev.field(0), lev.field(0)
W.set(rnd(n)); J.set(1.0 + 0.1 * rnd(n))
ov = [lev.field(0) for _ in range(3)]
av = [lev.field(0) for _ in range(3)]
for f in ov + av:
    f.set(rnd(n) - 0.5)
out = [lev.field(0) for _ in range(3)]
new = lev.field(0)
for lim in (False, True):
    t = timeit(lev, s, lambda: lev.advect_predict_scalar(out, W, ov, av, J, 1e-3, use_limiting=lim))
    print("predict_scalar 512^3 limiting=%d  %7.3f ms  %6.1f Mcell/s" % (lim, t, cells / t / 1e3))
t = timeit(lev, s, lambda: lev.advect_update(new, W, out, 1e-3))
print("advect_update 512^3  %7.3f ms  %6.0f GB/s by 48 B/cell" % (t, 48 * cells / t / 1e6))

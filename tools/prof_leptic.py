"""times the leptic solver (LevelLepticSolver) on a thin 1024 x 1024 x 64 domain (dx / H = 4), device-resident inputs"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from somar_b200 import capi, synthetic as syn

n = (1024, 1024, 64)
cells = n[0] * n[1] * n[2]
spec = syn.headline_grid(n=n, aniso=256.0)
m = syn.build_metric(spec, device="cuda")
rhs_t = syn.random_rhs(spec, m.jinv, device="cuda")
torch.cuda.synchronize()
for name, bc, kw in [("dirichlet top (vertical solves only)", [0, 0, 0, 0, 0, 1], {}),
                     ("all Neumann (excess + horizontal MG + fall-back MG)", [0] * 6, {}),
                     ("all Neumann, max_order 1 (no fall-back)", [0] * 6, {"max_order": 1})]:
    lev = capi.Level(spec.n, spec.dx, dim=3, is_diagonal=True, bc_type=bc, params=capi.default_params(), device=0)
    lev.set_metric_device([j[0].contiguous() for j in m.jgup], m.jinv, 1)
    lev.finalize()
    ls = capi.LepticSolver(lev, capi.leptic_default_params(**kw))
    phi, rhs = lev.field(0), lev.field(0)
    rhs.set(rhs_t)
    lev.sync()
    best = None
    for rep in range(3):
        lev.set_to_zero(phi)
        st = ls.solve(phi, rhs, homogeneous=True)
        if best is None or st["solve_ms"] < best["solve_ms"]:
            best = st
    nr = best["res_norms"]
    print(json.dumps({"case": name, "n": n, "dx_over_H": spec.dx[0] / (spec.dx[2] * n[2]), "solve_ms": round(best["solve_ms"], 2),
                      "orders": best["orders"], "exit_status": best["exit_status"], "horiz_solves": best["horiz_solves"],
                      "horiz_vcycles": best["horiz_iters_total"], "fallback_used": best["krylov_used"],
                      "launches": best["launches"], "rel_res_norms": ["%.2e" % (x / nr[0]) for x in nr],
                      "Mcell_orders_per_s": round(cells * best["orders"] / best["solve_ms"] / 1e3)}))
    ls.destroy(); lev.destroy()
    del ls, lev, phi, rhs
    torch.cuda.empty_cache()

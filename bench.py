#!/usr/bin/env python
"""bench.py -- fp64 mapped-Poisson pressure solve (semicoarsening multigrid V-cycles) on B200, BASELINE.json metric.

A "step" is ONE pressure solve of the synthetic stratified, stretched-grid problem: V-cycles until
the max-norm residual has dropped by 1e-10, pre/post/bottom smooths 4/4/2, LevelGSRB (the shipped decks' settings).
    value  = N_fine * (#V-cycles) / t_solve   [Mcell-updates/s], whole job, inputs resident in HBM
    e2e    = same metric through the C ABI with HOST buffers: pinned rhs -> device, solve, phi -> host
Strong scaling: the 1024 x 1024 x 256 grid is always cut into the same 8 y-slab boxes, spread over the N ranks.
`--impl reference` times the reference's own algorithm on the host cores (the CPU oracle; the reference itself
needs Chombo/Fortran/MPI and cannot be built here), on a bounded sample of the workload.
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SOLVE = dict(eps=1e-10, num_smooth_down=4, num_smooth_up=4, num_smooth_bottom=2, imin=0, imax=50)
NBOXES = 8
CPU_SAMPLE_GRID = (256, 256, 64)
CPU_SAMPLE_CYCLES = 2


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--grid", type=int, nargs=3, default=[1024, 1024, 256])
    ap.add_argument("--aniso", type=float, default=16.0)
    ap.add_argument("--map", default="stretched", choices=["stretched", "terrain", "cartesian"])
    ap.add_argument("--relax", type=int, default=1)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------------------
def cpu_reference_run(grid, aniso, map_name, relax, cycles, threads=None):
    """Times the oracle (port of the reference's CPU path) on `grid`, capped at `cycles` V-cycles."""
    import numpy as np
    from oracle import pyoracle as po
    from somar_b200 import synthetic as syn
    po.build()
    if threads:
        po.set_threads(threads)
    spec = syn.headline_grid(n=tuple(grid), aniso=aniso, map_name=map_name)
    m = syn.build_metric(spec)
    jg = [np.ascontiguousarray(j[0].numpy()) if m.ncomp == 1 else j.numpy() for j in m.jgup]
    rhs = syn.random_rhs(spec, m.jinv).numpy()
    # one y-slab box per host thread (the reference runs one MPI rank per core, each owning boxes); the oracle's
    # OpenMP loops run over boxes
    nb = 1
    while nb * 2 <= po.max_threads() and grid[1] // (nb * 2) >= 16:
        nb *= 2
    boxes = syn.split_boxes(spec.n, (1, nb, 1))
    orc = po.Oracle(spec.n, spec.dx, jg, m.jinv.numpy(), dim=3, is_diagonal=spec.is_diagonal, boxes=boxes,
                    params=po.default_params(relaxMode=relax, **SOLVE))
    secs, st = orc.timed_solve(rhs, ncycles_cap=cycles)
    cells = grid[0] * grid[1] * grid[2]
    return {"secs": secs, "iters": st["iters"], "mcups": cells * st["iters"] / secs / 1e6,
            "cores": min(nb, po.max_threads()), "boxes": nb}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    for i in range(a.warmup + a.steps):
        r = cpu_reference_run(CPU_SAMPLE_GRID, a.aniso, a.map, a.relax, CPU_SAMPLE_CYCLES)
        if i >= a.warmup:
            vals.append(r)
    secs = sum(v["secs"] for v in vals) / len(vals)
    mc = sum(v["mcups"] for v in vals) / len(vals)
    sample = (f"{CPU_SAMPLE_GRID[0]}x{CPU_SAMPLE_GRID[1]}x{CPU_SAMPLE_GRID[2]} {a.map} grid, aniso {a.aniso:g}, "
              f"{CPU_SAMPLE_CYCLES} V-cycles per step, OpenMP over {vals[0]['boxes']} y-slab boxes")
    line = {"impl": "reference", "metric": "fp64 Poisson V-cycle Mcell-updates/s", "value": mc, "unit": "Mcell-updates/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": secs * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(a, None),
            "cpu_baseline": {"value": mc, "unit": "Mcell-updates/s", "cores": vals[0]["cores"], "kind": "port",
                             "sample": sample},
            "e2e": {"value": mc, "unit": "Mcell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def workload_config(a, extra):
    c = {"workload": f"synthetic stratified stretched-grid Poisson solve {a.grid[0]}x{a.grid[1]}x{a.grid[2]} "
                     f"({a.map}, dx/dz={a.aniso:g}, homogeneous Neumann, eps 1e-10, pre/post/bottom 4/4/2, "
                     f"relax_mode {a.relax})",
         "grid": list(a.grid), "boxes": NBOXES, "partition": f"y-slabs over {a.gpus} GPU(s)",
         "l2_policy": "inputs larger than L2 (2.2 GB per field at depth 0 vs 126 MB L2)"}
    if extra:
        c.update(extra)
    return c


# ------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.path = f"/tmp/somar_clocks_{os.getpid()}.csv"
        self.f = open(self.path, "w")
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.close()
        sm, mx, reasons = [], None, set()
        for ln in open(self.path):
            t = [x.strip() for x in ln.split(",")]
            if len(t) < 9:
                continue
            try:
                sm.append(float(t[1]))
                mx = float(t[2])
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], t[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def byte_model(schedule, diagonal, relax, pre, post):
    """Algorithmic (compulsory) HBM bytes per V-cycle, SURVEY.md 8(d) / DESIGN.md: per cell of each depth."""
    gsrb_iter = (96 if relax == 1 else 88) if diagonal else (208 if relax == 1 else 200)
    resid = 56 if diagonal else 104
    total = 0.0
    for lv, s in enumerate(schedule[:-1]):
        r = s["ratio"][0] * s["ratio"][1] * s["ratio"][2]
        cells = s["n"][0] * s["n"][1] * s["n"][2]
        per_cell = (pre + post) * gsrb_iter + (resid + 8.0 / r) + (24 + 8.0 / r) + 8
        total += cells * per_cell
    n0 = schedule[0]["n"][0] * schedule[0]["n"][1] * schedule[0]["n"][2]
    total += n0 * (resid + 24 + 8)
    return total


def run_ours(a):
    import numpy as np
    import torch
    import torch.distributed as dist
    from somar_b200 import capi, synthetic as syn

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    comm_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        buf = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            buf.copy_(torch.tensor(list(capi.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(buf, 0)
        comm_id = bytes(buf.cpu().tolist())
    assert NBOXES % world == 0, "--gpus must divide 8"

    grid = tuple(a.grid)
    spec = syn.headline_grid(n=grid, aniso=a.aniso, map_name=a.map)
    boxes = syn.split_boxes(grid, (1, NBOXES, 1))
    box_rank = [b * world // NBOXES for b in range(NBOXES)]
    mine = [b for b in range(NBOXES) if box_rank[b] == rank]
    j0, j1 = boxes[mine[0]][0][1], boxes[mine[-1]][1][1] + 1
    params = capi.default_params(relaxMode=a.relax, **SOLVE)
    lev = capi.Level(grid, spec.dx, dim=3, is_diagonal=spec.is_diagonal, boxes=boxes, box_rank=box_rank, params=params,
                     device=local, rank=rank, nranks=world, comm_id=comm_id)
    m = syn.build_metric(spec, device=dev, yrange=(j0, j1))
    if m.ncomp == 1:
        lev.set_metric_device([j[0].contiguous() for j in m.jgup], m.jinv, 1)
    else:
        lev.set_metric_device(m.jgup, m.jinv, m.ncomp)
    lev.finalize()
    # rhs: seeded global field, zero dV-mean (every rank draws the same numbers and keeps its slab)
    gen = torch.Generator(device=dev)
    gen.manual_seed(12345)
    w = 1.0 / m.jinv
    parts = []
    for b in range(NBOXES):   # box by box so that no rank ever holds the global field
        r = torch.rand((grid[2], boxes[b][1][1] - boxes[b][0][1] + 1, grid[0]), dtype=torch.float64, device=dev,
                       generator=gen) * 2.0 - 1.0
        if b in mine:
            parts.append(r)
        del r
    rhs = torch.cat(parts, 1).contiguous()
    del parts
    sums = torch.stack([(rhs * w).sum(), w.sum()])
    if world > 1:
        dist.all_reduce(sums)
    rhs -= sums[0] / sums[1]
    del w, m
    torch.cuda.empty_cache()

    frhs, fphi = lev.field(0, rhs), lev.field(0)
    h_rhs = rhs.cpu().pin_memory()
    h_phi = torch.empty_like(h_rhs).pin_memory()
    del rhs
    stream = torch.cuda.ExternalStream(lev.stream(), device=dev)
    schedule = [lev.depth_info(d) for d in range(lev.ndepths)]
    n_fine = grid[0] * grid[1] * grid[2]
    local_cells = grid[0] * (j1 - j0) * grid[2]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        out = [fn() for _ in range(steps)]
        e1.record(stream)
        e1.synchronize()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), out

    def solve_resident():
        return lev.solve(fphi, frhs)

    def solve_e2e():
        frhs.set(h_rhs)                      # pinned host -> device (inside the timed region)
        st = lev.solve(fphi, frhs)
        fphi.get(h_phi)                      # device -> pinned host (synchronises)
        return st

    for _ in range(a.warmup):
        st = solve_resident()
    lev.profile(True)
    lev.profile_read()
    clocks = ClockSampler(local) if rank == 0 else None
    ms, stats = timed(solve_resident, a.steps)
    clk = clocks.stop() if clocks else None
    prof = lev.profile_read()
    lev.profile(False)
    iters = sum(s["iters"] for s in stats)
    launches = sum(s["launches"] for s in stats)
    value = n_fine * iters / (ms * 1e-3) / 1e6

    e2e = None
    if not a.no_e2e:
        solve_e2e()
        ms_e, stats_e = timed(solve_e2e, a.steps)
        it_e = sum(s["iters"] for s in stats_e)
        e2e = {"value": n_fine * it_e / (ms_e * 1e-3) / 1e6, "unit": "Mcell-updates/s",
               "h2d_bytes_per_step": int(h_rhs.numel() * 8 * world), "d2h_bytes_per_step": int(h_phi.numel() * 8 * world),
               "ms_per_step": ms_e / a.steps}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = {}
    pk_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk_path):
        peaks = json.load(open(pk_path))
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    k = prof["smooth0_kernel"]
    # algorithmic bytes per cell one launch of the depth-0 smoother kernel moves (a red+black GSRB
    # iteration is 96 B/cell diagonal, 208 full; LineGSRB 88 / 200; a colour pass is half an iteration)
    kname = lev.relax_kernel(0, a.relax)
    per_cell = {"k_gsrb_ortho": 48, "k_gsrb_fused": 96, "k_gsrb_async": 96, "k_gsrb_full": 104,
                "k_line_gsrb": 44 if spec.is_diagonal else 100}[kname]
    kdesc = {"k_gsrb_ortho": "k_gsrb_ortho<3> (depth-0 red/black colour pass, 2 launches per iteration)",
             "k_gsrb_fused": "k_gsrb_fused (depth-0 red+black iteration in one sweep)",
             "k_gsrb_async": "k_gsrb_async (depth-0 red+black iteration in one sweep, producer/consumer pipeline)",
             "k_gsrb_full": "k_gsrb_full<3> (depth-0 colour pass, full metric)",
             "k_line_gsrb": "k_line_gsrb<3> (depth-0 vertical-line colour pass)"}[kname]
    alg_bytes = per_cell * local_cells
    avg_ms = k["ms"] / max(k["count"], 1)
    achieved = alg_bytes / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        key = f"{kname}_{a.map}_{grid[0]}x{grid[1]}x{grid[2]}_gpus{world}"
        traffic = tj.get(key, {}).get("dram_bytes_per_launch")
    bv = byte_model(schedule, spec.is_diagonal, a.relax, SOLVE["num_smooth_down"], SOLVE["num_smooth_up"])
    vcycle_gbs = bv * iters / (ms * 1e-3) / 1e9
    line = {"metric": "fp64 Poisson V-cycle Mcell-updates/s", "value": value, "unit": "Mcell-updates/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(a, {
                "halo_path": lev.halo_path(),
                "vcycles_per_solve": iters / a.steps, "final_rel_residual": stats[-1]["final_rnorm"] / stats[-1]["initial_rnorm"],
                "exit_status": stats[-1]["exit_status"], "bottom_iters_per_solve": stats[-1]["bottom_iters_total"],
                "schedule": [{"n": list(s["n"]), "ratio": list(s["ratio"])} for s in schedule],
                "model_bytes_per_vcycle": bv, "vcycle_model_gbs": vcycle_gbs,
                "vcycle_hbm_frac_per_gpu": vcycle_gbs / world / peak,
                "phase_ms_per_solve": {n: v["ms"] / a.steps for n, v in prof.items()}}),
            "roofline": {"bound": "hbm", "kernel": kdesc, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "alg_bytes_per_launch": alg_bytes, "avg_launch_ms": avg_ms, "launches_timed": k["count"],
                         # the same launch time against the DRAM bytes ncu saw (profiles/traffic.json): the fused sweep
                         # moves fewer bytes than the SURVEY 8(d) model charges, so `frac` can exceed 1; this one cannot
                         "frac_of_peak_by_dram_traffic": (traffic / (avg_ms * 1e-3) / 1e9 / peak) if (traffic and avg_ms > 0) else None},
            "clocks": clk, "gpu_launches": int(launches), "pressure_solve_ms": ms / a.steps}
    if e2e:
        line["e2e"] = e2e
    if world == 1 and not a.no_cpu_baseline:
        r = cpu_reference_run(CPU_SAMPLE_GRID, a.aniso, a.map, a.relax, CPU_SAMPLE_CYCLES)
        line["cpu_baseline"] = {"value": r["mcups"], "unit": "Mcell-updates/s", "cores": r["cores"], "kind": "port",
                                "sample": f"{CPU_SAMPLE_GRID[0]}x{CPU_SAMPLE_GRID[1]}x{CPU_SAMPLE_GRID[2]} {a.map} grid, "
                                          f"{r['iters']} V-cycles in {r['secs']:.2f} s, OpenMP over {r['boxes']} y-slab boxes"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_OUT = None


def emit(line):
    _JSON_OUT.write(json.dumps(line) + "\n")
    _JSON_OUT.flush()


if __name__ == "__main__":
    # stdout carries exactly ONE JSON line: keep a private handle to it and point fd 1 at stderr, so that library
    # banners (NCCL prints its version on stdout) cannot get in front of it
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)

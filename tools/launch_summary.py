"""Groups an `ncu --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum` launch list by
kernel and grid: python tools/launch_summary.py launches.csv > profiles/<name>.txt"""
import csv
import sys
from collections import OrderedDict

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
h = rows[0]
ix = {n: h.index(n) for n in ("ID", "Kernel Name", "Grid Size", "Metric Name", "Metric Unit", "Metric Value")}
per = OrderedDict()
for r in rows[1:]:
    key = r[ix["ID"]]
    d = per.setdefault(key, {"name": r[ix["Kernel Name"]], "grid": r[ix["Grid Size"]]})
    v = float(r[ix["Metric Value"]].replace(",", ""))
    unit = r[ix["Metric Unit"]]
    m = r[ix["Metric Name"]]
    if m.startswith("gpu__time"):
        d["us"] = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)
    else:
        d[m] = v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
groups = OrderedDict()
for d in per.values():
    g = groups.setdefault((d["name"][:60], d["grid"]), {"n": 0, "us": 0.0, "b": 0.0})
    g["n"] += 1
    g["us"] += d.get("us", 0.0)
    g["b"] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
tot = sum(g["us"] for g in groups.values())
print(f"# {len(per)} launches, {tot / 1e3:.2f} ms in kernels (serialised, cold-cache ncu times: use the SHARES)")
for (name, grid), g in sorted(groups.items(), key=lambda kv: -kv[1]["us"]):
    bw = g["b"] / (g["us"] * 1e-6) / 1e9 if g["us"] else 0.0
    print(f"{name:62s} grid={grid:22s} n={g['n']:5d} total={g['us'] / 1e3:9.2f} ms share={100 * g['us'] / tot:5.1f}% "
          f"avg={g['us'] / g['n']:9.1f} us  dram={g['b'] / g['n'] / 1e6:9.1f} MB/launch {bw:8.1f} GB/s")

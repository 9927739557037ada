#!/bin/bash
mkdir -p gpurun_out
python tools/prof_terrain_solve.py > gpurun_out/r2_terrain_solve_plain.log 2>&1; echo "plain exit $?"; tail -3 gpurun_out/r2_terrain_solve_plain.log
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 6000 -c 1200 --csv --log-file gpurun_out/r2_terrain_bottom_launches.csv python tools/prof_terrain_solve.py > gpurun_out/r2_terrain_solve_ncu.log 2>&1; echo "ncu exit $?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(l for l in open("gpurun_out/r2_terrain_bottom_launches.csv") if l.startswith('"'))]
h = rows[0]; ki, vi, gi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Grid Size")
agg = collections.OrderedDict()
for r in rows[1:]:
    k = (r[ki][:60], r[gi])
    agg.setdefault(k, []).append(float(r[vi].replace(",", "")))
tot = sum(sum(v) for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print("%-62s %-14s n=%4d avg %8.1f us  share %4.1f%%" % (k[0], k[1], len(v), sum(v) / len(v) / 1e3, 100 * sum(v) / tot))
PY

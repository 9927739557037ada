#!/bin/bash
# cell-kind dispatch (FastView / DirView / ExView) in the rim kernels: parity, then bottom-depth and depth-0 timings
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -x -p no:cacheprovider -k "terrain or full or twisted or test_apply_and_residual or test_relax or test_restrict or leptic or baseline" > gpurun_out/r2_kind_tests.log 2>&1; echo "tests exit $?"; tail -5 gpurun_out/r2_kind_tests.log
python tools/prof_bottom_full.py > gpurun_out/r2_prof_bottom_full2.log 2>&1; echo "exit $?"; grep -v "^depth [0-9] {" gpurun_out/r2_prof_bottom_full2.log
python tools/prof_full.py > gpurun_out/r2_prof_full_kind.log 2>&1; echo "exit $?"; cat gpurun_out/r2_prof_full_kind.log
PROF_REPS=2 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_bottom_full_launches.csv python tools/prof_bottom_full.py > gpurun_out/r2_bottom_full_ncu.log 2>&1; echo "ncu exit $?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(l for l in open("gpurun_out/r2_bottom_full_launches.csv") if l.startswith('"'))]
h = rows[0]; ki, vi, gi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Grid Size")
for r in rows[-150:]:
    if "full" in r[ki] or "apply" in r[ki] or "restrict" in r[ki]:
        print(r[ki][:60], r[gi], r[vi])
PY

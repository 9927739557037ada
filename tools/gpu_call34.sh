#!/bin/bash
# full-metric instantiation of the persistent bottom kernel: a guarded first try on the terrain solves, then the whole
# single-GPU suite, smoke and the default bench line with whichever bottom path survived
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_baseline_configs.py tests/test_gpu_solve.py tests/test_projector.py -q -m gpu -x -p no:cacheprovider -k "terrain or dem or twisted or full" > gpurun_out/r2_fullpers_try.log 2>&1
rc=$?; echo "first try exit $rc"; tail -3 gpurun_out/r2_fullpers_try.log
if [ $rc -ne 0 ]; then export SOMAR_B200_BOTTOM_FULL_PERSISTENT=0; echo "FALLING BACK: SOMAR_B200_BOTTOM_FULL_PERSISTENT=0"; fi
timeout 400 python -m pytest tests -q -m gpu -x -p no:cacheprovider > gpurun_out/r2_final3_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r2_final3_tests.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 300 python bench.py > gpurun_out/r2_bench_final3_n1.json 2> gpurun_out/r2_bench_final3_n1.err; echo "bench exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/r2_bench_final3_n1.json"):
    try: d = json.loads(l)
    except Exception: continue
    print({k: d.get(k) for k in ("value", "ms_per_step", "gpu_launches")}, d.get("e2e", {}).get("value"))
    print("terrain", json.dumps(d.get("config", {}).get("secondary", {}).get("terrain"))[:1200])
PY

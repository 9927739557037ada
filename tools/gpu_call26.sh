#!/bin/bash
# inner/rim split of the full-metric kernels: parity suite, then the colour-pass timing with the split on and off
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -x -p no:cacheprovider -k "terrain or full or twisted or test_apply_and_residual or test_relax or leptic or baseline" > gpurun_out/r2_split_tests.log 2>&1; echo "tests exit $?"; tail -5 gpurun_out/r2_split_tests.log
python tools/prof_full.py > gpurun_out/r2_prof_full_split.log 2>&1; echo "split exit $?"; cat gpurun_out/r2_prof_full_split.log
SOMAR_B200_FULL_SPLIT=0 python tools/prof_full.py > gpurun_out/r2_prof_full_nosplit.log 2>&1; echo "nosplit exit $?"; cat gpurun_out/r2_prof_full_nosplit.log
python bench.py --steps 3 --warmup 3 > gpurun_out/r2_bench_split.json 2> gpurun_out/r2_bench_split.err; echo "bench exit $?"; python - <<'PY'
import json
for l in open("gpurun_out/r2_bench_split.json"):
    try: d = json.loads(l)
    except Exception: continue
    print({k: d.get(k) for k in ("value", "ms_per_step")}, d.get("e2e", {}).get("value"))
    print("terrain", json.dumps(d.get("config", {}).get("secondary", {}).get("terrain"))[:1500])
PY

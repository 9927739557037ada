#!/bin/bash
# full single-GPU suite, smoke, default bench line, bottom-depth timings
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x -p no:cacheprovider > gpurun_out/r2_final2_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r2_final2_tests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
python bench.py > gpurun_out/r2_bench_final2_n1.json 2> gpurun_out/r2_bench_final2_n1.err; echo "bench exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/r2_bench_final2_n1.json"):
    try: d = json.loads(l)
    except Exception: continue
    print({k: d.get(k) for k in ("value", "ms_per_step", "gpu_launches")}, d.get("e2e", {}).get("value"), d.get("roofline"))
    print("terrain", json.dumps(d.get("config", {}).get("secondary", {}).get("terrain"))[:1200])
PY
python tools/prof_bottom_full.py > gpurun_out/r2_prof_bottom_full4.log 2>&1; grep -v "^depth [0-9] {" gpurun_out/r2_prof_bottom_full4.log

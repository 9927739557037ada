#!/bin/bash
# bottom-solver chunk as a CUDA graph (one rank, full metric): parity subset, solve timing with the graph on / off, bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -x -p no:cacheprovider -k "terrain or full or twisted or baseline or solve or project" > gpurun_out/r2_graph_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r2_graph_tests.log
python tools/prof_terrain_solve.py 2>&1 | tail -2
SOMAR_B200_BOTTOM_GRAPH=0 python tools/prof_terrain_solve.py 2>&1 | tail -2
python bench.py > gpurun_out/r2_bench_graph_n1.json 2> gpurun_out/r2_bench_graph_n1.err; echo "bench exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/r2_bench_graph_n1.json"):
    try: d = json.loads(l)
    except Exception: continue
    print({k: d.get(k) for k in ("value", "ms_per_step", "gpu_launches")}, d.get("e2e", {}).get("value"))
    print("terrain", json.dumps(d.get("config", {}).get("secondary", {}).get("terrain"))[:1200])
PY

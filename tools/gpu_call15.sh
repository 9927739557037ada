#!/bin/bash
mkdir -p gpurun_out
{
timeout 300 python -m pytest tests/test_gpu_solve.py -x -q -m gpu -k "resident or bench_workload" 2>&1 | tail -4
SOMAR_B200_BOTTOM_PROF=1 python tools/prof_bottom.py 2>&1 | grep -i "probe" | tail -1
timeout 600 python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_g.json 2> gpurun_out/r2_bench_g.err; echo "exit $?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_g.json').read()); c=d['config']
print(round(d['value']), c['vcycles_per_solve'], c['bottom_iters_per_solve'], c['final_rel_residual'], c['phase_ms_per_vcycle']['bottom'], 'us/iter', c['phase_ms_per_vcycle']['bottom']*17/c['bottom_iters_per_solve']*1e3)"
} > gpurun_out/r2_call15.log 2>&1
tail -8 gpurun_out/r2_call15.log

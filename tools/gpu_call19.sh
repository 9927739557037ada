#!/bin/bash
mkdir -p gpurun_out
{
date
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -8
date
python tools/prof_advect.py 2>&1 | tail -4
timeout 600 python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/r2_bench_h.json 2> gpurun_out/r2_bench_h.err; echo "exit $?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_h.json').read()); c=d['config']
print(round(d['value']), d['e2e']['value'], c['vcycles_per_solve'], c['final_rel_residual'], c['phase_ms_per_vcycle'], d['roofline'])"
date
} > gpurun_out/r2_call19.log 2>&1
tail -20 gpurun_out/r2_call19.log

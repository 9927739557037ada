#!/bin/bash
mkdir -p gpurun_out
{
python -m pytest tests/test_gpu_solve.py -x -q -m gpu -k "resident" 2>&1 | tail -4
python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_d.json 2> gpurun_out/r2_bench_d.err; echo "bench exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_d.json').read()); c=d['config']
print(round(d['value']), c['vcycles_per_solve'], c['bottom_iters_per_solve'], c['phase_ms_per_vcycle'])"
} > gpurun_out/r2_call8.log 2>&1
tail -12 gpurun_out/r2_call8.log

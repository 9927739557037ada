#!/bin/bash
mkdir -p gpurun_out
{
timeout 300 python -m pytest tests/test_gpu_solve.py -x -q -m gpu -k "resident or bench_workload" 2>&1 | tail -4
for c in 8 4 0; do
SOMAR_B200_BOTTOM_CLUSTER=$c timeout 600 python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_cl$c.json 2> gpurun_out/r2_bench_cl.err; echo "cluster $c exit $?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_cl$c.json').read()); c=d['config']
print(round(d['value']), c['vcycles_per_solve'], c['bottom_iters_per_solve'], c['final_rel_residual'], c['phase_ms_per_vcycle']['bottom'], 'us/iter', c['phase_ms_per_vcycle']['bottom']*17/c['bottom_iters_per_solve']*1e3)"
done
} > gpurun_out/r2_call14.log 2>&1
tail -12 gpurun_out/r2_call14.log

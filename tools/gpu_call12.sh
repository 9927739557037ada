#!/bin/bash
mkdir -p gpurun_out
{
python -m pytest tests -q -m gpu -x 2>&1 | tail -6
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2_bench_f.json 2> gpurun_out/r2_bench_f.err; echo "bench exit $?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_f.json').read()); c=d['config']
print(round(d['value']), round(d['e2e']['value']), c['vcycles_per_solve'], c['strict3'])
print('terrain', json.dumps(c['secondary']['terrain'])[:900])
print('relax3', {k:v for k,v in c['secondary']['relax3'].items() if k not in ('residual_history','phase_ms_per_solve')})
print('sweep', c['sweep'])"
} > gpurun_out/r2_call12.log 2>&1
tail -14 gpurun_out/r2_call12.log

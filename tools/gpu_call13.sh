#!/bin/bash
mkdir -p gpurun_out
{
python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "line or relax" 2>&1 | tail -8
python bench.py --steps 1 --warmup 1 --relax 3 --no-secondary --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_relax3_compact.json 2> gpurun_out/r2_bench_relax3.err; echo "exit $?"
SOMAR_B200_LINE_COMPACT=0 python bench.py --steps 1 --warmup 1 --relax 3 --no-secondary --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_relax3_brick.json 2>> gpurun_out/r2_bench_relax3.err; echo "exit $?"
python -c "
import json
for f in ['r2_bench_relax3_compact','r2_bench_relax3_brick']:
    d=json.loads(open('gpurun_out/'+f+'.json').read()); c=d['config']
    print(f, round(d['value']), round(d['ms_per_step'],1), c['vcycles_per_solve'], c['final_rel_residual'], d['roofline']['kernel'][:20], d['roofline']['avg_launch_ms'], d['roofline']['frac'], c['phase_ms_per_vcycle'])"
} > gpurun_out/r2_call13.log 2>&1
tail -14 gpurun_out/r2_call13.log

#!/bin/bash
mkdir -p gpurun_out
{
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "variants" 2>&1 | tail -6
timeout 600 python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_e.json 2> gpurun_out/r2_bench_e.err; echo "bench exit $?"
SOMAR_B200_TMA=0 timeout 600 python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_e_notma.json 2>> gpurun_out/r2_bench_e.err; echo "bench exit $?"
python -c "
import json
for f in ['r2_bench_e','r2_bench_e_notma']:
    d=json.loads(open('gpurun_out/'+f+'.json').read()); c=d['config']
    print(f, round(d['value']), c['vcycles_per_solve'], c['final_rel_residual'], d['roofline']['avg_launch_ms'], c['phase_ms_per_vcycle'])"
} > gpurun_out/r2_call10.log 2>&1
tail -14 gpurun_out/r2_call10.log

#!/bin/bash
mkdir -p gpurun_out
{
date
timeout 900 python -m pytest tests/test_multi_rank.py -q -m gpu -k "8" 2>&1 | tail -6
date
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 2 --warmup 3 --no-secondary > gpurun_out/r2_bench_final_n8.json 2> gpurun_out/r2_bench_final_n8.err; echo "bench exit $?"
date
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_final_n8.json').read()); c=d['config']
print(round(d['value']), d['ms_per_step'], c['vcycles_per_solve'], c['final_rel_residual'], c['phase_ms_per_vcycle'], c.get('strict3'))"
} > gpurun_out/r2_call24.log 2>&1
tail -12 gpurun_out/r2_call24.log

#!/bin/bash
# final single-GPU evidence of round 2: tests, smoke, default bench line, reference arm, ncu launch list
mkdir -p gpurun_out
{
date
python -m pytest tests -q -m gpu 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
date
python bench.py > gpurun_out/r2_bench_final_n1.json 2> gpurun_out/r2_bench_final_n1.err; echo "bench exit $?"
date
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/r2_bench_final_reference.json 2> gpurun_out/r2_bench_final_reference.err; echo "reference exit $?"
date
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2600 --csv \
    --log-file gpurun_out/r2_launches_final.csv python bench.py --steps 1 --warmup 0 --no-secondary --no-e2e --no-cpu-baseline > gpurun_out/r2_ncu_launches_final.log 2>&1
date
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_final_n1.json').read()); print(round(d['value']), round(d['e2e']['value']), d['roofline']['frac'], d['cpu_baseline'], d['config'].get('converged'), d['clocks'])
d=json.loads(open('gpurun_out/r2_bench_final_reference.json').read()); print(d['value'], d['cpu_baseline'])"
} > gpurun_out/r2_call23.log 2>&1
tail -20 gpurun_out/r2_call23.log

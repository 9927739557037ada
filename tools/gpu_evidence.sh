#!/bin/bash
# round-2 evidence: default bench line (no profiler), reference arm, ncu launch list of the same workload, one full ncu capture
mkdir -p gpurun_out
python bench.py > gpurun_out/r2_bench_default_n1.json 2> gpurun_out/r2_bench_default_n1.err; echo "bench exit $?"
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err; echo "reference exit $?"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2600 --csv \
    --log-file gpurun_out/r2_launches.csv python bench.py --steps 1 --warmup 0 --no-secondary --no-e2e --no-cpu-baseline > gpurun_out/r2_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_gsrb_tma -s 1 -c 1 -f -o gpurun_out/r2_ncu_gsrb_tma \
    python bench.py --steps 1 --warmup 0 --no-secondary --no-e2e --no-cpu-baseline > gpurun_out/r2_ncu_full.log 2>&1
ls -la gpurun_out/r2_launches.csv gpurun_out/r2_ncu_gsrb_tma.ncu-rep
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_default_n1.json').read()); print(round(d['value']), round(d['e2e']['value']), d['roofline']['frac'], d['cpu_baseline'])
d=json.loads(open('gpurun_out/r2_bench_reference.json').read()); print(d['value'], d['cpu_baseline'])"

#!/bin/bash
# Development A/B runs of the depth-0 smoother variants on the GPU box (bench without the CPU and e2e legs).
# usage: tools/exp.sh "<FUSED> <PRODUCERS> <NOLOAD>" ...
for cfg in "$@"; do
  set -- $cfg
  export SOMAR_B200_FUSED=$1 SOMAR_B200_PRODUCERS=$2
  if [ "$3" = "1" ]; then export SOMAR_B200_NOLOAD=1; else unset SOMAR_B200_NOLOAD; fi
  timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_exp.json 2> gpurun_out/bench_exp.err
  tail -2 gpurun_out/bench_exp.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_exp.json'));print('fused/prod/noload $cfg ->', round(d['value'],1), 'Mcups,', round(d['roofline']['avg_launch_ms'],3), 'ms/launch', d['roofline']['kernel'][:14])"
done

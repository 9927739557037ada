#!/bin/bash
mkdir -p gpurun_out
{
python -m pytest tests/test_gpu_solve.py -x -q -m gpu -k "resident" 2>&1 | tail -15
python -m pytest tests -x -q -m gpu 2>&1 | tail -5
python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/r2_bench_b.json 2> gpurun_out/r2_bench_b.err; echo "bench exit $?"
SOMAR_B200_BOTTOM=2 python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_b_mode2.json 2>> gpurun_out/r2_bench_b.err; echo "bench exit $?"
} > gpurun_out/r2_call4.log 2>&1
tail -30 gpurun_out/r2_call4.log

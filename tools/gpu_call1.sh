#!/bin/bash
# round 2, first GPU call: the new entry point alone, the guide's TMA example, the whole GPU suite, one bench line
mkdir -p gpurun_out
{
echo "== no-alloc test alone"; python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -5
echo "== tma guide example"; timeout 60 tests/cuda/tma_guide_example; echo "exit $?"
echo "== tma probe mode 0"; (cd tests/cuda && nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 tma_probe.cu -o /tmp/tma_probe -lcuda 2>&1 | tail -2; timeout 60 /tmp/tma_probe 0; echo "exit $?")
echo "== full gpu suite"; python -m pytest tests -x -q -m gpu 2>&1 | tail -8
echo "== bench"; python bench.py --steps 2 --warmup 1 > gpurun_out/r2_bench_a.json 2> gpurun_out/r2_bench_a.err; echo "exit $?"; tail -5 gpurun_out/r2_bench_a.err
} > gpurun_out/r2_call1.log 2>&1
tail -40 gpurun_out/r2_call1.log

#!/bin/bash
mkdir -p gpurun_out
{
echo "== tma bisect"; for s in 0 1 2 3 4 5 6; do timeout 60 tests/cuda/tma_bisect $s; echo "exit $?"; done
echo "== full gpu suite"; python -m pytest tests -x -q -m gpu 2>&1 | tail -8
} > gpurun_out/r2_call2.log 2>&1
tail -40 gpurun_out/r2_call2.log

#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_leptic_solver.py tests/test_advect.py tests/test_leptic.py -q -m gpu 2>&1 | tail -40
} > gpurun_out/r2_call17.log 2>&1
tail -40 gpurun_out/r2_call17.log

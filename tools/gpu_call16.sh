#!/bin/bash
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_cc_projector.py tests/test_projector.py -q -m gpu 2>&1 | tail -30
} > gpurun_out/r2_call16.log 2>&1
tail -30 gpurun_out/r2_call16.log

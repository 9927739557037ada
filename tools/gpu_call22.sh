#!/bin/bash
mkdir -p gpurun_out
{
date
timeout 2400 python -m pytest tests/test_multi_rank.py -q -m gpu 2>&1 | tail -8
date
} > gpurun_out/r2_call22.log 2>&1
tail -10 gpurun_out/r2_call22.log

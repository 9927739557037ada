#!/bin/bash
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_advect.py tests/test_taylor_green.py tests/test_host_adapter.py -q -m gpu 2>&1 | tail -15
timeout 900 python -m pytest tests/test_multi_rank.py -q -m gpu -k "advection" 2>&1 | tail -30
} > gpurun_out/r2_call18.log 2>&1
tail -45 gpurun_out/r2_call18.log

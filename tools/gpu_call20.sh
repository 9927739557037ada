#!/bin/bash
mkdir -p gpurun_out
{
date
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -12
date
} > gpurun_out/r2_call20.log 2>&1
tail -14 gpurun_out/r2_call20.log

#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r2_call3.log 2>&1
tail -8 gpurun_out/r2_call3.log

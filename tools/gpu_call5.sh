#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -q -m gpu 2>&1 | tail -30 > gpurun_out/r2_call5.log
tail -30 gpurun_out/r2_call5.log

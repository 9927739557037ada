#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | tail -12 > gpurun_out/r2_call11.log
tail -12 gpurun_out/r2_call11.log

#!/bin/bash
mkdir -p gpurun_out
python tools/prof_bottom_full.py > gpurun_out/r2_prof_bottom_full.log 2>&1; echo "exit $?"; cat gpurun_out/r2_prof_bottom_full.log

#!/bin/bash
mkdir -p gpurun_out
python tools/prof_full.py > gpurun_out/r2_prof_full_plain.log 2>&1; echo "plain exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gsrb_full_rim -s 3 -c 1 -f -o gpurun_out/r2_ncu_gsrb_full_rim \
    python tools/prof_full.py > gpurun_out/r2_ncu_gsrb_full_rim.log 2>&1
ls -la gpurun_out/r2_ncu_gsrb_full_rim.ncu-rep; tail -3 gpurun_out/r2_ncu_gsrb_full_rim.log

#!/bin/bash
# one full ncu capture of the full-metric smoother (terrain family, 512x512x128): the weakest kernel of the round
mkdir -p gpurun_out
python tools/prof_full.py > gpurun_out/r2_prof_full.log 2>&1; echo "plain exit $?"; cat gpurun_out/r2_prof_full.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gsrb_full -s 3 -c 1 -f -o gpurun_out/r2_ncu_gsrb_full \
    python tools/prof_full.py > gpurun_out/r2_ncu_gsrb_full.log 2>&1
ls -la gpurun_out/r2_ncu_gsrb_full.ncu-rep; tail -3 gpurun_out/r2_ncu_gsrb_full.log

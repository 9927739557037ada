#!/bin/bash
# one ncu --set full capture of the full-metric persistent bottom kernel inside a terrain solve
mkdir -p gpurun_out
timeout 40 python tools/prof_terrain_solve.py > gpurun_out/r2_terrain_solve_final.log 2>&1; echo "plain exit $?"; tail -2 gpurun_out/r2_terrain_solve_final.log
timeout 45 ncu --set full --clock-control none -k regex:k_bicg_persistent -s 2 -c 1 -f -o gpurun_out/r2_ncu_bicg_persistent_full \
    python tools/prof_terrain_solve.py > gpurun_out/r2_ncu_bicg_persistent_full.log 2>&1
ls -la gpurun_out/r2_ncu_bicg_persistent_full.ncu-rep; tail -2 gpurun_out/r2_ncu_bicg_persistent_full.log

#!/bin/bash
mkdir -p gpurun_out
python tools/prof_bottom.py > gpurun_out/r2_prof_bottom.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_bicg_resident -s 3 -c 1 -f -o gpurun_out/r2_ncu_bicg_resident python tools/prof_bottom.py >> gpurun_out/r2_prof_bottom.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_gsrb_async -s 10 -c 1 -f -o gpurun_out/r2_ncu_gsrb_async python bench.py --steps 1 --warmup 1 --no-secondary --no-e2e --no-cpu-baseline >> gpurun_out/r2_prof_bottom.log 2>&1
tail -5 gpurun_out/r2_prof_bottom.log; ls -la gpurun_out/*.ncu-rep

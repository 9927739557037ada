#!/bin/bash
# last check of the shipped build (stencil_full.cuh now __host__ __device__): full-metric subset + smoke
mkdir -p gpurun_out
timeout 200 python -m pytest tests -q -m gpu -x -p no:cacheprovider -k "terrain or full or twisted or test_apply_and_residual or test_relax or test_restrict or baseline" > gpurun_out/r2_last_check.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r2_last_check.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1

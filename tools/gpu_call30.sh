#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -x -p no:cacheprovider -k "terrain or full or twisted or test_apply_and_residual or test_relax or test_restrict or baseline" > gpurun_out/r2_kind_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r2_kind_tests.log
python tools/prof_bottom_full.py > gpurun_out/r2_prof_bottom_full3.log 2>&1; echo "exit $?"; grep -v "^depth [0-9] {" gpurun_out/r2_prof_bottom_full3.log
python tools/prof_full.py > gpurun_out/r2_prof_full_kind.log 2>&1; echo "exit $?"; cat gpurun_out/r2_prof_full_kind.log

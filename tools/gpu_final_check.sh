#!/bin/bash
mkdir -p gpurun_out
{
for r in 1 2; do python -m pytest tests -q -m gpu 2>&1 | tail -4; done
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python -m pytest tests/test_zz_no_alloc.py -q -m gpu 2>&1 | tail -2
} > gpurun_out/r2_final_check.log 2>&1
tail -14 gpurun_out/r2_final_check.log

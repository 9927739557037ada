#!/bin/bash
mkdir -p gpurun_out
{ python -m pytest tests -q -m gpu 2>&1 | tail -12; python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3; } > gpurun_out/r2_call6.log 2>&1
tail -20 gpurun_out/r2_call6.log

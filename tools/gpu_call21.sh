#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=4 --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_worker.py nccl terrain > gpurun_out/r2_call21.log 2>&1
grep -n "MGPU\|Error\|error" gpurun_out/r2_call21.log | head -20

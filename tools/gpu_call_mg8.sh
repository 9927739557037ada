#!/bin/bash
mkdir -p gpurun_out
{
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 > gpurun_out/r2_bench_default_n8.json 2> gpurun_out/r2_bench_default_n8.err; echo "bench n=8 exit $?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --impl reference --steps 1 --warmup 0 > gpurun_out/r2_bench_reference_n8.json 2> gpurun_out/r2_bench_reference_n8.err; echo "reference n=8 exit $?"
timeout 900 python -m pytest tests/test_multi_rank.py -q -m gpu -k "8" 2>&1 | tail -5
} > gpurun_out/r2_call_mg8.log 2>&1
tail -12 gpurun_out/r2_call_mg8.log

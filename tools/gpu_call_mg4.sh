#!/bin/bash
mkdir -p gpurun_out
{
python -m pytest tests/test_multi_rank.py -q -m gpu 2>&1 | tail -25
for n in 4 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 2 --warmup 1 --no-secondary > gpurun_out/r2_bench_c_n$n.json 2> gpurun_out/r2_bench_c_n$n.err; echo "bench n=$n exit $?"
done
} > gpurun_out/r2_call_mg4.log 2>&1
tail -40 gpurun_out/r2_call_mg4.log

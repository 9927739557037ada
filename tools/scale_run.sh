N=$1
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/s13_v3_n$N.json 2> gpurun_out/s13_v3_n$N.err; tail -c 600 gpurun_out/s13_v3_n$N.json

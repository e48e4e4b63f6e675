#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --no-full > gpurun_out/r02_28_bench_n8.json 2> gpurun_out/r02_28_bench_n8.err; echo "bench n8 rc $?"
head -c 600 gpurun_out/r02_28_bench_n8.json

#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 400 python -m pytest tests/test_gpu_dist.py -x -q --timeout 200 > gpurun_out/r02_15_dist.log 2>&1; echo "dist rc $?"; tail -3 gpurun_out/r02_15_dist.log
timeout -k 5 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-full > gpurun_out/r02_15_bench_n2.json 2> gpurun_out/r02_15_bench_n2.err; echo "bench n2 rc $?"
tail -c 600 gpurun_out/r02_15_bench_n2.err

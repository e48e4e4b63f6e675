#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 300 python -m pytest tests/test_gpu_dist.py -m gpu -q --timeout 180 > gpurun_out/r02_27_pytest_dist.log 2>&1; echo "pytest dist rc $?"; tail -2 gpurun_out/r02_27_pytest_dist.log
timeout -k 5 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-full > gpurun_out/r02_27_bench_n2.json 2> gpurun_out/r02_27_bench_n2.err; echo "bench n2 rc $?"
head -c 1500 gpurun_out/r02_27_bench_n2.json

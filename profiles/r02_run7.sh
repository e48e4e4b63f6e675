#!/bin/bash
mkdir -p gpurun_out
bash profiles/ab4.sh default stay0 stay16 stay28 nt512:MIB_POOL_NT=512 nt512:MIB_POOL_NT=448 > gpurun_out/r02_7_ab.log 2>&1
timeout -k 5 420 python bench.py --steps 3 --warmup 3 > gpurun_out/r02_7_bench.json 2> gpurun_out/r02_7_bench.err
echo "bench rc $?" >> gpurun_out/r02_7_ab.log
timeout -k 5 240 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_7_bench_ref.json 2> gpurun_out/r02_7_bench_ref.err
echo "bench ref rc $?" >> gpurun_out/r02_7_ab.log
timeout -k 5 600 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/r02_7_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r02_7_ab.log
tail -5 gpurun_out/r02_7_pytest.log
echo done

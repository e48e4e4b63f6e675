#!/bin/bash
mkdir -p gpurun_out
( ldd oracle/_ref/scalar/ref_driver | grep -i "not found"; ls -la oracle/_ref/scalar | head -30; oracle/_ref/scalar/ref_driver; echo "rc $?" ) > gpurun_out/r02_8_refdiag.log 2>&1
timeout -k 5 300 python -m pytest tests/test_gpu_ocean.py -x -q -s --timeout 120 > gpurun_out/r02_8_ocean.log 2>&1
echo "ocean rc $?"
tail -30 gpurun_out/r02_8_ocean.log
timeout -k 5 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_8_bench_ref.json 2> gpurun_out/r02_8_bench_ref.err
tail -3 gpurun_out/r02_8_bench_ref.err

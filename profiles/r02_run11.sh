#!/bin/bash
mkdir -p gpurun_out
bash profiles/ab4.sh default a11 a31 f82 f44 > gpurun_out/r02_11_ab.log 2>&1
cat gpurun_out/r02_11_ab.log
timeout -k 5 400 python -m pytest tests/test_gpu_parity_plugins.py tests/test_gpu_pool.py tests/test_gpu_parity_adjoint.py -x -q --timeout 120 2>&1 | tail -5

#!/bin/bash
mkdir -p gpurun_out
bash profiles/ab4.sh default default:MIB_TILE=0 default:MIB_TILE=8 default:MIB_TILE=32 base nee1 nee2 nee2m1 > gpurun_out/r02_10_ab.log 2>&1
cat gpurun_out/r02_10_ab.log
timeout -k 5 400 python -m pytest tests/test_gpu_parity_plugins.py tests/test_gpu_pool.py tests/test_gpu_parity_adjoint.py -x -q --timeout 120 2>&1 | tail -5

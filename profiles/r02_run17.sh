#!/bin/bash
mkdir -p gpurun_out
bash profiles/ab4.sh default nodrain default nodrain > gpurun_out/r02_17_ab.log 2>&1
cat gpurun_out/r02_17_ab.log
timeout -k 5 300 python profiles/dist_balance.py 16 4 > gpurun_out/r02_17_balance.log 2>&1
cat gpurun_out/r02_17_balance.log
MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200_nodrain.so timeout -k 5 300 python profiles/dist_balance.py 16 4 2>&1 | tee gpurun_out/r02_17_balance_nodrain.log
timeout -k 5 400 python -m pytest tests/test_gpu_parity_plugins.py tests/test_gpu_pool.py tests/test_gpu_parity_adjoint.py -x -q --timeout 120 2>&1 | tail -3

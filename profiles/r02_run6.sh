#!/bin/bash
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -m gpu -x -q --timeout 90 > gpurun_out/r02_6_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r02_6_pytest.log
bash profiles/ab3.sh - MIB_VARIANT=0 MIB_POOL_NT=320 MIB_POOL_NT=256 MIB_FORWARD_KERNEL=thread > gpurun_out/r02_6_ab.log 2>&1
MIB_FORWARD_KERNEL=pool timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_forward_pool -c 1 -o gpurun_out/r02_6_c2_pool -f python profiles/prof_run.py c2 22 > gpurun_out/r02_6_ncu_c2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_forward_pool -c 1 -o gpurun_out/r02_6_c3pol_pool -f python profiles/prof_run.py c3pol 22 > gpurun_out/r02_6_ncu_c3pol.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_adjoint_pool -c 1 -o gpurun_out/r02_6_c4_apool -f python profiles/adj_run.py 4 > gpurun_out/r02_6_ncu_c4.log 2>&1
echo done

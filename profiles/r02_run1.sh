#!/bin/bash
# round 2, GPU call 1: correctness of the bank-owned-slot pools, A/B against the round-1 rings, CTA-size sweep, one ncu capture
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02_1_smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_1_pytest.log 2>&1; echo "pytest rc $?" | tee -a gpurun_out/r02_1_pytest.log
bash profiles/ab.sh default ring full24 > gpurun_out/r02_1_ab.log 2>&1
( for v in default ring full24; do
    if [ "$v" = "default" ]; then export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200.so; else export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200_$v.so; fi
    echo "== $v"; timeout 300 python profiles/prof_run.py c3pol 22 2>&1 | tail -1 | cut -c1-60; timeout 300 python profiles/adj_run.py 16 2>&1 | head -2 | tail -1
  done ) > gpurun_out/r02_1_ab2.log 2>&1
unset MIB_B200_LIB
( for nt in 128 192 256 320 384; do
    echo "== NT $nt"
    MIB_POOL_NT=$nt MIB_FORWARD_KERNEL=pool timeout 300 python profiles/prof_run.py c2 26 2>&1 | tail -1 | grep -oE "[0-9.]+ ms [0-9.]+ Msamples"
    MIB_POOL_NT=$nt MIB_FORWARD_KERNEL=pool timeout 300 python profiles/prof_run.py c3pol 22 2>&1 | tail -1 | cut -c1-60
    MIB_POOL_NT=$nt timeout 300 python profiles/adj_run.py 16 2>&1 | head -2 | tail -1
  done ) > gpurun_out/r02_1_nt.log 2>&1
MIB_FORWARD_KERNEL=pool timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_forward_pool -c 1 -o gpurun_out/r02_1_c2_pool_bos -f python profiles/prof_run.py c2 22 > gpurun_out/r02_1_ncu_c2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_adjoint_pool -c 1 -o gpurun_out/r02_1_c4_apool_bos -f python profiles/adj_run.py 4 > gpurun_out/r02_1_ncu_c4.log 2>&1
echo done

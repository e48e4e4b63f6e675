#!/bin/bash
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_forward_pool -c 1 -o gpurun_out/r02_13_c2_pool -f python profiles/prof_run.py c2 22 > gpurun_out/r02_13_ncu_c2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_forward_pool -c 1 -o gpurun_out/r02_13_c3pol_pool -f python profiles/prof_run.py c3pol 22 > gpurun_out/r02_13_ncu_c3pol.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_adjoint_pool -c 1 -o gpurun_out/r02_13_c4_apool -f python profiles/adj_run.py 4 > gpurun_out/r02_13_ncu_c4.log 2>&1
sha256sum mitsuba3_nasa_b200/libmi_b200.so > gpurun_out/r02_13_lib.sha
echo done

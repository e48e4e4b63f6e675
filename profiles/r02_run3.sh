#!/bin/bash
# round 2, GPU call 3: adjoint pool (bank-owned slots) debugging with the checked build, short timeouts
mkdir -p gpurun_out
L=$PWD/mitsuba3_nasa_b200
( for lib in libmi_b200_dbg.so libmi_b200.so; do
    for v in mono pol; do
      echo "== $lib $v"; MIB_B200_LIB=$L/$lib timeout -k 5 60 python profiles/adj_run.py 1 32 $v 2>&1 | tail -6 | cut -c1-300; echo "rc $?"
    done
  done ) > gpurun_out/r02_3_adj_debug.log 2>&1
timeout -k 10 900 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/r02_3_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r02_3_pytest.log
echo done

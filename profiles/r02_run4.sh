#!/bin/bash
mkdir -p gpurun_out
L=$PWD/mitsuba3_nasa_b200
run() { sz="$1"; shift; echo "== $sz $*"; env "$@" timeout -k 3 30 python profiles/adj_debug.py $sz 2>&1 | grep -v "^  File\|^    " | tail -8 | cut -c1-300; echo "rc ${PIPESTATUS[0]}"; }
( run "16 8 12 mono" MIB_B200_LIB=$L/libmi_b200_dbg.so MIB_POOL_NT=384
  run "16 32 12 mono" X=1
  run "64 16 32 pol" X=1
  echo "== adj_run 4 spp (C4 scene)"; timeout -k 3 60 python profiles/adj_run.py 4 2>&1 | tail -3 | cut -c1-200
) > gpurun_out/r02_4_adj_debug.log 2>&1
echo done

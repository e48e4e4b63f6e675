#!/bin/bash
mkdir -p gpurun_out
export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200_stats.so
( for w in "c2 26" "c3pol 22" "c3 22" "c1"; do echo "== $w"; timeout -k 3 60 python profiles/prof_run.py $w 2>&1 | tail -3; done
  echo "== c4"; timeout -k 3 90 python profiles/adj_run.py 16 2>&1 | tail -4 ) > gpurun_out/r02_12_stats.log 2>&1
cat gpurun_out/r02_12_stats.log

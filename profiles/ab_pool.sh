#!/bin/bash
# A/B of pool-kernel variants on the heterogeneous workloads: profiles/ab_pool.sh <suffix> ...
for v in "$@"; do
  if [ "$v" = "default" ]; then export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200.so; else export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200_$v.so; fi
  export MIB_FORWARD_KERNEL=pool
  c3=$(timeout 300 python profiles/prof_run.py c3 22 2>&1 | tail -1 | awk '{print $2, $4}')
  c3p=$(timeout 300 python profiles/prof_run.py c3pol 22 2>&1 | tail -1 | awk '{print $2, $4}')
  c2=$(timeout 300 python profiles/prof_run.py c2 24 2>&1 | tail -1 | grep -oE "[0-9.]+ ms [0-9.]+ Msamples")
  c1=$(timeout 300 python profiles/prof_run.py c1 2>&1 | tail -1 | awk '{print $2, $4}')
  echo "variant $v pool | c3 mono ms,Ms/s: $c3 | c3 polarized+Mie ms,Ms/s: $c3p | c2: $c2 | c1 ms,Ms/s: $c1"
done

#!/bin/bash
# A/B timing of library variants: profiles/ab.sh "<suffix list>"  (suffix "" = default build)
for v in "$@"; do
  if [ "$v" = "default" ]; then export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200.so; else export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200_$v.so; fi
  c3=$(python profiles/prof_run.py c3 22 2>&1 | tail -1 | awk '{print $2, $4}')
  c2=$(python profiles/prof_run.py c2 26 2>&1 | tail -1 | grep -oE "[0-9.]+ ms [0-9.]+ Msamples")
  c1=$(python profiles/prof_run.py c1 2>&1 | tail -1 | awk '{print $2, $4}')
  echo "variant $v | c3 ms,Ms/s: $c3 | c2: $c2 | c1 ms,Ms/s: $c1"
done

#!/bin/bash
# A/B timing of library variants: profiles/ab.sh <suffix> ...   ("default" = the shipped libmi_b200.so; other
# suffixes name libmi_b200_<suffix>.so built with `make B=build_<suffix> OUT=../libmi_b200_<suffix>.so EXTRA=-D...`).
# Each variant is timed with the path-pool kernel and with the one-thread-one-path kernel (MIB_FORWARD_KERNEL).
for v in "$@"; do
  if [ "$v" = "default" ]; then export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200.so; else export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200_$v.so; fi
  for k in pool thread; do
    export MIB_FORWARD_KERNEL=$k
    c3=$(timeout 300 python profiles/prof_run.py c3 22 2>&1 | tail -1 | awk '{print $2, $4}')
    c2=$(timeout 300 python profiles/prof_run.py c2 26 2>&1 | tail -1 | grep -oE "[0-9.]+ ms [0-9.]+ Msamples")
    c1=$(timeout 300 python profiles/prof_run.py c1 2>&1 | tail -1 | awk '{print $2, $4}')
    echo "variant $v kernel $k | c3 ms,Ms/s: $c3 | c2: $c2 | c1 ms,Ms/s: $c1"
  done
done

#!/bin/bash
# A/B of library variants with optional env: profiles/ab4.sh "suffix[:ENV=VAL]" ...   ("default" = libmi_b200.so)
for spec in "$@"; do
  v=${spec%%:*}; e=""; [ "$spec" != "$v" ] && e=${spec#*:}
  ( if [ "$v" = "default" ]; then export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200.so; else export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200_$v.so; fi
    [ -n "$e" ] && export $e
    c3=$(timeout -k 3 60 python profiles/prof_run.py c3 22 2>&1 | tail -1 | awk '{print $4}')
    c3p=$(timeout -k 3 60 python profiles/prof_run.py c3pol 22 2>&1 | tail -1 | awk '{print $4}')
    c2=$(timeout -k 3 60 python profiles/prof_run.py c2 26 2>&1 | tail -1 | grep -oE "[0-9.]+ Msamples" | grep -oE "[0-9.]+")
    c1=$(timeout -k 3 60 python profiles/prof_run.py c1 2>&1 | tail -1 | awk '{print $4}')
    c4=$(timeout -k 3 90 python profiles/adj_run.py 16 2>&1 | head -2 | tail -1 | grep -oE "[0-9.]+ Msamples" | grep -oE "[0-9.]+")
    printf "%-22s | c3 %7.1f | c3pol %7.1f | c2 %7.1f | c1 %7.1f | c4 adjoint %6.2f  (Msamples/s)\n" "$spec" ${c3:-0} ${c3p:-0} ${c2:-0} ${c1:-0} ${c4:-0} )
done

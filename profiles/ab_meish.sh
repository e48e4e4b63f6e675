mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_all_meish.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/gpu_all_meish.log
for rep in 1 2; do
for v in base default; do
  if [ "$v" = "default" ]; then export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200.so; else export MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200_$v.so; fi
  unset MIB_FORWARD_KERNEL
  c2=$(timeout 120 python profiles/prof_run.py c2 26 2>&1 | tail -1)
  c3=$(timeout 120 python profiles/prof_run.py c3 22 2>&1 | tail -1)
  c3p=$(timeout 120 python profiles/prof_run.py c3pol 22 2>&1 | tail -1)
  c1=$(timeout 120 python profiles/prof_run.py c1 2>&1 | tail -1)
  adj=$(timeout 120 python profiles/adj_run.py 16 2>&1 | grep "^c4" | tail -1)
  echo "[$rep] $v | c2: $c2 | c3: $c3 | c3pol: $c3p | c1: $c1 | $adj"
done
done 2>&1 | tee gpurun_out/ab_meish.log

#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_14_smoke.log 2>&1; echo "smoke rc $?"
timeout -k 5 700 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/r02_14_pytest.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r02_14_pytest.log
timeout -k 5 300 python bench.py --impl reference > gpurun_out/r02_14_bench_ref.json 2> gpurun_out/r02_14_bench_ref.err; echo "ref rc $?"
timeout -k 5 600 python bench.py > gpurun_out/r02_14_bench.json 2> gpurun_out/r02_14_bench.err; echo "bench rc $?"
timeout -k 5 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_14_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-full > gpurun_out/r02_14_ncu_bench.log 2>&1; echo "ncu rc $?"
echo done

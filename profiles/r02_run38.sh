#!/bin/bash
# final measurement of round 2: smoke, full GPU suite, both bench arms, launch list, ncu --set full captures of the three top
# kernels (for roofline.traffic), coverage-kernel throughput
mkdir -p gpurun_out
timeout -k 5 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_38_smoke.log 2>&1; echo "smoke rc $?"
timeout -k 5 900 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/r02_38_pytest.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r02_38_pytest.log
timeout -k 5 300 python bench.py --impl reference > gpurun_out/r02_38_bench_ref.json 2> gpurun_out/r02_38_bench_ref.err; echo "ref rc $?"
timeout -k 5 600 python bench.py > gpurun_out/r02_38_bench.json 2> gpurun_out/r02_38_bench.err; echo "bench rc $?"
timeout -k 5 300 python profiles/coverage_bench.py 4096 > gpurun_out/r02_38_coverage.jsonl 2> gpurun_out/r02_38_coverage.err; echo "coverage rc $?"
timeout -k 5 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_38_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-full > gpurun_out/r02_38_ncu_bench.log 2>&1; echo "ncu rc $?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_forward_pool -c 1 -o gpurun_out/r02_38_c2_pool -f python profiles/prof_run.py c2 22 > gpurun_out/r02_38_ncu_c2.log 2>&1; echo "ncu c2 rc $?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_forward_pool -c 1 -o gpurun_out/r02_38_c3pol_pool -f python profiles/prof_run.py c3pol 22 > gpurun_out/r02_38_ncu_c3pol.log 2>&1; echo "ncu c3pol rc $?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_adjoint_pool -c 1 -o gpurun_out/r02_38_c4_apool -f python profiles/adj_run.py 4 > gpurun_out/r02_38_ncu_c4.log 2>&1; echo "ncu c4 rc $?"
cat gpurun_out/r02_38_bench.json | head -c 3000
echo done

#!/bin/bash
mkdir -p gpurun_out
timeout 300 compute-sanitizer --tool memcheck --print-limit 5 python profiles/adj_run.py 1 32 > gpurun_out/r02_2_memcheck_adj.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_2_pytest.log 2>&1; echo "pytest rc $?" | tee -a gpurun_out/r02_2_pytest.log
echo done

#!/bin/bash
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -m gpu -x -q --timeout 90 > gpurun_out/r02_5_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r02_5_pytest.log
bash profiles/ab2.sh default ring nosteal steal8 full24 wagg > gpurun_out/r02_5_ab.log 2>&1
echo done

#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 900 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/r02_20_pytest.log 2>&1; echo "pytest rc $?"; tail -15 gpurun_out/r02_20_pytest.log

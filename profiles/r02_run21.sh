#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests/test_gpu_volpathmis.py -m gpu -q --timeout 120 -s > gpurun_out/r02_21_pytest.log 2>&1; echo "pytest rc $?"; grep -E "lanes|p-value|passed|failed|Error" gpurun_out/r02_21_pytest.log | tail -60

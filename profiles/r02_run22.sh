#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 1200 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/r02_22_pytest.log 2>&1; echo "pytest rc $?"; grep -E "passed|failed|FAILED|Error" gpurun_out/r02_22_pytest.log | tail -30

#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests/test_gpu_parity_adjoint.py -m gpu -q --timeout 180 -s > gpurun_out/r02_31_pytest.log 2>&1; echo "pytest rc $?"; grep -E "quotient|passed|failed|Error" gpurun_out/r02_31_pytest.log | tail

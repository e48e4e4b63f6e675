#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 900 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/r02_32_pytest.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r02_32_pytest.log; grep -E "FAILED|Error" gpurun_out/r02_32_pytest.log | head
timeout -k 5 600 python bench.py --no-cpu --no-full > gpurun_out/r02_32_bench.json 2> gpurun_out/r02_32_bench.err; echo "bench rc $?"
python - <<'PY'
import json
b = json.load(open('gpurun_out/r02_32_bench.json'))
print('C2', round(b['value'], 1), 'adjoint', round(b['adjoint']['value'], 1), 'adjoint e2e', round(b['adjoint']['e2e']['value'], 1))
PY
MIB_PRB_PRIMAL_KERNEL=thread timeout -k 5 300 python -m pytest tests/test_gpu_parity_forward.py tests/test_gpu_ad.py -m gpu -q --timeout 180 > gpurun_out/r02_32_pytest_thread.log 2>&1; echo "pytest (thread primal) rc $?"; tail -2 gpurun_out/r02_32_pytest_thread.log

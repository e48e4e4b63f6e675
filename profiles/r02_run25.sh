#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests/test_gpu_mesh.py tests/test_gpu_reference_scenes.py tests/test_gpu_volpathmis.py tests/test_gpu_plastic.py -m gpu -q --timeout 180 > gpurun_out/r02_25_pytest.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r02_25_pytest.log
echo "== shipped lib (mesh search cut at the nearest hit)"
timeout -k 5 300 python profiles/coverage_bench.py 1024 > gpurun_out/r02_25_coverage.jsonl 2> gpurun_out/r02_25_coverage.err; echo "coverage rc $?"
python - <<'PY'
import json
for l in open('gpurun_out/r02_25_coverage.jsonl'):
    d = json.loads(l); print(d['scene'], d['integrator'], d['kernel_ms'], d['msamples_per_s'])
PY
exit 0
MIB_B200_LIB=$PWD/mitsuba3_nasa_b200/libmi_b200_minb2.so timeout -k 5 300 python profiles/coverage_bench.py 1024 > gpurun_out/r02_25_coverage_minb2.jsonl 2> gpurun_out/r02_25_coverage_minb2.err; echo "coverage rc $?"
python - <<'PY'
import json
for l in open('gpurun_out/r02_25_coverage_minb2.jsonl'):
    d = json.loads(l); print(d['scene'], d['integrator'], d['kernel_ms'], d['msamples_per_s'])
PY

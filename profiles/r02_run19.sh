#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 300 python -m pytest tests/test_gpu_mesh.py tests/test_gpu_ocean.py -x -q -s --timeout 120 > gpurun_out/r02_19_mesh.log 2>&1; echo "mesh rc $?"; tail -25 gpurun_out/r02_19_mesh.log
bash profiles/ab4.sh default > gpurun_out/r02_19_ab.log 2>&1; cat gpurun_out/r02_19_ab.log

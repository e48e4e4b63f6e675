#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 500 python profiles/dist_balance.py 16 64 16 4 1 > gpurun_out/r02_16_balance.log 2>&1
cat gpurun_out/r02_16_balance.log

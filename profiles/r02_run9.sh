#!/bin/bash
mkdir -p gpurun_out
bash profiles/ab4.sh default s14 s8 s0 > gpurun_out/r02_9_ab.log 2>&1
cat gpurun_out/r02_9_ab.log

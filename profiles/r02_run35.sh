#!/bin/bash
mkdir -p gpurun_out
for leaf in 1 2 4 8; do
  echo "== MIB_BVH_LEAF=$leaf"
  MIB_BVH_LEAF=$leaf timeout -k 5 300 python profiles/coverage_bench.py 512 > gpurun_out/r02_35_coverage_leaf$leaf.jsonl 2> gpurun_out/r02_35_coverage_leaf$leaf.err
  python - <<PY
import json
for l in open('gpurun_out/r02_35_coverage_leaf$leaf.jsonl'):
    d = json.loads(l)
    if d['integrator'] == 'volpath': print(d['scene'], d['kernel_ms'], d['msamples_per_s'])
PY
done

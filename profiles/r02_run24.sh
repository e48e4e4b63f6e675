#!/bin/bash
mkdir -p gpurun_out
timeout -k 5 300 python profiles/coverage_bench.py 1024 > gpurun_out/r02_24_coverage.jsonl 2> gpurun_out/r02_24_coverage.err; echo "coverage rc $?"
cut -c1-330 gpurun_out/r02_24_coverage.jsonl
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_forward -c 2 -o gpurun_out/r02_24_cov_thread -f python - > gpurun_out/r02_24_ncu.log 2>&1 <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import mitsuba3_nasa_b200 as mi
mi.set_variant('llvm_ad_rgb')
f = 'tests/golden/participating_media/test_homo_null_envmap_volpath.xml'
sc = mi.load_file(f, spp=64)
mi.render(sc, seed=0, spp=64)
mi.render(sc, seed=1, spp=256)
PY
echo "ncu rc $?"

"""Throughput of the one-thread-one-path coverage kernels on the fork's own render-test scenes: each of the ten
covered participating-media scenes rendered with volpath and with volpathmis at 48 x 48 x `spp` samples through mi.render (device
film, one D2H of the 48 x 48 image), second call timed (`ms`: wall clock of the call; `kernel_ms`: CUDA-event time of the kernel, the throughput's denominator).  Prints one JSON line per (scene, integrator).  Not a bench line of
bench.py: these kernels exist for coverage of the reference's plugin set, the headline workloads run on the pool kernels.

    python profiles/coverage_bench.py [spp]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mitsuba3_nasa_b200 as mi  # noqa: E402

PM = os.path.join(ROOT, "tests", "golden", "participating_media")
STEMS = ['test_hetero_sigmat', 'test_hetero_sigmat_albedo', 'test_homo_null_envmap', 'test_homo_dielectric_envmap',
         'test_homo_roughdielectric_spot', 'test_homo_area_light', 'test_homo_obj_inside_medium', 'test_surf_area_light',
         'test_homo_spectral_sigmat', 'test_hetero_spectral_sigmat']


def main():
    spp = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    mi.set_variant('llvm_ad_rgb')
    for stem in STEMS:
        f = os.path.join(PM, stem + '_volpathmis.xml')
        if not os.path.exists(f):
            f = os.path.join(PM, stem + '_volpath.xml')
        for integrator in ('volpath', 'volpathmis'):
            scene = mi.load_file(f, spp=spp, integrator=integrator)
            mi.render(scene, seed=0, spp=64)
            t = time.perf_counter()
            img = np.asarray(mi.render(scene, seed=1, spp=spp))
            dt = time.perf_counter() - t
            c = scene.device.counters()
            n = 48 * 48 * spp
            print(json.dumps({"scene": stem, "integrator": integrator, "spp": spp, "samples": n, "ms": round(dt * 1e3, 2),
                              "kernel_ms": round(c['kernel_ms'], 2), "msamples_per_s": round(n / (c['kernel_ms'] * 1e-3) / 1e6, 1),
                              "per_sample": {k: round(c[k] / n, 2) for k in ('n_lookups', 'n_real', 'n_null', 'n_nee_steps', 'n_surface')},
                              "mean": round(float(img.mean()), 5)}), flush=True)


if __name__ == "__main__":
    main()

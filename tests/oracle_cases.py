"""Small fixed-seed oracle runs shared by tests/test_oracle_regression.py and tests/golden/make_oracle_regression.py.
One thread each: the oracle merges per-thread partial sums in completion order, which moves the last bits of a gradient."""
import numpy as np

from mitsuba3_nasa_b200 import scene as S, scenes
from orc import OracleScene


def _grid(n, seed, lo=0.0):
    return np.random.default_rng(seed).uniform(lo, 1.0, (n, n, n, 1)).astype(np.float32)


def run_all():
    out = {}
    # volpath, rgb, homogeneous HG cube: developed film
    sc = S.load_dict(scenes.c1_homogeneous_cube(res=12, spp=8, max_depth=16), "llvm_ad_rgb")
    out["c1_rgb_film"] = OracleScene(sc).render(3, 8, nthreads=1)[1]
    # volpath, mono, gaussian filter
    sc = S.load_dict(scenes.c1_homogeneous_cube(res=10, spp=4, max_depth=8, rfilter='gaussian'), "llvm_ad_mono")
    out["c1_mono_gauss_film"] = OracleScene(sc).render(1, 4, nthreads=1)[1]
    # volpathaos, polarized Rayleigh slab: Stokes vector and per-lane values, sun and constant sky
    for sun in (True, False):
        sc = S.load_dict(scenes.c2_rayleigh_slab(spp=4, depolarization=0.03, sun=sun), "llvm_ad_mono_polarized")
        o = OracleScene(sc)
        out["c2_stokes_sun%d" % sun] = o.render_1(5, 1 << 12, nthreads=1)
        out["c2_lanes_sun%d" % sun] = o.sample_lanes(5, 1 << 12, 0, 512, nthreads=1)
    # volpath / volpathaos on a heterogeneous grid, Rayleigh and HG phase
    sc = S.load_dict(scenes.c3_cloud(res=8, spp=4, max_depth=12, grid=_grid(10, 2), optical_depth=6.0, phase={'type': 'rayleigh'}),
                     "llvm_ad_mono_polarized")
    out["c3_pol_lanes"] = OracleScene(sc).sample_lanes(2, 4, 0, 8 * 8 * 4, nthreads=1)
    sc = S.load_dict(scenes.c3_cloud(res=8, spp=4, max_depth=12, grid=_grid(10, 3), optical_depth=8.0, phase={'type': 'hg', 'g': 0.6},
                                     integrator='volpathaos'), "llvm_ad_mono")
    out["c3_mono_aos_film"] = OracleScene(sc).render(4, 4, nthreads=1)[1]
    # adjoints: prbvolpath (mono) and prb_volpathaos (polarized) w.r.t. a density grid
    for integ, variant, key in (('prbvolpath', 'llvm_ad_mono', 'prb_mono'), ('prb_volpathaos', 'llvm_ad_mono_polarized', 'prb_pol')):
        g = _grid(6, 7, lo=0.05)
        d = scenes.c3_cloud(res=6, spp=8, max_depth=10, integrator=integ, grid=g, optical_depth=4.0, phase={'type': 'rayleigh'})
        d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
        sc = S.load_dict(d, variant)
        o = OracleScene(sc)
        o.render_1_backward(0, 8, [1.0, 0.5, 0.0, 0.0], nthreads=1)
        out[key + "_grad"] = o.grad_get(sc.desc.media[0].sigma_t.buffer, g.size)
    return out

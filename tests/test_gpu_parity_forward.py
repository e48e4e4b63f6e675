"""GPU parity tests: CUDA kernels (through the C ABI) vs the CPU oracle on the same seeds.

Tolerances: per-lane radiance / Stokes components within 1e-4 relative (north_star);
the count of lanes whose decision sequence flipped (rel. err > 1e-3) is reported and bounded.
Device math is bit-exact against the oracle's restatement of Dr.Jit's polynomials.
"""
import numpy as np
import pytest

from mitsuba3_nasa_b200 import scene as S, scenes, _lib
from orc import OracleScene, oracle_lib, fp
import ctypes as C

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4


def _lane_compare(sc_dict, variant, spp, count, seed=0, lane_begin=0, flip_frac=2e-3):
    sc = S.load_dict(sc_dict, variant)
    dev = _lib.DeviceScene(sc, 0)
    orc = OracleScene(sc)
    g = dev.sample_lanes(seed, spp, lane_begin, count).astype(np.float64)
    o = orc.sample_lanes(seed, spp, lane_begin, count).astype(np.float64)
    scale = np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)
    rel = np.abs(g - o) / scale
    bad = (rel > REL_TOL).any(axis=1)
    gc, oc = dev.counters(), orc.counters()
    print("lanes %d: max rel %.3g, lanes over tol %d, counters gpu %s oracle %s" %
          (count, rel.max(), bad.sum(), {k: gc[k] for k in ('n_real', 'n_null', 'n_nee_steps', 'n_surface', 'n_lookups')},
           {k: oc[k] for k in ('n_real', 'n_null', 'n_nee_steps', 'n_surface', 'n_lookups')}))
    assert bad.mean() <= flip_frac, "too many lanes differ: %d of %d (max rel %.3g)" % (bad.sum(), count, rel.max())
    return rel, gc, oc


@pytest.mark.parametrize("op,name,lo,hi", [(1, "exp", -90, 90), (2, "log", 1e-30, 1e4), (3, "cbrt", -10, 10), (4, "asin", -1, 1)])
def test_device_math_bit_exact(op, name, lo, hi):
    rng = np.random.default_rng(op)
    x = rng.uniform(lo, hi, 200000).astype(np.float32)
    got = _lib.debug_math(op, x)
    ref = np.empty_like(x)
    getattr(oracle_lib(), "orc_" + name)(fp(x), fp(ref), C.c_size_t(x.size))
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_device_sincos_bit_exact():
    x = np.random.default_rng(7).uniform(-200, 200, 200000).astype(np.float32)
    got = _lib.debug_math(0, x).reshape(-1, 2)
    s, c = np.empty_like(x), np.empty_like(x)
    oracle_lib().orc_sincos(fp(x), fp(s), fp(c), C.c_size_t(x.size))
    assert np.array_equal(got[:, 0].view(np.uint32), s.view(np.uint32))
    assert np.array_equal(got[:, 1].view(np.uint32), c.view(np.uint32))


def test_c1_rgb_volpath_lanes():
    _lane_compare(scenes.c1_homogeneous_cube(res=32, spp=16), "llvm_ad_rgb", 16, 32 * 32 * 16)


def test_c1_mono_volpath_lanes():
    _lane_compare(scenes.c1_homogeneous_cube(res=16, spp=16), "llvm_ad_mono", 16, 16 * 16 * 16)


def test_c1_spectral_extinction_branch():
    d = scenes.c1_homogeneous_cube(res=16, spp=16)
    d['medium1']['has_spectral_extinction'] = True
    _lane_compare(d, "llvm_ad_rgb", 16, 16 * 16 * 16)


def test_c2_polarized_volpathaos_lanes():
    rel, gc, oc = _lane_compare(scenes.c2_rayleigh_slab(spp=4), "llvm_ad_mono_polarized", 20000, 20000)
    for k in ('n_real', 'n_nee_steps', 'n_surface'):
        assert abs(gc[k] - oc[k]) <= 2e-3 * oc[k]


def test_c2_polarized_volpath_constant_emitter():
    _lane_compare(scenes.c2_rayleigh_slab(spp=4, integrator='volpath', sun=False), "llvm_ad_mono_polarized", 20000, 20000)


def test_c2_render_1_stokes():
    sc = S.load_dict(scenes.c2_rayleigh_slab(spp=4), "llvm_ad_mono_polarized")
    dev = _lib.DeviceScene(sc, 0)
    n = 1 << 17
    got = dev.render_1(3, n)
    ref = OracleScene(sc).render_1(3, n)
    assert np.abs(got - ref).max() <= REL_TOL * abs(ref[0])


def test_c1_film_develop():
    sc = S.load_dict(scenes.c1_homogeneous_cube(res=24, spp=8), "llvm_ad_rgb")
    dev = _lib.DeviceScene(sc, 0)
    img = dev.render(5, 8)
    raw_ref, ref = OracleScene(sc).render(5, 8)
    assert img.shape == (24, 24, 3)
    assert np.abs(img - ref).max() <= REL_TOL * np.abs(ref).max()
    raw = dev.render(5, 8, develop=False)
    assert np.allclose(raw[..., 3], 8.0)


def test_lane_range_shards_sum_to_whole():
    """Sharding by lane range (multi-GPU layout): two half renders add up to the full film."""
    sc = S.load_dict(scenes.c1_homogeneous_cube(res=16, spp=8), "llvm_ad_rgb")
    dev = _lib.DeviceScene(sc, 0)
    full = dev.render(2, 8, develop=False)
    total = 16 * 16 * 8
    dev.set_lane_range(0, total // 2)
    a = dev.render(2, 8, develop=False)
    dev.set_lane_range(total // 2, total)
    b = dev.render(2, 8, develop=False)
    assert np.allclose(a + b, full, rtol=1e-6, atol=1e-6)


def test_heterogeneous_grid_lanes():
    grid = scenes.cloud_grid(32)
    cost, val = scenes.hg_table(0.8, 257)
    phase = {'type': 'tabphase', 'cost': cost, 'values': val}
    d = scenes.c3_cloud(res=16, spp=8, grid=grid, phase=phase, optical_depth=8.0)
    rel, gc, oc = _lane_compare(d, "llvm_ad_mono", 8, 16 * 16 * 8, flip_frac=5e-3)
    assert gc['n_lookups'] > 0


def test_heterogeneous_polarized_tabphase_lanes():
    grid = scenes.cloud_grid(32)
    cost = np.linspace(-1, 1, 181).astype(np.float32)
    m11 = (1.0 + 0.5 * cost ** 2 + 0.8 * np.exp(4 * (cost - 1))).astype(np.float32)
    m12 = (-0.4 * (1 - cost ** 2)).astype(np.float32)
    m33 = (0.9 * cost).astype(np.float32)
    m34 = (0.1 * (1 - cost ** 2)).astype(np.float32)
    # tabphase_polarized returns the UNNORMALISED matrix with the NORMALISED pdf
    # (tabphase_polarized.cpp:161-164,201-202): tables must integrate to 1 over the sphere
    k = np.float32(1.0 / (2 * np.pi * np.trapezoid(m11.astype(np.float64), cost.astype(np.float64))))
    m11, m12, m33, m34 = m11 * k, m12 * k, m33 * k, m34 * k
    phase = {'type': 'tabphase_polarized', 'cost': cost, 'm11': m11, 'm12': m12, 'm33': m33, 'm34': m34}
    d = scenes.c3_cloud(res=12, spp=8, grid=grid, phase=phase, optical_depth=6.0)
    _lane_compare(d, "llvm_ad_mono_polarized", 8, 12 * 12 * 8, flip_frac=5e-3)


@pytest.mark.parametrize("rfilter", ["gaussian", "tent"])
@pytest.mark.parametrize("spp", [8, 2])          # >= 4: coalesced put, < 4: general put (integrator.cpp:320)
@pytest.mark.parametrize("variant", ["llvm_ad_rgb", "llvm_ad_mono"])
def test_reconstruction_filters(rfilter, spp, variant):
    """ImageBlock::put with a reconstruction filter (imageblock.cpp:261-530) + weight division (hdrfilm.cpp:304-418).
    Cropped film so that the crop offset arithmetic is exercised too."""
    d = scenes.c1_homogeneous_cube(res=24, spp=spp, max_depth=8, rfilter=rfilter)
    d['sensor']['film'].update({'crop_offset_x': 3, 'crop_offset_y': 5, 'crop_width': 17, 'crop_height': 13})
    if 'mono' in variant:
        d['sensor']['film']['pixel_format'] = 'luminance'
    sc = S.load_dict(d, variant)
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    g_raw = dev.render(2, spp, develop=False)
    g_dev = dev.render(2, spp, develop=True)
    o_raw, o_dev = orc.render(2, spp, nthreads=4)
    assert g_dev.shape == o_dev.shape == (13, 17, 3 if 'rgb' in variant else 1)
    # float atomics in arbitrary order vs double accumulation
    assert np.allclose(g_raw, o_raw, rtol=2e-5, atol=1e-6)
    assert np.allclose(g_dev, o_dev, rtol=REL_TOL, atol=1e-6)


def test_polarized_film_rejects_filters():
    d = scenes.c2_rayleigh_slab(spp=4)
    d['sensor']['film']['rfilter'] = {'type': 'gaussian'}
    sc = S.load_dict(d, "llvm_ad_mono_polarized")
    dev = _lib.DeviceScene(sc, 0)
    with pytest.raises(RuntimeError, match="box filter"):
        dev.render(0, 4)
    assert np.isfinite(dev.render_1(0, 64)).all()        # render_1 never touches the film filter

"""GPU tests of the path-pool kernels (k_forward_pool / k_adjoint_pool) at sizes where the stage rings wrap many times.

A lane's result depends only on the lane index (seed = TEA(base_seed + seed, lane), integrator.cpp:323-340), not on which
slot / warp / CTA walked it.  So the pool kernel and the one-thread-one-path kernel must agree on every event counter
EXACTLY and on the film / gradients up to the order of the float atomics -- a size-independent property that is checked
here at 10^5 .. 10^6 lanes (the lane-by-lane comparisons against the oracle in test_gpu_parity_*.py run at sizes the
oracle finishes in seconds and go through whichever kernel the scene selects).
"""
import os

import numpy as np
import pytest

from mitsuba3_nasa_b200 import scene as S, scenes, _lib

pytestmark = pytest.mark.gpu

COUNTERS = ('n_samples', 'n_lookups', 'n_real', 'n_null', 'n_nee_steps', 'n_surface')


class kernel:
    """Force a kernel family for the duration of a block (the library reads the variable at every launch)."""

    def __init__(self, var, value):
        self.var, self.value = var, value

    def __enter__(self):
        self.old = os.environ.get(self.var)
        os.environ[self.var] = self.value

    def __exit__(self, *a):
        if self.old is None:
            del os.environ[self.var]
        else:
            os.environ[self.var] = self.old


def _mie_free_phase():
    cost, val = scenes.hg_table(0.8, 257)
    return {'type': 'tabphase', 'cost': cost, 'values': val}


def _pol_phase():
    cost = np.linspace(-1, 1, 181).astype(np.float32)
    m11 = (1.0 + 0.5 * cost ** 2 + 0.8 * np.exp(4 * (cost - 1))).astype(np.float32)
    m12 = (-0.4 * (1 - cost ** 2)).astype(np.float32)
    m33 = (0.9 * cost).astype(np.float32)
    m34 = (0.1 * (1 - cost ** 2)).astype(np.float32)
    k = np.float32(1.0 / (2 * np.pi * np.trapezoid(m11.astype(np.float64), cost.astype(np.float64))))
    return {'type': 'tabphase_polarized', 'cost': cost, 'm11': m11 * k, 'm12': m12 * k, 'm33': m33 * k, 'm34': m34 * k}


FORWARD_CASES = {
    'c1_rgb_volpath': lambda: (scenes.c1_homogeneous_cube(res=96, spp=64), 'llvm_ad_rgb', 64),
    'c1_rgb_gaussian_filter': lambda: (scenes.c1_homogeneous_cube(res=64, spp=16, rfilter='gaussian'), 'llvm_ad_rgb', 16),
    'c2_polarized_volpathaos_film': lambda: (scenes.c2_rayleigh_slab(spp=4), 'llvm_ad_mono_polarized', 1 << 13),   # 1x1 film: all samples land in ONE fp32 pixel
    'c2_polarized_volpath_constant': lambda: (scenes.c2_rayleigh_slab(spp=4, integrator='volpath', sun=False), 'llvm_ad_mono_polarized', 1 << 13),
    'c3_mono_cloud': lambda: (scenes.c3_cloud(res=128, spp=32, grid=scenes.cloud_grid(64), phase=_mie_free_phase()), 'llvm_ad_mono', 32),
    'c3_polarized_cloud': lambda: (scenes.c3_cloud(res=96, spp=32, grid=scenes.cloud_grid(48), phase=_pol_phase(), optical_depth=10.0),
                                   'llvm_ad_mono_polarized', 32),
    'c3_rgb_cloud_spectral': lambda: (scenes.c3_cloud(res=96, spp=16, grid=scenes.cloud_grid(32), phase={'type': 'hg', 'g': 0.6}), 'llvm_ad_rgb', 16),
}


@pytest.mark.parametrize("case", sorted(FORWARD_CASES))
def test_forward_pool_equals_thread_kernel(case):
    d, variant, spp = FORWARD_CASES[case]()
    sc = S.load_dict(d, variant)
    dev = _lib.DeviceScene(sc, 0)
    out = {}
    for k in ('pool', 'thread'):
        with kernel('MIB_FORWARD_KERNEL', k):
            raw = np.asarray(dev.render(11, spp, develop=False), dtype=np.float64)
            c = dev.counters()
            out[k] = (raw, {kk: c[kk] for kk in COUNTERS})
    assert out['pool'][1] == out['thread'][1], "event counters differ: a lane was lost, duplicated or walked differently"
    assert out['pool'][1]['n_samples'] == raw.shape[0] * raw.shape[1] * spp
    scale = np.abs(out['thread'][0]).max()
    # float atomics in a different order (spp partial sums per pixel): relative to the largest pixel
    assert np.abs(out['pool'][0] - out['thread'][0]).max() <= 5e-5 * scale


def test_forward_pool_render_1_sums():
    sc = S.load_dict(scenes.c2_rayleigh_slab(spp=4), 'llvm_ad_mono_polarized')
    dev = _lib.DeviceScene(sc, 0)
    res = {}
    for k in ('pool', 'thread'):
        with kernel('MIB_FORWARD_KERNEL', k):
            res[k] = (dev.render_1(5, 1 << 21), dev.counters())
    assert all(res['pool'][1][kk] == res['thread'][1][kk] for kk in COUNTERS)
    # fp64 sums of the same 2^21 floats in a different order
    assert np.abs(res['pool'][0] - res['thread'][0]).max() <= 1e-12


def test_forward_pool_lane_range_and_tiny_wavefronts():
    """Empty and ragged inputs: lane ranges that are not multiples of the chunk / pool size, a single lane, no lane."""
    sc = S.load_dict(scenes.c3_cloud(res=32, spp=8, grid=scenes.cloud_grid(16), phase=_mie_free_phase()), 'llvm_ad_mono')
    dev = _lib.DeviceScene(sc, 0)
    total = 32 * 32 * 8
    with kernel('MIB_FORWARD_KERNEL', 'pool'):
        full = dev.render(4, 8, develop=False)
        acc = np.zeros_like(full)
        for b, e in ((0, 1), (1, 777), (777, 777), (777, 4099), (4099, total)):
            if b == e:
                continue                          # an empty range means "whole wavefront" in the ABI (mi_b200.h)
            dev.set_lane_range(b, e)
            acc += dev.render(4, 8, develop=False)
            assert dev.counters()['n_samples'] == e - b
        dev.set_lane_range(0, 0)
    assert np.allclose(acc, full, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("variant,integ", [("llvm_ad_mono", "prbvolpath"), ("llvm_ad_rgb", "prbvolpath"),
                                           ("llvm_ad_mono_polarized", "prb_volpathaos")])
def test_adjoint_pool_equals_thread_kernel(variant, integ):
    grid = scenes.cloud_grid(24)
    phase = _pol_phase() if 'polarized' in variant else _mie_free_phase()
    d = scenes.c3_cloud(res=64, spp=32, grid=grid, phase=phase, optical_depth=8.0, integrator=integ, max_depth=16)
    sc = S.load_dict(d, variant)
    dev = _lib.DeviceScene(sc, 0)
    slot = sc.desc.media[0].sigma_t.buffer
    rng = np.random.default_rng(5)
    nch = 1 if variant.endswith('mono') else (3 if 'rgb' in variant else 4)
    grad_img = rng.uniform(0.5, 1.5, (64, 64, nch)).astype(np.float32)
    out = {}
    for k in ('pool', 'thread'):
        with kernel('MIB_ADJOINT_KERNEL', k):
            dev.grad_zero()
            dev.render_backward(3, 32, grad_img)
            c = dev.counters()
            out[k] = (dev.grad_get(slot, grid.size).astype(np.float64), {kk: c[kk] for kk in COUNTERS})
    assert out['pool'][1] == out['thread'][1]
    scale = np.abs(out['thread'][0]).max()
    assert scale > 0
    assert np.abs(out['pool'][0] - out['thread'][0]).max() <= 1e-4 * scale      # float atomics in a different order


@pytest.mark.parametrize("fk", ["pool", "thread"])
def test_lane_stripes_add_up_to_the_whole_film(fk):
    """mib_scene_set_lane_stripes (multi-GPU load balance): the ranks' striped renders sum to the whole film, every lane
    is walked exactly once (counters add up), for a wavefront that is not a multiple of the stripe chunk."""
    sc = S.load_dict(scenes.c3_cloud(res=40, spp=8, grid=scenes.cloud_grid(16), phase=_mie_free_phase()), 'llvm_ad_mono')
    dev = _lib.DeviceScene(sc, 0)
    with kernel('MIB_FORWARD_KERNEL', fk):
        full = np.asarray(dev.render(6, 8, develop=False), dtype=np.float64)
        c_full = dev.counters()
        acc, cnt = np.zeros_like(full), {k: 0 for k in COUNTERS}
        world = 3
        for r in range(world):
            dev.set_lane_stripes(r, world, 8 * 24)                 # 24 pixels per chunk: 1600 pixels = 66 chunks + 16 pixels
            acc += dev.render(6, 8, develop=False)
            c = dev.counters()
            for k in COUNTERS:
                cnt[k] += c[k]
        dev.set_lane_stripes(0, 1, 1)
    assert cnt == {k: c_full[k] for k in COUNTERS}
    assert np.allclose(acc, full, rtol=1e-5, atol=1e-6)


def test_lane_stripes_adjoint():
    grid = scenes.cloud_grid(16)
    d = scenes.c3_cloud(res=32, spp=16, grid=grid, phase=_mie_free_phase(), optical_depth=6.0, integrator='prbvolpath', max_depth=12)
    sc = S.load_dict(d, 'llvm_ad_mono')
    dev = _lib.DeviceScene(sc, 0)
    slot = sc.desc.media[0].sigma_t.buffer
    g_img = np.ones((32, 32, 1), dtype=np.float32)
    dev.grad_zero(); dev.render_backward(2, 16, g_img)
    whole = dev.grad_get(slot, grid.size).astype(np.float64)
    dev.grad_zero()
    for r in range(2):
        dev.set_lane_stripes(r, 2, 16 * 10)
        dev.render_backward(2, 16, g_img)                          # gradients accumulate over the two "ranks"
    dev.set_lane_stripes(0, 1, 1)
    parts = dev.grad_get(slot, grid.size).astype(np.float64)
    assert np.abs(parts - whole).max() <= 1e-4 * np.abs(whole).max()

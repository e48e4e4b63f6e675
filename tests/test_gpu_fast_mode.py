"""GPU tests of the backend's fast mode: majorant supergrid + DDA free-flight sampling.

The fork has no counterpart (its heterogeneous medium uses one global majorant, heterogeneous.cpp:163-185), and the fast
mode does not reproduce the reference's random sequence.  What is checked:
  * the supergrid is a valid, tight majorant: equal to a numpy max-pool over the texels a look-up in the cell can touch;
  * fast mode is an unbiased estimator of the same integral: images / Stokes sums / gradients agree with parity mode
    (majorant_resolution = 0, itself checked lane by lane against the oracle) within Monte Carlo confidence bounds;
  * it does what it is for: fewer tentative collisions per sample.
"""
import numpy as np
import pytest

from mitsuba3_nasa_b200 import scene as S, scenes, _lib

pytestmark = pytest.mark.gpu


def _phase():
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


@pytest.mark.parametrize("n,M", [(32, 8), (40, 16), (24, 5)])
def test_supergrid_is_the_max_pool(n, M):
    grid = scenes.cloud_grid(n)
    d = scenes.c3_cloud(res=8, spp=4, grid=grid, phase=_phase(), majorant_resolution=M)
    sc = S.load_dict(d, 'llvm_ad_mono')
    dev = _lib.DeviceScene(sc, 0)
    got = dev.supergrid(0)
    scale = sc.desc.media[0].scale
    g = grid.reshape(n, n, n)
    ref = np.zeros((M, M, M), dtype=np.float64)

    def rng(c):
        lo = max(int(np.floor(np.float32(c) * np.float32(n) / np.float32(M) - np.float32(.5))) - 1, 0)
        hi = min(int(np.floor(np.float32(c + 1) * np.float32(n) / np.float32(M) - np.float32(.5))) + 2, n - 1)
        return lo, hi + 1
    for z in range(M):
        for y in range(M):
            for x in range(M):
                (z0, z1), (y0, y1), (x0, x1) = rng(z), rng(y), rng(x)
                ref[z, y, x] = g[z0:z1, y0:y1, x0:x1].max()
    ref *= scale
    assert np.allclose(got, ref, rtol=3e-6, atol=0)
    assert got.max() <= sc.desc.media[0].scale * grid.max() * (1 + 2e-6)
    # a majorant everywhere: trilinear values never exceed the cell's entry (checked on the texel centres)
    cell = np.minimum((np.arange(n) + 0.5) / n * M, M - 1e-6).astype(int)
    assert (g * scale <= got[np.ix_(cell, cell, cell)] * (1 + 1e-6)).all()


def _film(d, variant, seed, spp):
    sc = S.load_dict(d, variant)
    dev = _lib.DeviceScene(sc, 0)
    img = np.asarray(dev.render(seed, spp, develop=True), dtype=np.float64)
    return img, dev.counters()


@pytest.mark.parametrize("variant,phase", [("llvm_ad_mono", _phase), ("llvm_ad_mono_polarized", _pol_phase)])
def test_fast_mode_agrees_with_parity_mode(variant, phase):
    grid = scenes.cloud_grid(48)
    res, spp = 24, 2048
    kw = dict(res=res, spp=spp, grid=grid, phase=phase(), optical_depth=12.0, max_depth=32)
    ref, c_ref = _film(scenes.c3_cloud(**kw), variant, 1, spp)
    ref2, _ = _film(scenes.c3_cloud(**kw), variant, 2, spp)             # second parity-mode render: the noise level
    fast, c_fast = _film(scenes.c3_cloud(majorant_resolution=12, **kw), variant, 3, spp)
    sigma = np.abs(ref - ref2)[..., 0].mean() / np.sqrt(2.0)            # mean abs deviation of one render, per pixel
    # per-pixel differences fast - parity behave like parity - parity (no bias): compare the means over the film
    assert abs((fast - ref)[..., 0].mean()) <= 4.0 * sigma / np.sqrt(res * res) * 1.5
    assert np.abs(fast - ref)[..., 0].mean() <= 2.0 * np.abs(ref2 - ref)[..., 0].mean()
    assert c_fast['n_samples'] == c_ref['n_samples']
    # the point of the mode: fewer tentative collisions / shadow steps (grid look-ups) per sample
    assert c_fast['n_lookups'] < 0.85 * c_ref['n_lookups']
    # real scatters are a property of the medium, not of the majorant
    assert abs(c_fast['n_real'] - c_ref['n_real']) <= 0.02 * c_ref['n_real']


def test_fast_mode_render_1_stokes():
    grid = scenes.cloud_grid(32)
    kw = dict(res=4, spp=4, grid=grid, phase=_pol_phase(), optical_depth=8.0, max_depth=32)
    n = 1 << 20
    a = _lib.DeviceScene(S.load_dict(scenes.c3_cloud(**kw), 'llvm_ad_mono_polarized'), 0).render_1(1, n // 16)
    b = _lib.DeviceScene(S.load_dict(scenes.c3_cloud(**kw), 'llvm_ad_mono_polarized'), 0).render_1(2, n // 16)
    f = _lib.DeviceScene(S.load_dict(scenes.c3_cloud(majorant_resolution=8, **kw), 'llvm_ad_mono_polarized'), 0).render_1(3, n // 16)
    noise = np.abs(a - b).max()
    assert np.abs(f - a).max() <= 4.0 * max(noise, 1e-4 * abs(a[0]))


def test_fast_mode_adjoint_gradient():
    grid = scenes.cloud_grid(16)
    kw = dict(res=32, spp=512, grid=grid, phase=_phase(), optical_depth=6.0, max_depth=16, integrator='prbvolpath')
    out = []
    for seed, M in ((1, 0), (2, 0), (3, 6)):
        sc = S.load_dict(scenes.c3_cloud(majorant_resolution=M, **kw), 'llvm_ad_mono')
        dev = _lib.DeviceScene(sc, 0)
        dev.grad_zero()
        dev.render_backward(seed, 512, np.ones((32, 32, 1), dtype=np.float32))
        out.append(dev.grad_get(sc.desc.media[0].sigma_t.buffer, grid.size).astype(np.float64))
    ref, ref2, fast = out
    noise = np.linalg.norm(ref - ref2)
    assert np.linalg.norm(ref) > 0
    assert np.linalg.norm(fast - ref) <= 1.6 * noise                   # fast - parity looks like parity - parity
    assert abs(fast.sum() - ref.sum()) <= 4.0 * max(abs(ref.sum() - ref2.sum()), 1e-3 * abs(ref.sum()))

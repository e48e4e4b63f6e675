"""BASELINE.json's configurations at FULL size, checked through size-independent properties (the oracle cannot walk 1e8
lanes in test time): sharding linearity (lane ranges add up to the whole wavefront -- the multi-GPU layout of SURVEY 8e),
exact sample counts, determinism of the event counters, agreement between the two kernel families, gradient accumulation."""
import os

import numpy as np
import pytest

import mitsuba3_nasa_b200 as mi
from mitsuba3_nasa_b200 import scenes

pytestmark = pytest.mark.gpu
COUNTERS = ('n_samples', 'n_lookups', 'n_real', 'n_null', 'n_nee_steps', 'n_surface')


def test_c2_full_size_1e8_samples_shards_and_kernel_families():
    """configs[1]: plane-parallel Rayleigh slab, mono_polarized, distant sensor, 1e8 samples, Stokes IQUV."""
    mi.set_variant('llvm_ad_mono_polarized')
    scene = mi.load_dict(scenes.c2_rayleigh_slab(spp=4, integrator='volpathaos', max_depth=64))
    dev = scene.device
    n = 100_000_000
    whole = dev.render_1(0, n, normalize=False)
    c_whole = dev.counters()
    assert c_whole['n_samples'] == n
    parts, cnt = np.zeros(4), {k: 0 for k in COUNTERS}
    for b, e in ((0, 37_000_001), (37_000_001, 80_000_000), (80_000_000, n)):       # ragged shards, as ranks would take them
        dev.set_lane_range(b, e)
        parts += dev.render_1(0, n, normalize=False)
        c = dev.counters()
        for k in COUNTERS:
            cnt[k] += c[k]
    dev.set_lane_range(0, 0)
    assert cnt == {k: c_whole[k] for k in COUNTERS}                  # every lane walked exactly once, identically
    assert np.abs(parts - whole).max() <= 1e-9 * abs(whole[0])       # fp64 sums of the same floats
    stokes = whole / n
    # physics of the slab (sun and view at 30 deg from zenith, opposite azimuths): I > 0, Q < 0, U = V = 0 within noise
    assert 0.09 < stokes[0] < 0.105 and -0.03 < stokes[1] < -0.015
    assert abs(stokes[2]) < 5e-5 and stokes[3] == 0.0
    os.environ['MIB_FORWARD_KERNEL'] = 'pool'
    try:
        pool = dev.render_1(0, n, normalize=False)
        c_pool = dev.counters()
    finally:
        del os.environ['MIB_FORWARD_KERNEL']
    assert {k: c_pool[k] for k in COUNTERS} == {k: c_whole[k] for k in COUNTERS}
    assert np.abs(pool - whole).max() <= 1e-9 * abs(whole[0])


def test_c3_full_size_film_1024x1024x256():
    """configs[2]: heterogeneous 256^3 cloud, tabulated Mie phase from a lognormal size distribution, 1024x1024, 256 spp."""
    mi.set_variant('llvm_ad_mono')
    scene = mi.load_dict(scenes.c3_cloud())                           # res 1024, spp 256, grid 256^3, Mie x lognormal table
    dev = scene.device
    raw = np.asarray(dev.render(0, 256, develop=False), dtype=np.float64)
    c = dev.counters()
    assert c['n_samples'] == 1024 * 1024 * 256
    # box filter: weight = number of samples that fell into the pixel = spp, except where fp32 rounds posx + jitter up
    # to the next pixel (jitter within 3e-5 of 1 at x ~ 1000; the reference's ImageBlock::put floors the same float)
    # (and a few samples of the last column / row round up to x = 1024 and leave the film, as in the reference)
    total = 1024 * 1024 * 256
    assert raw.shape == (1024, 1024, 2) and total - 200 <= raw[..., 1].sum() <= total and np.abs(raw[..., 1] - 256.0).max() <= 3
    img = raw[..., 0] / raw[..., 1]
    assert np.isfinite(img).all() and img.min() >= 0.0
    # a non-absorbing-ish cloud (albedo 0.99) under a unit constant emitter: every pixel close to, and never above, 1
    assert 0.9 < img.mean() <= 1.0 + 1e-3 and img.max() < 4.0       # (single samples may exceed 1: NEE against a peaked Mie phase)
    assert abs(img[:8, :8].mean() - 1.0) < 1e-6                       # corner pixels miss the cube: exactly the emitter
    # determinism: the same seed gives the same event counters (and a different seed does not)
    dev.set_lane_range(1024 * 256 * 480, 1024 * 256 * 544)            # 64 rows through the middle of the cloud
    dev.render(0, 256, develop=False, to_host=False); a = dev.counters()
    dev.render(0, 256, develop=False, to_host=False); b = dev.counters()
    dev.render(1, 256, develop=False, to_host=False); d = dev.counters()
    dev.set_lane_range(0, 0)
    assert all(a[k] == b[k] for k in COUNTERS) and any(a[k] != d[k] for k in COUNTERS)


def test_c4_full_size_adjoint_512x512x64_accumulates():
    """configs[3]: prb adjoint w.r.t. the 256^3 density grid, mono_polarized, 512x512, 64 spp."""
    mi.set_variant('llvm_ad_mono_polarized')
    scene = mi.load_dict(scenes.c4_cloud_adjoint())
    dev = scene.device
    slot = scene.flat.desc.media[0].sigma_t.buffer
    n = 256 ** 3
    dev.grad_zero()
    dev.render_1_backward(0, 64, [1.0, 0.0, 0.0, 0.0])
    c = dev.counters()
    assert c['n_samples'] == 512 * 512 * 64
    g1 = dev.grad_get(slot, n).astype(np.float64)
    assert np.isfinite(g1).all() and np.abs(g1).max() > 0
    # gradients accumulate until zeroed (dr.grad semantics): a second identical pass doubles them up to atomic order
    dev.render_1_backward(0, 64, [1.0, 0.0, 0.0, 0.0])
    g2 = dev.grad_get(slot, n).astype(np.float64)
    assert np.abs(g2 - 2.0 * g1).max() <= 1e-4 * np.abs(g1).max()
    # linearity in the adjoint radiance: d/dI scaled by 3 scales the gradient by 3
    dev.grad_zero()
    dev.render_1_backward(0, 64, [3.0, 0.0, 0.0, 0.0])
    g3 = dev.grad_get(slot, n).astype(np.float64)
    assert np.abs(g3 - 3.0 * g1).max() <= 1e-4 * np.abs(g1).max()

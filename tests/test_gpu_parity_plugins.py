"""GPU parity, plugin coverage: every scene of the integrator-level pin (tests/scalar_cases.py -- the scenes on which the
oracle reproduces the reference's own volpath / volpathaos, tests/test_oracle_vs_scalar_ref.py) rendered lane by lane by the
CUDA kernels and by the oracle under the llvm_ad_* seeding, in BOTH kernel families (one thread per path, path pool).
Covers what round 1 left without a GPU test: sphere, radiancemeter, nearest filter, repeat / mirror wrap, isotropic phase,
hide_emitters, 3-channel albedo grids, spectral extinction, and rng_arg_order = right_to_left.

Also: 1e5-lane slices of the REAL C3 and C4 scenes (256^3 grid, 1801-node Mie x lognormal table, optical depth 20) against
the oracle -- the BASELINE configs at their own size, not a toy version of them.

Tolerance: 1e-4 of the lane's largest component (north_star); lanes whose decision sequence flipped are counted and bounded
(measured: bit-equal on every scene)."""
import os

import numpy as np
import pytest

from mitsuba3_nasa_b200 import _abi as A, scene as S, scenes, _lib
from orc import OracleScene
import scalar_cases

pytestmark = pytest.mark.gpu

LLVM = {'scalar_rgb': 'llvm_ad_rgb', 'scalar_mono': 'llvm_ad_mono', 'scalar_mono_polarized': 'llvm_ad_mono_polarized'}
# (the ocean-interface scenes have their own file: tests/test_gpu_ocean.py)
CASES = {n: c for n, c in scalar_cases.cases().items() if c['kind'] == 'samples' and not n.startswith('ocean_')}


def _compare(sc, seed, spp, lane_begin, count, flip_frac=2e-3, counters=True):
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    g = dev.sample_lanes(seed, spp, lane_begin, count).astype(np.float64)
    gc = dev.counters()
    o = orc.sample_lanes(seed, spp, lane_begin, count).astype(np.float64)
    oc = orc.counters()
    scale = np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)
    bad = (np.abs(g - o) / scale > 1e-4).any(axis=1)
    exact = np.array_equal(g, o)
    print("lanes %d: bit-equal %s, over tolerance %d" % (count, exact, bad.sum()))
    assert bad.mean() <= flip_frac, "%d of %d lanes differ" % (bad.sum(), count)
    if counters and not bad.any():
        for k in ('n_real', 'n_null', 'n_nee_steps', 'n_surface', 'n_lookups', 'n_samples'):
            assert gc[k] == oc[k], (k, gc[k], oc[k])


@pytest.fixture(params=["thread", "pool"])
def kernel(request):
    old = os.environ.get("MIB_FORWARD_KERNEL")
    os.environ["MIB_FORWARD_KERNEL"] = request.param
    yield request.param
    if old is None:
        del os.environ["MIB_FORWARD_KERNEL"]
    else:
        os.environ["MIB_FORWARD_KERNEL"] = old


@pytest.mark.parametrize("name", sorted(CASES))
def test_pinned_scenes_lane_by_lane(name, kernel):
    case = CASES[name]
    sc = S.load_dict(case['scene'], LLVM[case['variant']])
    W, H = sc.desc.sensor.crop_width, sc.desc.sensor.crop_height
    spp = min(case['spp'], 64)
    _compare(sc, case['seed'], spp, 0, W * H * spp)


@pytest.mark.parametrize("name", ["c1_mono_seed3", "het_pol_tab", "area_emitter"])
def test_rng_arg_order_right_to_left(name, kernel):
    """the draw order of phase->sample(ctx, mei, next_1d(), next_2d()) as gcc evaluates it (the order the scalar
    reference build of tests/test_oracle_vs_scalar_ref.py runs with)."""
    case = CASES[name]
    sc = S.load_dict(case['scene'], LLVM[case['variant']])
    sc.desc.integrator.rng_arg_order = A.RNG_ARGS_RIGHT_TO_LEFT
    W, H = sc.desc.sensor.crop_width, sc.desc.sensor.crop_height
    _compare(sc, 1, case['spp'], 0, W * H * case['spp'])


def test_fixed_max_value_is_the_majorant(kernel):
    """grid.cpp:296-297: a gridvolume given `max_value` uses it as the majorant as it is (more null collisions, different
    random sequence than with the data's own maximum): lane by lane against the oracle."""
    d = scenes.c3_cloud(res=8, spp=8, grid=np.random.default_rng(9).uniform(0.0, 1.0, (6, 6, 6, 1)).astype(np.float32),
                        phase={'type': 'isotropic'}, optical_depth=4.0)
    d['object']['interior']['sigma_t']['max_value'] = 1.5
    sc = S.load_dict(d, "llvm_ad_mono")
    assert sc.desc.media[0].sigma_t.fixed_max == 1
    _compare(sc, 0, 8, 0, 8 * 8 * 8)
    e = scenes.c3_cloud(res=8, spp=8, grid=np.random.default_rng(9).uniform(0.0, 1.0, (6, 6, 6, 1)).astype(np.float32),
                        phase={'type': 'isotropic'}, optical_depth=4.0)
    a = _lib.DeviceScene(sc, 0)
    a.sample_lanes(0, 8, 0, 512)
    b = _lib.DeviceScene(S.load_dict(e, "llvm_ad_mono"), 0)
    b.sample_lanes(0, 8, 0, 512)
    assert a.counters()['n_null'] > b.counters()['n_null']


@pytest.mark.parametrize("variant", ["llvm_ad_mono", "llvm_ad_mono_polarized"])
def test_c3_real_scene_lane_slices(variant):
    """BASELINE configs[2] as specified: 1024 x 1024 x 256 spp film over the 256^3 cloud, Mie x lognormal table with 1801
    nodes, majorant optical depth 20.  Two slices of the wavefront (image centre: through the cloud; a corner: mostly
    background) against the oracle, 2e4 lanes each."""
    sc = S.load_dict(scenes.c3_cloud(), variant)
    spp = 256
    centre = (512 * 1024 + 500) * spp
    for begin, count in ((centre, 20000), (3 * spp, 4000)):
        _compare(sc, 0, spp, begin, count)


def test_c4_real_scene_primal_lane_slice():
    """BASELINE configs[3]: the prb_volpathaos primal pass on the real C4 scene (512 x 512 x 64 spp)."""
    sc = S.load_dict(scenes.c4_cloud_adjoint(), "llvm_ad_mono_polarized")
    _compare(sc, 0, 64, (256 * 512 + 250) * 64, 20000)


def test_c4_real_scene_gradient_slice():
    """BASELINE configs[3]: gradient of a lane slice of the real C4 scene (render_1_backward, dL/dI = 1) w.r.t. the 256^3
    density grid, against the oracle.  Tolerance: 1e-3 of the largest gradient entry (north_star; float atomics vs the
    oracle's double accumulation)."""
    sc = S.load_dict(scenes.c4_cloud_adjoint(), "llvm_ad_mono_polarized")
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    begin, count = (256 * 512 + 250) * 64, 8192
    dev.set_lane_range(begin, begin + count)
    dev.grad_zero()
    dev.render_1_backward(0, 64, [1.0, 0.0, 0.0, 0.0])
    orc.render_1_backward(0, 64, [1.0, 0.0, 0.0, 0.0], lanes=(begin, begin + count))
    gc, oc = dev.counters(), orc.counters()
    for k in ('n_real', 'n_null', 'n_nee_steps', 'n_surface', 'n_lookups'):
        assert gc[k] == oc[k], (k, gc[k], oc[k])
    slot = sc.desc.media[0].sigma_t.buffer
    n = sc.buffers[slot].size
    g, o = dev.grad_get(slot, n).astype(np.float64), orc.grad_get(slot, n)
    assert np.abs(o).max() > 0
    assert np.abs(g - o).max() <= 1e-3 * np.abs(o).max(), (np.abs(g - o).max(), np.abs(o).max())


def test_three_channel_albedo_grid_gradient_in_a_mono_variant():
    """a 3-channel grid in a 1-channel variant is its luminance; the adjoint must scatter the luminance
    weights to all three channels (adjoint.cuh sink_volume)."""
    from mitsuba3_nasa_b200.transforms import ScalarTransform4f as T
    alb = np.random.default_rng(8).uniform(0.4, 0.9, size=(3, 3, 3, 3)).astype(np.float32)
    d = scenes.c1_homogeneous_cube(res=8, spp=16, max_depth=10)
    d['integrator'] = {'type': 'prbvolpath', 'max_depth': 10, 'rr_depth': 5}
    d['medium1']['albedo'] = {'type': 'gridvolume', 'data': alb, 'to_world': T.translate([-1, -1, -1]) @ T.scale(2.0)}
    d['sensor']['film']['pixel_format'] = 'luminance'
    sc = S.load_dict(d, 'llvm_ad_mono')
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    dev.grad_zero()
    dev.render_1_backward(3, 16, [1.0, 0, 0, 0])
    orc.render_1_backward(3, 16, [1.0, 0, 0, 0])
    slot = sc.desc.media[0].albedo.buffer
    g, o = dev.grad_get(slot, alb.size).astype(np.float64), orc.grad_get(slot, alb.size)
    assert np.all(np.abs(o).reshape(-1, 3).sum(axis=0) > 0)
    assert np.abs(g - o).max() <= 1e-3 * np.abs(o).max()

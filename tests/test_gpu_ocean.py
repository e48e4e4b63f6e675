"""GPU parity of the atmosphere-ocean interface: volpathaos over a smooth / rough dielectric ocean surface
with refractive next-event estimation (volpathaos.cpp:232-290, 299-455; dielectric.cpp, roughdielectric.cpp), walked by
k_forward<SPEC, 4> (csrc/docean.cuh).

The scenes are those of the integrator-level pin (tests/scalar_cases.py ocean_*), on which the oracle reproduces the
reference's own scalar build sample by sample (tests/test_oracle_vs_scalar_ref.py).  Here the CUDA kernel must reproduce
the oracle lane by lane under the llvm_ad_* seeding, in both draw orders of (next_1d(), next_2d()).

Tolerance: 1e-4 of the lane's largest component (north_star); lanes whose decision sequence flipped (a comparison decided
the other way by a last-bit difference of a transcendental) are counted and bounded by 0.2 %."""
import numpy as np
import pytest

from mitsuba3_nasa_b200 import _abi as A, scene as S, scenes, _lib
from orc import OracleScene
import scalar_cases

pytestmark = pytest.mark.gpu

LLVM = {'scalar_rgb': 'llvm_ad_rgb', 'scalar_mono': 'llvm_ad_mono', 'scalar_mono_polarized': 'llvm_ad_mono_polarized'}
CASES = {n: c for n, c in scalar_cases.cases().items() if c['kind'] == 'samples' and n.startswith('ocean_')}


def _compare(sc, seed, spp, count, flip_frac=2e-3):
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    g = dev.sample_lanes(seed, spp, 0, count).astype(np.float64)
    gc = dev.counters()
    o = orc.sample_lanes(seed, spp, 0, count).astype(np.float64)
    oc = orc.counters()
    assert np.array_equal(np.isnan(g), np.isnan(o))
    g, o = np.nan_to_num(g), np.nan_to_num(o)
    scale = np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)
    bad = (np.abs(g - o) / scale > 1e-4).any(axis=1)
    print("lanes %d: bit-equal %s, over tolerance %d, non-zero lanes %d" % (count, np.array_equal(g, o), bad.sum(), (o[:, 0] != 0).sum()))
    assert (o[:, 0] != 0).mean() > 0.2
    assert bad.mean() <= flip_frac, "%d of %d lanes differ" % (bad.sum(), count)
    if not bad.any():
        for k in ('n_real', 'n_null', 'n_nee_steps', 'n_surface', 'n_lookups', 'n_samples'):
            assert gc[k] == oc[k], (k, gc[k], oc[k])
    return g


@pytest.mark.parametrize("order", [A.RNG_ARGS_LEFT_TO_RIGHT, A.RNG_ARGS_RIGHT_TO_LEFT])
@pytest.mark.parametrize("name", sorted(CASES))
def test_ocean_scenes_lane_by_lane(name, order):
    case = CASES[name]
    sc = S.load_dict(case['scene'], LLVM[case['variant']])
    sc.desc.integrator.rng_arg_order = order
    _compare(sc, case['seed'], 4096, 4096)


def test_ocean_rgb_variant():
    """No scalar pin of its own (the pinned ocean scenes are mono / mono_polarized): the rgb instantiation against the oracle."""
    sc = S.load_dict(scenes.aos_ocean(rough=True), "llvm_ad_rgb")
    _compare(sc, 7, 2048, 2048)


def test_render_1_agrees_with_the_reference_render_1():
    """SamplingIntegrator::render_1 of the unmodified reference (scalar_mono_polarized, 2e4 samples; the fixture of
    tests/test_oracle_vs_scalar_ref.py) against 4e6 GPU samples with independent seeding: same expectation.  The reference
    run's own Monte-Carlo error (2e4 samples) sets the tolerance: 5 % of I."""
    import os
    golden = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scalar_ref.npz"))
    ref = golden['ocean_rough_pol_render1'].reshape(4, 4)[:, 0].astype(np.float64)
    sc = S.load_dict(scenes.aos_ocean(rough=True), "llvm_ad_mono_polarized")
    dev = _lib.DeviceScene(sc, 0)
    assert dev.forward_kernel() == "thread"
    stokes = dev.render_1(0, 4000000)[:4]
    print("reference", ref, "gpu", stokes)
    assert np.abs(stokes - ref).max() <= 0.05 * ref[0]


def test_a_diffuse_bounce_that_keeps_the_sign_of_dz_counts_as_a_refraction():
    """volpathaos.cpp:262-285 applies its "refraction occurred" test to every non-null BSDF event: after a diffuse bounce off a
    surface that is not horizontal the next-event estimation asks the diffuse BSDF for a transmission lobe and gets nothing
    (diffuse.cpp:108-111).  Pinned on the reference's scalar build (tests/scalar_cases.py mesh_aos_pol_sun); here with an
    analytic sphere: the kernel that carries the interface state must be the one that runs, and it must equal the oracle."""
    from mitsuba3_nasa_b200.transforms import ScalarTransform4f as T
    d = scenes.c2_rayleigh_slab(spp=4)
    d['ball'] = {'type': 'sphere', 'center': [0.0, 0.0, 0.5], 'radius': 0.3,
                 'bsdf': {'type': 'diffuse', 'reflectance': 0.8}}
    sc = S.load_dict(d, "llvm_ad_mono_polarized")
    dev = _lib.DeviceScene(sc, 0)
    assert dev.forward_kernel() == "thread"
    g = dev.sample_lanes(1, 4096, 0, 4096).astype(np.float64)
    o = OracleScene(sc).sample_lanes(1, 4096, 0, 4096).astype(np.float64)
    bad = (np.abs(g - o) > 1e-4 * np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)).any(axis=1)
    print("over tolerance", bad.sum())
    assert bad.mean() <= 2e-3
    # and the plain C2 scene (horizontal diffuse ground only) stays on the pool kernel
    assert _lib.DeviceScene(S.load_dict(scenes.c2_rayleigh_slab(spp=4), "llvm_ad_mono_polarized"), 0).forward_kernel() == "pool"


@pytest.mark.parametrize("variant", ["llvm_ad_mono_polarized", "llvm_ad_rgb"])
@pytest.mark.parametrize("rough", [True, False])
def test_dielectrics_under_volpath(variant, rough):
    """volpath.cpp:281-326 with a dielectric / roughdielectric BSDF (no refractive next-event estimation there: a smooth
    interface hides the sun from everything below it, a rough one is sampled by its BSDF with MIS): the general BSDF
    interface of the one-thread-one-path kernel against the oracle."""
    d = scenes.aos_ocean(rough=rough)
    d['integrator']['type'] = 'volpath'
    d['emitter'] = {'type': 'constant', 'radiance': 1.0}
    _compare(S.load_dict(d, variant), 4, 4096, 4096)


@pytest.mark.parametrize("integrator", ["prb_volpathaos"])
def test_other_integrators_refuse_dielectrics(integrator):
    d = scenes.aos_ocean()
    d['integrator']['type'] = integrator
    sc = S.load_dict(d, "llvm_ad_mono_polarized")
    with pytest.raises(RuntimeError, match="dielectric"):
        _lib.DeviceScene(sc, 0)


@pytest.mark.parametrize("name", sorted(n for n in scalar_cases.cases() if n.startswith('ad_rough_dielectric_')))
def test_primal_of_the_reference_adjoint_test_configurations(name):
    """The four active configurations of the fork's own adjoint test (src/integrators/tests/test_ad_volpathaos.py:235-338): a
    directly visible rough dielectric plane with `sample_visible = false` (microfacet.h:246-293: normals drawn from D cos(theta_m),
    roughdielectric.cpp:402-408 for the weight), lit by an area light or a directional one -- primal side, volpathaos.  The oracle
    equals the reference's scalar build on them sample by sample with 0 flips (tests/test_oracle_vs_scalar_ref.py ad_*); the
    adjoint of the interface (gradients w.r.t. eta / alpha) is not built."""
    case = scalar_cases.cases()[name]
    sc = S.load_dict(case['scene'], LLVM[case['variant']])
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    assert dev.forward_kernel() == "thread"
    g = dev.sample_lanes(3, 4096, 0, 4096).astype(np.float64)
    o = orc.sample_lanes(3, 4096, 0, 4096).astype(np.float64)
    bad = (np.abs(g - o) > 1e-4 * np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-9)).any(axis=1)
    print("%s: bit-equal %s, over tolerance %d, non-zero lanes %d" % (name, np.array_equal(g, o), bad.sum(), (o[:, 0] != 0).sum()))
    assert (o[:, 0] != 0).any() and bad.mean() <= 2e-3


def test_eta_and_alpha_are_scene_parameters():
    """mi.traverse exposes the keys the reference's adjoint test perturbs (test_ad_volpathaos.py:238,264: 'plane.bsdf.eta',
    'plane.bsdf.alpha.value'; roughdielectric.cpp:224-236); params.update() re-commits the scene, so the finite-difference side of
    that test runs on this backend.  They are not differentiable here (the adjoint of the interface is not built)."""
    import copy
    import mitsuba3_nasa_b200 as mi
    mi.set_variant('llvm_ad_mono_polarized')
    d = copy.deepcopy(scalar_cases.cases()['ad_rough_dielectric_directional']['scene'])
    scene = mi.load_dict(d)
    p = mi.traverse(scene)
    assert {'plane.bsdf.eta', 'plane.bsdf.alpha.value'} <= set(p.keys())
    assert abs(p['plane.bsdf.eta'] - 1.34) < 1e-6 and abs(p['plane.bsdf.alpha.value'] - 0.1) < 1e-7
    before = np.asarray(mi.render_1(scene, seed=1, spp=4096)).copy()
    p['plane.bsdf.eta'] = 1.5
    p['plane.bsdf.alpha.value'] = 0.2
    p.update()
    after = np.asarray(mi.render_1(scene, seed=1, spp=4096))
    d['plane']['bsdf'].update(int_ior=1.5, alpha=0.2)
    fresh = np.asarray(mi.render_1(mi.load_dict(d), seed=1, spp=4096))
    assert np.array_equal(after, fresh) and not np.array_equal(after, before)
    with pytest.raises(RuntimeError, match="not differentiable"):
        p.enable_grad('plane.bsdf.eta')

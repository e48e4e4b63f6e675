"""GPU parity of volpathmis, the spectral-MIS volumetric path tracer (src/integrators/volpathmis.cpp with
use_spectral_mis, its default), walked by the one-thread-one-path kernel of csrc/dmis.cuh (k_forward MODE 5).

The oracle's volpathmis (oracle/owalkmis.inc) reproduces the reference's scalar build sample by sample on the scenes of
tests/scalar_cases.py mis_* and on the fork's two *_spectral_sigmat_volpathmis render-test scenes (tests/
test_oracle_vs_scalar_ref.py, tests/test_reference_scenes.py); here the kernel must reproduce the oracle lane by lane under
the llvm_ad_* seeding: tolerance 1e-4 of the lane's largest component, at most 0.2 % of the lanes flipped, event counters
equal when no lane flipped.  And the reference's own render tests (src/render/tests/test_renders.py:179-236): its volpathmis
scenes, plus every volpath scene re-rendered with volpathmis as its INTEGRATOR_MAPPING (:44-48) asks, against the reference's
images by its Z-test."""
import os

import numpy as np
import pytest

import mitsuba3_nasa_b200 as mi
from mitsuba3_nasa_b200 import scene as S, scenes, _lib
from orc import OracleScene
import scalar_cases
from test_reference_scenes import load, z_test_against_reference, PM
from test_gpu_mesh import _compare

pytestmark = pytest.mark.gpu
MIS_CASES = [n for n in scalar_cases.cases() if n.startswith('mis_')]
VOLPATH_STEMS = ['test_hetero_sigmat', 'test_hetero_sigmat_albedo', 'test_homo_null_envmap', 'test_homo_dielectric_envmap',
                 'test_homo_roughdielectric_spot', 'test_homo_area_light', 'test_homo_obj_inside_medium', 'test_surf_area_light']
MIS_STEMS = ['test_homo_spectral_sigmat', 'test_hetero_spectral_sigmat']


@pytest.mark.parametrize("name", MIS_CASES)
@pytest.mark.parametrize("variant", ["llvm_ad_rgb", "llvm_ad_mono"])
def test_lane_by_lane(name, variant):
    case = scalar_cases.cases()[name]
    sc = S.load_dict(case['scene'], variant)
    sn = sc.desc.sensor
    spp = max(case['spp'], 16)
    _compare(sc, 7, spp, sn.crop_width * sn.crop_height * spp)


def test_right_to_left_argument_order():
    """rng_arg_order: the order gcc evaluates phase->sample(ctx, mei, next_1d(), next_2d()) in."""
    from mitsuba3_nasa_b200 import _abi as A
    sc = S.load_dict(scalar_cases.cases()['mis_area_emitter']['scene'], 'llvm_ad_rgb')
    sc.desc.integrator.rng_arg_order = A.RNG_ARGS_RIGHT_TO_LEFT
    _compare(sc, 1, 16, 10 * 10 * 16)


def test_dielectric_sphere_and_mesh():
    d = scalar_cases.mesh_scene(hetero=True, res=12)
    d['integrator']['type'] = 'volpathmis'
    d['ball'] = {'type': 'sphere', 'center': [0.0, 0.0, 1.6], 'radius': 0.4, 'bsdf': {'type': 'roughdielectric', 'alpha': 0.2}}
    _compare(S.load_dict(d, "llvm_ad_rgb"), 4, 16, 12 * 12 * 16)


def test_film_equals_the_oracle_film():
    sc = S.load_dict(scalar_cases.cases()['mis_het_rgb_albedo_grid']['scene'], "llvm_ad_rgb")
    img = _lib.DeviceScene(sc, 0).render(1, 16)
    _, ref = OracleScene(sc).render(1, 16)
    assert np.abs(img - ref).max() <= 1e-4 * np.abs(ref).max()


def test_polarized_variants_are_refused():
    d = scenes.c1_homogeneous_cube(res=4, spp=1)
    d['integrator']['type'] = 'volpathmis'
    with pytest.raises(Exception, match="does not support polarized mode"):
        S.load_dict(d, "llvm_ad_mono_polarized")


def _xml(stem):
    f = os.path.join(PM, stem + '_volpathmis.xml')
    return f if os.path.exists(f) else os.path.join(PM, stem + '_volpath.xml')


@pytest.mark.parametrize("stem", MIS_STEMS + VOLPATH_STEMS)
def test_reference_scene_lanes_equal_the_oracle(stem):
    sc = load(_xml(stem), 'llvm_ad_rgb', spp=4, integrator='volpathmis')
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    n = 48 * 48 * 4
    g = dev.sample_lanes(0, 4, 0, n).astype(np.float64)
    o = orc.sample_lanes(0, 4, 0, n).astype(np.float64)
    bad = (np.abs(g - o) > 1e-4 * np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)).any(axis=1)
    print("%s: lanes %d, bit-equal %s, over tolerance %d" % (stem, n, np.array_equal(g, o), bad.sum()))
    assert bad.mean() <= 2e-3


@pytest.mark.parametrize("stem", MIS_STEMS + VOLPATH_STEMS)
def test_render_passes_the_reference_z_test(stem):
    mi.set_variant('llvm_ad_rgb')
    spp = int(2e6) // (48 * 48)
    scene = mi.load_file(_xml(stem), spp=spp, integrator='volpathmis')
    img = np.asarray(mi.render(scene, seed=0, spp=spp))
    ok, pmin, alpha = z_test_against_reference(img, spp, stem)
    print("%s: min p-value %.3g, Sidak level %.3g" % (stem, pmin, alpha))
    assert ok

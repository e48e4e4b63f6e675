"""GPU parity of triangle meshes and the checkerboard texture (`obj` shapes, src/shapes/obj.cpp,
include/mitsuba/render/mesh.h:431-453, src/render/mesh.cpp:1368-1537, src/textures/checkerboard.cpp), walked by the forward
one-thread-one-path kernels (csrc/dmesh.cuh).  Scenes: those of the integrator-level pin (tests/scalar_cases.py mesh_*), on
which the oracle reproduces the reference's scalar build sample by sample; here the kernels must reproduce the oracle lane by
lane under the llvm_ad_* seeding.  Tolerance 1e-4 of the lane's largest component, at most 0.2 % of the lanes flipped."""
import os

import numpy as np
import pytest

from mitsuba3_nasa_b200 import _abi as A, scene as S, scenes, _lib
from orc import OracleScene
import scalar_cases

pytestmark = pytest.mark.gpu
MESHES = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "meshes")


def _compare(sc, seed, spp, count, flip_frac=2e-3):
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    assert dev.forward_kernel() == "thread"
    g = dev.sample_lanes(seed, spp, 0, count).astype(np.float64)
    gc = dev.counters()
    o = orc.sample_lanes(seed, spp, 0, count).astype(np.float64)
    oc = orc.counters()
    scale = np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)
    bad = (np.abs(g - o) / scale > 1e-4).any(axis=1)
    print("lanes %d: bit-equal %s, over tolerance %d" % (count, np.array_equal(g, o), bad.sum()))
    assert bad.mean() <= flip_frac, "%d of %d lanes differ" % (bad.sum(), count)
    if not bad.any():
        for k in ('n_real', 'n_null', 'n_nee_steps', 'n_surface', 'n_lookups', 'n_samples'):
            assert gc[k] == oc[k], (k, gc[k], oc[k])


@pytest.mark.parametrize("variant", ["llvm_ad_rgb", "llvm_ad_mono", "llvm_ad_mono_polarized"])
@pytest.mark.parametrize("hetero", [False, True])
def test_mesh_scene_lane_by_lane(variant, hetero):
    sc = S.load_dict(scalar_cases.mesh_scene(hetero=hetero, res=16), variant)
    _compare(sc, 2, 16, 16 * 16 * 16)


def test_mesh_scene_volpathaos_polarized_with_sun():
    d = scalar_cases.mesh_scene(res=12)
    d['integrator']['type'] = 'volpathaos'
    d['emitter'] = {'type': 'directional', 'direction': [0.3, 0.2, -1.0], 'irradiance': 2.0}
    d['medium1']['phase'] = {'type': 'rayleigh'}
    _compare(S.load_dict(d, "llvm_ad_mono_polarized"), 5, 16, 12 * 12 * 16)


def test_film_of_a_mesh_scene_equals_the_oracle_film():
    sc = S.load_dict(scalar_cases.mesh_scene(res=16), "llvm_ad_rgb")
    img = _lib.DeviceScene(sc, 0).render(1, 16)
    _, ref = OracleScene(sc).render(1, 16)
    assert np.abs(img - ref).max() <= 1e-4 * np.abs(ref).max()


def test_checkerboard_on_a_rectangle():
    d = scenes.c1_homogeneous_cube(res=12, spp=8, max_depth=8)
    from mitsuba3_nasa_b200.transforms import ScalarTransform4f as T
    d['floor'] = {'type': 'rectangle', 'to_world': T.translate([0, 0, -1.5]) @ T.scale(4.0),
                  'bsdf': {'type': 'diffuse', 'reflectance': {'type': 'checkerboard', 'color0': 0.1, 'color1': 0.9, 'to_uv': T.scale([6, 6, 6])}}}
    _compare(S.load_dict(d, "llvm_ad_rgb"), 0, 8, 12 * 12 * 8)


@pytest.mark.parametrize("name,variant", [("envmap_homog_rgb", "llvm_ad_rgb"), ("envmap_homog_rgb", "llvm_ad_mono_polarized"),
                                          ("envmap_mesh_mono", "llvm_ad_mono"), ("envmap_mesh_mono", "llvm_ad_rgb")])
def test_environment_map_lane_by_lane(name, variant):
    """src/emitters/envmap.cpp: importance sampling through the hierarchical luminance warp, bilinear texel look-up, MIS of the
    rays that leave the scene.  The oracle reproduces the reference's scalar build on these scenes sample by sample
    (tests/test_oracle_vs_scalar_ref.py envmap_*)."""
    case = scalar_cases.cases()[name]
    sc = S.load_dict(case['scene'], variant)
    sn = sc.desc.sensor
    _compare(sc, 3, 16, sn.crop_width * sn.crop_height * 16)


def test_environment_map_in_volpathaos():
    d = scalar_cases.cases()["envmap_homog_rgb"]['scene']
    d['integrator']['type'] = 'volpathaos'
    d['medium1']['phase'] = {'type': 'rayleigh'}
    _compare(S.load_dict(d, "llvm_ad_mono_polarized"), 1, 16, 12 * 12 * 16)


@pytest.mark.parametrize("integrator", ["prbvolpath"])
def test_adjoint_integrators_refuse_meshes(integrator):
    d = scalar_cases.mesh_scene(res=8)
    d['integrator']['type'] = integrator
    with pytest.raises(RuntimeError, match="meshes"):
        _lib.DeviceScene(S.load_dict(d, "llvm_ad_rgb"), 0)

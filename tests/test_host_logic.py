"""Host-side mirror of the reference's scene interface: mi.load_dict flattening, plugin/property
errors, parameter keys, transforms, the .vol container.  No GPU and no oracle involved.

Reference behaviour followed: src/core/python/xml_v.cpp:112-146 (load_dict), src/core/plugin.cpp:174-211
(unknown plugin), src/core/xml.cpp (unreferenced property error), src/render/volumegrid.cpp:29-123 and
src/render/tests/test_volumegrid.py:8-50 (.vol round trip), src/volumes/grid.cpp:156-299 (gridvolume
parameters), include/mitsuba/core/transform.h (look_at / perspective conventions)."""
import os

import numpy as np
import pytest

from mitsuba3_nasa_b200 import _abi as A
from mitsuba3_nasa_b200 import scene as S
from mitsuba3_nasa_b200 import scenes
from mitsuba3_nasa_b200.transforms import ScalarTransform4f as T
from mitsuba3_nasa_b200.volumegrid import VolumeGrid


def test_c1_flattening():
    f = S.load_dict(scenes.c1_homogeneous_cube(res=32, spp=16), "llvm_ad_rgb")
    d = f.desc
    assert d.abi_version == A.MIB_ABI_VERSION and d.spectrum == A.SPEC_RGB
    assert d.n_shapes == 1 and d.n_media == 1 and d.n_emitters == 1 and d.n_bsdfs == 1
    assert d.shapes[0].type == A.SHAPE_CUBE and d.shapes[0].interior_medium == 0 and d.shapes[0].exterior_medium == -1
    m = d.media[0]
    assert m.type == A.MEDIUM_HOMOGENEOUS and m.sigma_t.type == A.VOLUME_CONST and abs(m.sigma_t.value[0] - 2.0) < 1e-7
    assert m.albedo.channels == 3 and np.allclose(list(m.albedo.value), [0.9, 0.8, 0.7])
    assert d.phases[m.phase].type == A.PHASE_HG and abs(d.phases[m.phase].g - 0.7) < 1e-7
    assert d.integrator.type == A.INTEG_VOLPATH and d.integrator.max_depth == 64 and d.integrator.rr_depth == 5
    assert d.sensor.crop_width == 32 and f.sample_count == 16


def test_defaults_follow_the_reference():
    """medium defaults: albedo 0.75, sigma_t 1, scale 1, isotropic phase, sample_emitters (medium.cpp:12-32,
    homogeneous.cpp:137-147); prb integrators: max_depth 6, rr_depth 5 (common.py:32-41)."""
    d = scenes.c1_homogeneous_cube(res=8, spp=4)
    d['medium1'] = {'type': 'homogeneous'}
    d['integrator'] = {'type': 'prbvolpath'}
    f = S.load_dict(d, "llvm_ad_mono")
    m = f.desc.media[0]
    assert abs(m.albedo.value[0] - 0.75) < 1e-7 and abs(m.sigma_t.value[0] - 1.0) < 1e-7 and m.scale == 1.0
    assert f.desc.phases[m.phase].type == A.PHASE_ISOTROPIC and m.sample_emitters == 1 and m.has_spectral_extinction == 1
    assert f.desc.integrator.max_depth == 6 and f.desc.integrator.rr_depth == 5


def test_forward_reference_and_unknown_reference():
    d = scenes.c1_homogeneous_cube(res=8, spp=4)
    # 'object' refers to 'medium1', move the medium after the object (forward reference)
    med = d.pop('medium1')
    d['zz_medium'] = med
    d['object']['interior'] = {'type': 'ref', 'id': 'zz_medium'}
    f = S.load_dict(d, "llvm_ad_rgb")
    assert f.desc.shapes[0].interior_medium == 0
    d['object']['interior'] = {'type': 'ref', 'id': 'nope'}
    with pytest.raises(S.PluginError, match="unknown id"):
        S.load_dict(d, "llvm_ad_rgb")


def test_unknown_plugin_and_unreferenced_property():
    d = scenes.c1_homogeneous_cube(res=8, spp=4)
    d['integrator'] = {'type': 'path'}
    with pytest.raises(S.PluginError, match="not found"):
        S.load_dict(d, "llvm_ad_rgb")
    d = scenes.c1_homogeneous_cube(res=8, spp=4)
    d['medium1']['sigma_tt'] = 3.0
    with pytest.raises(S.PluginError, match="unreferenced propert"):
        S.load_dict(d, "llvm_ad_rgb")


def test_parameter_validation_messages():
    d = scenes.c1_homogeneous_cube(res=8, spp=4)
    d['medium1']['phase'] = {'type': 'hg', 'g': 1.0}
    with pytest.raises(S.PluginError, match="asymmetry parameter"):
        S.load_dict(d, "llvm_ad_rgb")
    d = scenes.c1_homogeneous_cube(res=8, spp=4)
    d['integrator']['rr_depth'] = 0
    with pytest.raises(S.PluginError, match="rr_depth"):
        S.load_dict(d, "llvm_ad_rgb")
    d = scenes.c4_cloud_adjoint(res=8, spp=4, grid_res=8, polarized=True, phase={'type': 'rayleigh'})
    with pytest.raises(S.PluginError, match="polarized"):
        S.load_dict(d, "llvm_ad_mono")


def test_fast_mode_property():
    """Backend extension: integrator property 'majorant_resolution' (0 = the reference's global majorant)."""
    d = scenes.c3_cloud(res=8, spp=4, grid_res=8, phase={'type': 'isotropic'})
    assert S.load_dict(d, "llvm_ad_mono").desc.integrator.majorant_res == 0
    d = scenes.c3_cloud(res=8, spp=4, grid_res=8, phase={'type': 'isotropic'}, majorant_resolution=16)
    assert S.load_dict(d, "llvm_ad_mono").desc.integrator.majorant_res == 16
    d['integrator']['majorant_resolution'] = 1000
    with pytest.raises(S.PluginError, match="majorant_resolution"):
        S.load_dict(d, "llvm_ad_mono")


def test_gridvolume_layout_and_param_keys():
    grid = np.arange(2 * 3 * 4, dtype=np.float32).reshape(2, 3, 4, 1)      # [z, y, x, c]
    f = S.load_dict(scenes.c3_cloud(res=8, spp=4, grid=grid, phase={'type': 'isotropic'}), "llvm_ad_mono")
    v = f.desc.media[0].sigma_t
    assert v.type == A.VOLUME_GRID and list(v.res) == [4, 3, 2] and v.channels == 1
    assert np.array_equal(f.buffers[v.buffer], grid.reshape(-1))
    assert np.allclose(list(v.bbox_min), [-1, -1, -1]) and np.allclose(list(v.bbox_max), [1, 1, 1])
    keys = set(f.params)
    assert 'object.interior_medium.sigma_t.data' in keys
    assert 'object.interior_medium.albedo.value.value' in keys
    assert 'object.interior_medium.scale' in keys
    # scale chosen so that the majorant optical depth across the cube is 20
    assert abs(f.desc.media[0].scale * grid.max() * 2.0 - 20.0) < 1e-4


def test_vol_file_round_trip(tmp_path):
    """src/render/tests/test_volumegrid.py:8-50."""
    rng = np.random.default_rng(0)
    a = rng.random((4, 8, 16, 3)).astype(np.float32)
    g = VolumeGrid(a)
    assert np.allclose(np.asarray(g), a) and np.isclose(g.max(), a.max())
    assert g.size() == [16, 8, 4] and g.channel_count() == 3
    assert VolumeGrid(a, False).max() == 0.0
    path = os.path.join(str(tmp_path), "out.vol")
    g.write(path)
    assert os.path.getsize(path) == 3 + 1 + 5 * 4 + 6 * 4 + a.size * 4
    h = VolumeGrid(path)
    assert np.array_equal(np.asarray(h), a) and h.size() == [16, 8, 4] and h.channel_count() == 3
    assert np.allclose(h.max_per_channel(), [a[..., c].max() for c in range(3)])
    with open(path, "r+b") as fh:
        fh.seek(3)
        fh.write(b"\x02")
    with pytest.raises(RuntimeError, match="version 3"):
        VolumeGrid(path)
    with open(path, "r+b") as fh:
        fh.write(b"XOL")
    with pytest.raises(RuntimeError, match="Invalid volume file"):
        VolumeGrid(path)


def test_gridvolume_from_file_and_grid_bbox(tmp_path):
    a = np.random.default_rng(1).random((3, 3, 3)).astype(np.float32)
    g = VolumeGrid(a)
    g.set_bbox([-2, -2, -2], [2, 2, 2])
    path = os.path.join(str(tmp_path), "g.vol")
    g.write(path)
    d = scenes.c3_cloud(res=8, spp=4, grid=a[..., None], phase={'type': 'isotropic'})
    d['object']['interior']['sigma_t'] = {'type': 'gridvolume', 'filename': path, 'use_grid_bbox': True}
    f = S.load_dict(d, "llvm_ad_mono")
    v = f.desc.media[0].sigma_t
    assert np.allclose(list(v.bbox_min), [-2, -2, -2]) and np.allclose(list(v.bbox_max), [2, 2, 2])
    assert np.array_equal(f.buffers[v.buffer], a.reshape(-1))
    d['object']['interior']['sigma_t'] = {'type': 'gridvolume', 'data': a[..., None], 'use_grid_bbox': True}
    with pytest.raises(S.PluginError, match="use_grid_bbox"):
        S.load_dict(d, "llvm_ad_mono")


def test_transform_conventions():
    m = T.look_at(origin=[0, 0, 5], target=[0, 0, 0], up=[0, 1, 0])
    assert np.allclose(m @ [0, 0, 0], [0, 0, 5])                      # camera origin
    assert np.allclose((m @ [0, 0, 1]) - (m @ [0, 0, 0]), [0, 0, -1])  # +z looks at the target
    r = T.rotate([0, 0, 1], 90)
    assert np.allclose(r @ [1, 0, 0], [0, 1, 0], atol=1e-12)
    assert np.allclose((T.translate([1, 2, 3]) @ T.scale(2.0)).inverse() @ [3, 6, 9], [1, 2, 3])


def test_variants_and_seed_helper():
    import mitsuba3_nasa_b200 as mi
    with pytest.raises(ImportError):
        mi.set_variant('cuda_ad_spectral')
    mi.set_variant('scalar_spectral', 'llvm_ad_mono_polarized')
    assert mi.variant() == 'llvm_ad_mono_polarized'
    assert mi.sample_tea_32(1, 1)[1] >> 9 | 0x3f800000 == np.array([0.5424730777740479 + 1], np.float32).view(np.uint32)[0]


def test_film_wrapper_writes_the_last_developed_image(tmp_path):
    """api.Film (hdrfilm.cpp: size / bitmap / write) on a stand-in scene -- no device involved.  component_format defaults to
    float16 as in the reference; the file is PIZ-compressed like Bitmap::write's output."""
    from types import SimpleNamespace
    from mitsuba3_nasa_b200 import api
    from mitsuba3_nasa_b200.exr import read_exr
    flat = S.load_dict(scenes.c1_homogeneous_cube(res=16, spp=4), "llvm_ad_rgb")
    assert flat.film_component_format == 'float32'                 # the C1 scene asks for float32
    d = scenes.c1_homogeneous_cube(res=16, spp=4)
    del d['sensor']['film']['component_format']
    flat16 = S.load_dict(d, "llvm_ad_rgb")
    assert flat16.film_component_format == 'float16'
    d['sensor']['film']['component_format'] = 'float8'
    with pytest.raises(S.PluginError, match='component_format'):
        S.load_dict(d, "llvm_ad_rgb")
    img = np.random.default_rng(0).uniform(0, 2, (16, 16, 3)).astype(np.float32)
    for f, want in ((flat, img), (flat16, img.astype(np.float16).astype(np.float32))):
        fake = SimpleNamespace(flat=f, _last_image=None)
        sensor = api.Sensor(fake)
        film = sensor.film()
        assert film.size() == (16, 16) and film.crop_size() == (16, 16) and sensor.crop_width == 16
        with pytest.raises(RuntimeError):
            film.bitmap()
        fake._last_image = img
        p = str(tmp_path / "film.exr")
        film.write(p)
        got, names = read_exr(p)
        assert names == ["B", "G", "R"] and np.array_equal(got[..., ::-1], want)

"""The fork's own render tests on the CUDA path (src/render/tests/test_renders.py:179-236): the
participating-media scenes this backend covers (eight of the nine `volpath` ones) -- `obj` meshes, `envmap`, `checkerboard`, heterogeneous `.vol` grids, a smooth
`dielectric` and a `roughdielectric` sphere filled with a medium, a `spot` light, a mesh area light behind a
`thindielectric` shell, `plastic` and `roughplastic` objects, the `moment` wrapper -- loaded from their XML files with mi.load_file, rendered at the reference's budget of 2e6 samples per image and
compared with the reference's own images by its own Z-test.  Plus lane-by-lane parity with the oracle on the same scenes."""
import os

import numpy as np
import pytest

import mitsuba3_nasa_b200 as mi
from mitsuba3_nasa_b200 import scene as S, _lib
from orc import OracleScene
import scalar_cases
from test_reference_scenes import load, z_test_against_reference, PM

pytestmark = pytest.mark.gpu
STEMS = ['test_hetero_sigmat', 'test_hetero_sigmat_albedo', 'test_homo_null_envmap', 'test_homo_dielectric_envmap', 'test_homo_roughdielectric_spot', 'test_homo_area_light',
         'test_homo_obj_inside_medium', 'test_surf_area_light']


@pytest.mark.parametrize("stem", STEMS)
def test_render_passes_the_reference_z_test(stem):
    mi.set_variant('llvm_ad_rgb')
    spp = int(2e6) // (48 * 48)
    scene = mi.load_file(os.path.join(PM, stem + '_volpath.xml'), spp=spp)
    img = np.asarray(mi.render(scene, seed=0, spp=spp))
    ok, pmin, alpha = z_test_against_reference(img, spp, stem)
    print("%s: min p-value %.3g, Sidak level %.3g" % (stem, pmin, alpha))
    assert ok


@pytest.mark.parametrize("stem", STEMS)
def test_lanes_equal_the_oracle(stem):
    sc = load(os.path.join(PM, stem + '_volpath.xml'), 'llvm_ad_rgb', spp=4)
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    n = 48 * 48 * 4
    g = dev.sample_lanes(0, 4, 0, n).astype(np.float64)
    o = orc.sample_lanes(0, 4, 0, n).astype(np.float64)
    bad = (np.abs(g - o) > 1e-4 * np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)).any(axis=1)
    print("%s: lanes %d, bit-equal %s, over tolerance %d" % (stem, n, np.array_equal(g, o), bad.sum()))
    assert bad.mean() <= 2e-3


def test_the_reference_test_body_runs_as_written(tmp_path):
    """test_renders.py:205-213 with `mi` = this package: load_file(scene, spp=, integrator=), integrator().render(scene, seed=0,
    spp=, develop=True), sensors()[0].film().bitmap(raw=False) -- then its Z-test; and film().write() gives an EXR that reads back."""
    from mitsuba3_nasa_b200 import exr
    mi.set_variant('llvm_ad_rgb')
    stem = 'test_homo_null_envmap'
    spp = int(2e6) // (48 * 48)
    scene = mi.load_file(os.path.join(PM, stem + '_volpath.xml'), spp=spp, integrator='volpath')
    scene.integrator().render(scene, seed=0, spp=spp, develop=True)
    film = scene.sensors()[0].film()
    img = np.asarray(film.bitmap(raw=False))
    assert tuple(film.size()) == (48, 48) and img.shape[:2] == (48, 48)
    ok, pmin, alpha = z_test_against_reference(img, spp, stem)
    assert ok, (pmin, alpha)
    out = os.path.join(str(tmp_path), "render.exr")
    film.write(out)
    back, names = exr.read_exr(out)
    rgb = back[..., [names.index(c) for c in 'RGB']]
    assert np.allclose(rgb, img[..., :3], rtol=1e-6, atol=1e-7)

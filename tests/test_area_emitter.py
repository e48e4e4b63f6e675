"""Area emitters (src/emitters/area.cpp:83-186 on rectangles, Shape::sample_direction / pdf_direction shape.cpp:358-388,
Rectangle::sample_position rectangle.cpp:166-180) in volpath.

CPU: the oracle's two estimators of the same integral agree -- with emitter sampling (NEE + MIS against the light's
pdf_direction) and without (light found by hitting it only); this pins the pdf / eval / one-sidedness against each other.
GPU: lane-by-lane parity of both kernel families with the oracle."""
import numpy as np
import pytest

from mitsuba3_nasa_b200 import scene as S, scenes
from mitsuba3_nasa_b200.transforms import ScalarTransform4f as T
from orc import OracleScene


def area_scene(res=8, spp=64, sample_emitters=True, second_emitter=False, grid=None):
    d = scenes.c1_homogeneous_cube(res=res, spp=spp, max_depth=12) if grid is None else \
        scenes.c3_cloud(res=res, spp=spp, max_depth=12, grid=grid, optical_depth=4.0, phase={'type': 'hg', 'g': 0.5})
    del d['emitter']
    # a 3 x 3 light at x = 3 facing the medium (-x), outside the camera's view: reached by NEE and by scattered paths
    d['light'] = {'type': 'rectangle', 'to_world': T.translate([3, 0, 0]) @ T.rotate([0, 1, 0], -90) @ T.scale(1.5),
                  'bsdf': {'type': 'diffuse', 'reflectance': 0.2},
                  'emitter': {'type': 'area', 'radiance': 4.0}}
    if second_emitter:
        d['sky'] = {'type': 'constant', 'radiance': 0.25}
    med = d['medium1'] if 'medium1' in d else d['object']['interior']
    med['sample_emitters'] = sample_emitters
    return d


def test_flattening_and_validation():
    sc = S.load_dict(area_scene(), 'llvm_ad_rgb')
    e = sc.desc.emitters[0]
    sh = sc.desc.shapes[e.shape]
    assert sc.desc.n_emitters == 1 and sh.emitter == 0 and abs(sh.inv_surface_area - 1 / 9.0) < 1e-6
    assert np.allclose(sh.frame_n[:], [-1, 0, 0], atol=1e-6)
    assert 'light.emitter.radiance.value' in sc.params
    d = area_scene()
    d['light']['type'] = 'sphere'
    with pytest.raises(S.PluginError, match="rectangle"):
        S.load_dict(d, 'llvm_ad_rgb')


@pytest.mark.parametrize("second", [False, True])
def test_oracle_nee_and_hit_estimators_agree(second):
    means = []
    for se in (True, False):
        sc = S.load_dict(area_scene(res=6, spp=4000, sample_emitters=se, second_emitter=second), 'llvm_ad_mono')
        raw, _ = OracleScene(sc).render(3 if se else 4, 4000, nthreads=8)
        raw = np.asarray(raw, dtype=np.float64)
        means.append((raw[..., 0].sum() / raw[..., -1].sum(), raw[..., 0].std()))
    a, b = means[0][0], means[1][0]
    assert a > 0.02 and b > 0.02
    assert abs(a - b) <= 0.04 * max(a, b), (a, b)         # 144k samples each; the hit-only estimator is the noisy one


@pytest.mark.gpu
@pytest.mark.parametrize("variant", ["llvm_ad_rgb", "llvm_ad_mono", "llvm_ad_mono_polarized"])
@pytest.mark.parametrize("kernel", ["thread", "pool"])
def test_gpu_lanes_match_the_oracle(variant, kernel, monkeypatch):
    from mitsuba3_nasa_b200 import _lib
    monkeypatch.setenv('MIB_FORWARD_KERNEL', kernel)
    d = area_scene(res=12, spp=16, second_emitter=(variant == "llvm_ad_mono"))
    sc = S.load_dict(d, variant)
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    n = 12 * 12 * 16
    g = dev.sample_lanes(5, 16, 0, n).astype(np.float64)
    o = orc.sample_lanes(5, 16, 0, n).astype(np.float64)
    scale = np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)
    bad = (np.abs(g - o) / scale > 1e-4).any(axis=1)
    assert np.abs(o).max() > 0 and bad.mean() <= 5e-3, "lanes over tolerance: %d of %d" % (bad.sum(), n)
    gc, oc = dev.counters(), orc.counters()
    for k in ('n_real', 'n_nee_steps', 'n_surface'):
        assert abs(gc[k] - oc[k]) <= 5e-3 * max(oc[k], 1)


@pytest.mark.gpu
def test_gpu_heterogeneous_cloud_lit_by_an_area_light():
    from mitsuba3_nasa_b200 import _lib
    sc = S.load_dict(area_scene(res=12, spp=16, grid=scenes.cloud_grid(16)), 'llvm_ad_mono')
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    n = 12 * 12 * 16
    g = dev.sample_lanes(1, 16, 0, n).astype(np.float64)
    o = orc.sample_lanes(1, 16, 0, n).astype(np.float64)
    bad = (np.abs(g - o) / np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6) > 1e-4).any(axis=1)
    assert np.abs(o).max() > 0 and bad.mean() <= 5e-3

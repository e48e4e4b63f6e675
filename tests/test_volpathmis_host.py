"""volpathmis on the host side: the plugin's properties (src/integrators/volpathmis.cpp:86-112), the
oracle's restatement against the reference's own render test, and the properties of the weight matrices that do not need a
reference (a grey medium makes every row of p_over_f equal, so volpathmis and volpath estimate the same image)."""
import os

import numpy as np
import pytest

from mitsuba3_nasa_b200 import _abi as A, scene as S, scenes
from mitsuba3_nasa_b200.scene import PluginError
from orc import OracleScene
from test_reference_scenes import load, z_test_against_reference, PM


def _cube(**kw):
    d = scenes.c1_homogeneous_cube(res=6, spp=4)
    d['integrator'].update(type='volpathmis', **kw)
    return d


def test_integrator_type_and_defaults():
    sc = S.load_dict(_cube(max_depth=7, rr_depth=3), 'llvm_ad_rgb')
    it = sc.desc.integrator
    assert (it.type, it.max_depth, it.rr_depth) == (A.INTEG_VOLPATHMIS, 7, 3)
    assert S.load_dict(_cube(use_spectral_mis=True), 'llvm_ad_mono').desc.integrator.type == A.INTEG_VOLPATHMIS


def test_polarized_variant_is_refused_like_the_reference():
    with pytest.raises(PluginError, match="does not support polarized mode"):      # volpathmis.cpp:134-136
        S.load_dict(_cube(), 'llvm_ad_mono_polarized')


def test_use_spectral_mis_reaches_the_descriptor():
    assert S.load_dict(_cube(), 'llvm_ad_rgb').desc.integrator.spectral_mis == 1                     # volpathmis.cpp:93
    assert S.load_dict(_cube(use_spectral_mis=False), 'llvm_ad_rgb').desc.integrator.spectral_mis == 0


def test_grey_medium_same_expectation_as_volpath():
    d = scenes.c1_homogeneous_cube(res=4, spp=1, max_depth=8)
    a = OracleScene(S.load_dict(d, 'llvm_ad_rgb')).sample_lanes(0, 4096, 0, 16 * 4096).astype(np.float64)
    d['integrator']['type'] = 'volpathmis'
    b = OracleScene(S.load_dict(d, 'llvm_ad_rgb')).sample_lanes(0, 4096, 0, 16 * 4096).astype(np.float64)
    ma, mb = a[:, :3].mean(), b[:, :3].mean()
    se = np.hypot(a[:, :3].mean(axis=1).std(), b[:, :3].mean(axis=1).std()) / np.sqrt(len(a))
    print("volpath %.5f volpathmis %.5f (standard error %.5f)" % (ma, mb, se))
    assert abs(ma - mb) < 4 * se


def test_oracle_passes_the_reference_render_test_of_volpathmis():
    stem, spp = 'test_homo_spectral_sigmat', int(2e6) // (48 * 48)
    sc = load(os.path.join(PM, stem + '_volpathmis.xml'), 'llvm_ad_rgb', spp=spp)
    assert sc.desc.integrator.type == A.INTEG_VOLPATHMIS
    _, img = OracleScene(sc).render(0, spp)
    ok, pmin, alpha = z_test_against_reference(img, spp, stem)
    print("min p-value %.3g, Sidak level %.3g" % (pmin, alpha))
    assert ok

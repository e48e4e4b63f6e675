"""The reference's statistical image comparison (mitsuba3_nasa_b200/ztest.py <- src/render/tests/test_renders.py:159-229):
known values of the Z-test, the Sidak-corrected acceptance rule, and an end-to-end use on CPU -- two oracle renders of one
scene with different seeds must agree, a render of a changed scene must not.  Reference images of the fork's test-suite go
through the same function (read with read_exr; skipped without the reference checkout)."""
import os

import numpy as np
import pytest

from mitsuba3_nasa_b200 import scene as S, scenes
from mitsuba3_nasa_b200.exr import read_exr
from mitsuba3_nasa_b200.ztest import z_test, sidak_alpha, images_agree
from orc import OracleScene


def test_z_test_known_values():
    # z = |m - r| sqrt(n / var): 0 -> p = 1; 1.959964 -> p = 0.05; 2.575829 -> p = 0.01
    ref, var = np.zeros((1, 3)), np.full((1, 3), 4.0)
    mean = np.array([[0.0, 1.959964 * 2 / 10, 2.575829 * 2 / 10]])
    p = z_test(mean, 100, ref, var)
    assert np.allclose(p, [[1.0, 0.05, 0.01]], atol=1e-6)
    # variances are floored at 1e-4 (test_renders.py:162)
    assert np.allclose(z_test(np.array([0.01]), 1, np.array([0.0]), np.array([0.0])), z_test(np.array([0.01]), 1, np.array([0.0]), np.array([1e-4])))
    assert np.isclose(sidak_alpha(0.01, 1), 0.01) and np.isclose(sidak_alpha(0.01, 2304), 1 - 0.99 ** (1 / 2304))


def test_acceptance_rule():
    rng = np.random.default_rng(0)
    ref = rng.uniform(0.1, 1, (48, 48, 3))
    var = rng.uniform(0.01, 0.1, (48, 48, 3))
    spp = 800
    good = ref + rng.normal(size=ref.shape) * np.sqrt(var / spp)
    ok, pmin, alpha = images_agree(good, spp, ref, var)
    assert ok and alpha < 1e-5
    biased = good * 1.05                                            # a 5 % brightness error at 800 spp is far outside the noise
    assert not images_agree(biased, spp, ref, var)[0]
    few = good.copy()
    few[:2, :2] += 10.0                                             # 4 broken pixels of 2304 (< 0.25 %) are tolerated, as in the reference
    assert images_agree(few, spp, ref, var)[0]
    few[:4, :4] += 10.0                                             # 16 are not
    assert not images_agree(few, spp, ref, var)[0]


def test_two_oracle_renders_agree_and_a_changed_scene_does_not():
    spp = 64
    sc = S.load_dict(scenes.c1_homogeneous_cube(res=12, spp=spp, max_depth=8), "llvm_ad_rgb")
    orc = OracleScene(sc)
    # reference + per-sample variance from many small renders (the `moment` integrator's role)
    renders = np.stack([orc.render(100 + k, spp)[1] for k in range(24)])
    ref, var = renders.mean(axis=0), renders.var(axis=0, ddof=1) * spp
    img = orc.render(7, spp)[1]
    ok, pmin, alpha = images_agree(img, spp, ref, var)
    assert ok, (pmin, alpha)
    d = scenes.c1_homogeneous_cube(res=12, spp=spp, max_depth=8)
    d['medium1']['sigma_t'] = 6.0
    other = OracleScene(S.load_dict(d, "llvm_ad_rgb")).render(7, spp)[1]
    assert not images_agree(other, spp, ref, var)[0]


REFS = "/root/reference/resources/data/tests/scenes/participating_media/refs"


@pytest.mark.skipif(not os.path.isdir(REFS), reason="reference checkout absent (build container only)")
def test_reference_images_pass_against_themselves_and_fail_against_each_other():
    a, _ = read_exr(os.path.join(REFS, "test_hetero_sigmat_ref_rgb.exr"))
    a_var, _ = read_exr(os.path.join(REFS, "test_hetero_sigmat_ref_rgb_var.exr"))
    b, _ = read_exr(os.path.join(REFS, "test_hetero_sigmat_albedo_ref_rgb.exr"))
    spp = int(2e6) // (48 * 48)                                     # the sample budget of test_renders.py:205-208
    assert a_var.shape == a.shape                                   # colour-converted variances dip below zero: hence the floor
    noisy = a + np.random.default_rng(1).normal(size=a.shape) * np.sqrt(np.maximum(a_var, 1e-4) / spp)
    assert images_agree(noisy, spp, a, a_var)[0]
    assert not images_agree(b, spp, a, a_var)[0]

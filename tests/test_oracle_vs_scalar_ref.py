"""Integrator-level pin of the oracle on the reference's OWN integrators.

tests/golden/scalar_ref.npz holds what the unmodified volpath.cpp / volpathaos.cpp (+ media, phase functions, volumes,
shapes, emitters, sensors, sampler) of /root/reference return, sample by sample, in a patched *scalar* build
(oracle/build_ref_scalar.sh: scalar_rgb, scalar_mono, scalar_mono_polarized; generator: tests/golden/make_scalar_golden.py).
Here the oracle runs the same scenes with the scalar variants' seeding (one PCG32 stream per pixel, sampler.cpp:131-133,
integrator.cpp:1103-1123) and must reproduce every sample.

Tolerance: 1e-4 relative to the largest value of the case, at most 0.5 % of the samples allowed to "flip" (a float
comparison decided the other way by a last-bit difference; the scalar build sums dot products left to right and is built
without fma contraction where the oracle follows the llvm variants' fma chains).  Measured: see the table this test prints.

The scalar reference is built with gcc, which evaluates `phase->sample(ctx, mei, next_1d(), next_2d())` right to left
(SURVEY F7), so the oracle runs with rng_arg_order = RIGHT_TO_LEFT here; the kernels' default stays left to right (the order
of the Python integrators and of clang builds).  Likewise the scalar variants re-project a rectangle hit onto the rectangle
(rectangle.cpp:549-553) where the differentiable variants -- the kernels' parity target -- keep p = ray(t); the oracle has
a test-only switch for that branch (orc_set_scalar_shapes), without which the distant sensor of C2 (ray origin 2.8e4 units
away) places the entry point differently in the last bits and every path through the 1-unit slab differs."""
import os

import numpy as np
import pytest

from mitsuba3_nasa_b200 import _abi as A, scene as S
from orc import OracleScene
import scalar_cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scalar_ref.npz")
LLVM = {'scalar_rgb': 'llvm_ad_rgb', 'scalar_mono': 'llvm_ad_mono', 'scalar_mono_polarized': 'llvm_ad_mono_polarized'}
CASES = scalar_cases.cases()


def _oracle(case):
    sc = S.load_dict(case['scene'], LLVM[case['variant']])
    sc.desc.integrator.rng_arg_order = A.RNG_ARGS_RIGHT_TO_LEFT
    o = OracleScene(sc)
    o.lib.orc_set_scalar_shapes(1)      # rectangle.cpp:549-553: the branch only the non-differentiable variants take
    return sc, o


@pytest.fixture(autouse=True)
def _restore_shape_mode():
    yield
    from orc import oracle_lib
    oracle_lib().orc_set_scalar_shapes(0)


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


@pytest.mark.parametrize("name", [n for n, c in CASES.items() if c['kind'] == 'samples'])
def test_samples_equal_the_reference(name, golden):
    case = CASES[name]
    sc, o = _oracle(case)
    W, H = sc.desc.sensor.crop_width, sc.desc.sensor.crop_height
    pix, seeds = scalar_cases.pixel_seeds(W, H, case['seed'])
    got = o.sample_pixels_scalar(pix, seeds, case['spp']).astype(np.float64)
    ref = golden[name].astype(np.float64)
    nch = {'scalar_rgb': 3, 'scalar_mono': 1, 'scalar_mono_polarized': 4}[case['variant']]
    got, ref = got[..., :nch], ref[..., :nch]
    assert got.shape == ref.shape
    assert np.array_equal(np.isnan(got), np.isnan(ref))        # (the reference's own NaN samples, SURVEY F4, are reproduced)
    got, ref = np.nan_to_num(got), np.nan_to_num(ref)
    scale = np.abs(ref).max()
    err = np.abs(got - ref).max(axis=-1) / scale
    flips = err > 1e-4
    print("%-32s samples %6d  max rel err (non-flipped) %.2e  flips %d" % (name, err.size, err[~flips].max(), flips.sum()))
    assert scale > 0
    assert flips.mean() <= 5e-3, "%d of %d samples differ from the reference" % (flips.sum(), flips.size)


@pytest.mark.parametrize("name", [n for n, c in CASES.items() if c['kind'] == 'render1'])
def test_render_1_equals_the_reference(name, golden):
    """SamplingIntegrator::render_1 (integrator.cpp:399-743), 2e4 samples, one stream."""
    case = CASES[name]
    sc, o = _oracle(case)
    s = o.sample_pixels_scalar([0], [case['seed']], case['spp']).astype(np.float64)[0]
    stokes = s.sum(axis=0) / case['spp']
    ref = golden[name].reshape(4, 4)[:, 0].astype(np.float64)
    print("render_1 reference", ref, "oracle", stokes)
    assert np.abs(stokes - ref).max() <= 1e-4 * np.abs(ref).max()

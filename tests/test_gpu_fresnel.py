"""Device Fresnel and microfacet blocks (csrc/dfresnel.cuh, csrc/dmicrofacet.cuh, through the C ABI: mib_debug_fresnel,
mib_debug_microfacet) against the oracle restatement (oracle/ofresnel.h, oracle/omicrofacet.h), which tests/test_oracle_vs_ref.py pins on the reference's own headers.  Tolerance: bit-exact (the same
fused multiply-adds, IEEE division and square root on both sides), as for the device math in test_gpu_parity_forward.py.

First GPU run: round 2 driver run GPUTEST_r01 (11 XPASS); the expected-failure markers of round 1 are gone.
"""
import ctypes as C

import numpy as np
import pytest

import blocks
from mitsuba3_nasa_b200 import _lib
from orc import oracle_lib

pytestmark = pytest.mark.gpu

N = 20000


@pytest.mark.parametrize("op,name,seed", [(0, "fresnel", 30), (1, "fresnel_polarized", 31), (2, "specular_reflection", 32),
                                          (3, "specular_transmission", 34)])
def test_device_fresnel_blocks_equal_the_oracle(op, name, seed):
    want = blocks.BLOCKS[name][0](oracle_lib(), "orc_", N)
    ci = blocks._cos_theta(seed, N)
    for k, eta in enumerate(blocks._ETAS):
        got = {"o%d" % k: _lib.debug_fresnel(op, ci, eta)}
        ok, msg = blocks.compare(name, got, {"o%d" % k: want["o%d" % k]}, None)
        assert ok, msg


def test_device_refract_equals_the_oracle():
    want = blocks.b_refract(oracle_lib(), "orc_", N)
    rng = np.random.default_rng(33)
    wi, m = blocks._unit(rng, N), blocks._unit(rng, N)
    x = np.ascontiguousarray(np.concatenate([wi, m], axis=1))
    for k, eta in enumerate(blocks._ETAS):
        ok, msg = blocks.compare("refract", {"o": _lib.debug_fresnel(4, x, eta)}, {"o": want["o%d" % k]}, None)
        assert ok, msg


# ---- microfacet distribution (csrc/dmicrofacet.cuh, mib_debug_microfacet) against oracle/omicrofacet.h ----
@pytest.mark.parametrize("op,name,lo,hi,seed", [(2, "erf", -6, 6, 42), (3, "erfinv", -0.999999, 0.999999, 43), (4, "exp2", -130, 130, 44)])
def test_device_special_functions_equal_the_oracle(op, name, lo, hi, seed):
    x = np.random.default_rng(seed).uniform(lo, hi, N).astype(np.float32)
    want = blocks.BLOCKS[name][0](oracle_lib(), "orc_", N)["y"]
    ok, msg = blocks.compare(name, {"y": _lib.debug_microfacet(op, x).ravel()}, {"y": want}, None)
    assert ok, msg


def test_device_microfacet_eval_equals_the_oracle():
    want = blocks.b_microfacet_eval(oracle_lib(), "orc_", N)
    rng = np.random.default_rng(40)
    wi = blocks._upper(rng, N)
    m = blocks._unit(rng, N) * np.float32(0.3) + np.array([0, 0, 1], np.float32)
    m = (m / np.linalg.norm(m, axis=1, keepdims=True)).astype(np.float32)
    x = np.ascontiguousarray(np.concatenate([wi, m], axis=1))
    for k, (t, au, av) in enumerate(blocks._MICROFACETS):
        ok, msg = blocks.compare("microfacet_eval", {"o": _lib.debug_microfacet(0, x, t, au, av)}, {"o": want["o%d" % k]}, None)
        assert ok, msg


def test_device_microfacet_sample_equals_the_oracle():
    want = blocks.b_microfacet_sample(oracle_lib(), "orc_", N)
    rng = np.random.default_rng(41)
    wi = blocks._upper(rng, N)
    s = rng.uniform(0, 1, (N, 2)).astype(np.float32)
    x = np.ascontiguousarray(np.concatenate([wi, s], axis=1))
    for k, (t, au, av) in enumerate(blocks._MICROFACETS):
        o = _lib.debug_microfacet(1, x, t, au, av)
        got = {"m": np.ascontiguousarray(o[:, :3]), "p": np.ascontiguousarray(o[:, 3] / (1 + np.abs(o[:, 3])))}
        ok, msg = blocks.compare("microfacet_sample", got, {"m": want["m%d" % k], "p": want["p%d" % k]}, None)
        assert ok, msg

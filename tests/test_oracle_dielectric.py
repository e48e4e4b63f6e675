"""Oracle restatement of SmoothDielectric::sample (oracle/ofresnel.h <- src/bsdfs/dielectric.cpp:247-371), the ocean-interface
BSDF of SURVEY 8f rank 2.  The plugin itself cannot be compiled from the reference headers alone (it lives in a plugin .cpp on
top of libmitsuba), so this level is pinned by (a) its building blocks -- fresnel, specular_reflection / _transmission,
rotate_mueller_basis, stokes_basis are each bit-checked against the reference headers in tests/test_oracle_vs_ref.py -- and
(b) the properties the reference's own src/bsdfs/tests/test_dielectric.py checks: Snell's law and the mirror direction,
lobe probabilities r_i / 1 - r_i, eta and sampled_type per lobe, unit weights in unpolarized variants (times eta_ti^2 for
transmitted radiance), and agreement of the polarized intensity entry with the unpolarized weight."""
import ctypes as C

import numpy as np
import pytest

from orc import oracle_lib, fp

ETA = 1.33


def _sample(wi, s1, mask=3, eta=ETA):
    wi = np.ascontiguousarray(wi, np.float32)
    s1 = np.ascontiguousarray(s1, np.float32)
    out = np.empty((len(s1), 26), np.float32)
    oracle_lib().orc_dielectric_sample(fp(wi), C.c_float(eta), fp(s1), C.c_int(mask), fp(out), C.c_size_t(len(s1)))
    o = out.astype(np.float64)
    return dict(wo=o[:, 0:3], pdf=o[:, 3], eta=o[:, 4], type=o[:, 5].astype(int), w=o[:, 6], M=o[:, 7:23].reshape(-1, 4, 4))


def _fresnel(ci, eta=ETA):
    ci = np.ascontiguousarray(ci, np.float32)
    out = np.empty((len(ci), 4), np.float32)
    oracle_lib().orc_fresnel(fp(ci), C.c_float(eta), fp(out), C.c_size_t(len(ci)))
    return out.astype(np.float64)


def _dirs(n, seed):
    rng = np.random.default_rng(seed)
    v = rng.normal(size=(n, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    v = v[np.abs(v[:, 2]) > 0.02]          # stay off the exactly-grazing singularity
    return v.astype(np.float32)


def test_directions_lobes_and_unpolarized_weights():
    wi = _dirs(4000, 1)
    s1 = np.random.default_rng(2).uniform(0, 1, len(wi)).astype(np.float32)
    r = _sample(wi, s1)
    F = _fresnel(wi[:, 2])
    refl = s1 <= F[:, 0].astype(np.float32)
    assert refl.any() and (~refl).any()
    assert (r['type'][refl] == 0x20).all() and (r['type'][~refl] == 0x40).all()
    assert np.allclose(r['pdf'], np.where(refl, F[:, 0], 1 - F[:, 0]), atol=1e-7)
    assert np.allclose(r['eta'], np.where(refl, 1.0, F[:, 2]))
    # mirror direction / Snell's law in the local frame (normal = +z)
    assert np.allclose(r['wo'][refl], wi[refl] * [-1, -1, 1], atol=1e-7)
    t = ~refl
    assert np.allclose(np.linalg.norm(r['wo'][t], axis=1), 1, atol=2e-6)
    assert (np.sign(r['wo'][t, 2]) == -np.sign(wi[t, 2])).all()
    sin_i, sin_t = np.hypot(wi[t, 0], wi[t, 1]), np.hypot(r['wo'][t, 0], r['wo'][t, 1])
    n_i = np.where(wi[t, 2] > 0, 1.0, ETA)
    assert np.allclose(n_i * sin_i, (ETA + 1.0 - n_i) * sin_t, atol=3e-6)
    # unpolarized weight: 1 for reflection, eta_ti^2 for transmitted radiance
    assert np.allclose(r['w'], np.where(refl, 1.0, F[:, 3] ** 2), rtol=1e-6)
    # polarized weight: the intensity entry equals the unpolarized weight (R00 = r_i, T00 = 1 - r_i)
    assert np.allclose(r['M'][:, 0, 0], r['w'], rtol=1e-4)          # fp32: T00 / (1 - r_i) near grazing incidence
    # the blocks keep |M01| <= M00 (physical Mueller matrices), nothing leaks into NaN
    assert np.isfinite(r['M']).all() and (np.abs(r['M'][:, 0, 1]) <= r['M'][:, 0, 0] * (1 + 1e-6)).all()


def test_total_internal_reflection_always_reflects():
    wi = np.array([[np.sqrt(1 - 0.3 ** 2), 0.0, -0.3]] * 64, np.float32)       # from below the water, beyond the critical angle
    s1 = np.linspace(0, 1, 64).astype(np.float32)
    r = _sample(wi, s1)
    assert (r['type'] == 0x20).all() and np.allclose(r['pdf'], 1) and np.allclose(r['w'], 1)
    assert np.allclose(r['M'][:, 0, 0], 1, atol=1e-6)
    assert (np.abs(r['M'][:, 2, 3]) > 1e-3).all()          # TIR retards s against p: a genuine U <-> V coupling


def test_lobe_masks():
    """ctx.type_mask as volpathaos' refractive NEE sets it (transmission only): pdf 1, weight = (1 - r_i) eta_ti^2."""
    wi = _dirs(500, 3)
    wi = wi[wi[:, 2] > 0]                                   # from the air side: a transmitted direction always exists
    s1 = np.random.default_rng(4).uniform(0, 1, len(wi)).astype(np.float32)
    F = _fresnel(wi[:, 2])
    t = _sample(wi, s1, mask=2)
    assert (t['type'] == 0x40).all() and (t['pdf'] == 1).all()
    assert np.allclose(t['w'], (1 - F[:, 0]) * F[:, 3] ** 2, rtol=1e-6) and np.allclose(t['M'][:, 0, 0], t['w'], rtol=1e-4)
    rr = _sample(wi, s1, mask=1)
    assert (rr['type'] == 0x20).all() and (rr['pdf'] == 1).all() and np.allclose(rr['w'], F[:, 0], rtol=1e-6)
    none = _sample(wi, s1, mask=0)
    assert (none['pdf'] == 0).all() and (none['w'] == 0).all() and (none['M'] == 0).all()


def test_expected_weight_conserves_energy():
    """E over the lobe choice of the intensity weight = r_i + (1 - r_i) eta_ti^2; without the radiance scaling that is 1."""
    wi = np.array([[0.6, 0.0, 0.8]] * 20000, np.float32)
    s1 = ((np.arange(20000) + 0.5) / 20000).astype(np.float32)
    r = _sample(wi, s1)
    F = _fresnel(wi[:1, 2])[0]
    scale = np.where(r['type'] == 0x40, F[3] ** 2, 1.0)
    assert abs((r['M'][:, 0, 0] / scale).mean() - 1) < 1e-4
    assert abs(r['M'][:, 0, 0].mean() - (F[0] + (1 - F[0]) * F[3] ** 2)) < 1e-4


def test_normal_incidence_has_no_basis_rotation():
    """wi = n: the s axes are undefined (collinear), the reference falls back to (1, 0, 0) on both sides; reflection at
    normal incidence flips the handedness: diag(1, 1, -1, -1) times r_i / pdf = 1."""
    r = _sample(np.array([[0, 0, 1.0]] * 2, np.float32), np.array([0.0, 0.9], np.float32))
    assert r['type'][0] == 0x20 and r['type'][1] == 0x40
    assert np.allclose(np.abs(np.diag(r['M'][0])), 1, atol=1e-6) and np.allclose(r['M'][0] - np.diag(np.diag(r['M'][0])), 0, atol=1e-7)
    assert r['M'][0][2, 2] * r['M'][0][3, 3] > 0 and r['M'][0][1, 1] > 0
    t = r['M'][1] / r['M'][1][0, 0]
    assert np.allclose(np.abs(np.diag(t)), 1, atol=1e-6)

"""Oracle restatement of the fork's RoughDielectric (oracle/oroughdielectric.h <- src/bsdfs/roughdielectric.cpp:248-603), the
BSDF of the wind-roughened ocean surface in the fork's atmosphere-ocean tests (test_ad_volpathaos.py:234-338: Beckmann,
alpha = 0.1).  The plugin is a .cpp on top of libmitsuba and cannot be compiled from the headers alone; this level is
pinned by its building blocks (fresnel, specular_reflection / _transmission, rotate_mueller_basis, the microfacet
distribution: tests/test_oracle_vs_ref.py, against the reference headers) and by the relations between sample(), eval()
and pdf() that the reference's own BSDF tests rely on (chi^2 consistency of sample vs pdf, test_roughdielectric.py)."""
import ctypes as C

import numpy as np
import pytest

from orc import oracle_lib, fp

ETA = 1.33


def _sample(wi, s3, mask=3, type_=0, alpha=0.1, eta=ETA):
    wi, s3 = np.ascontiguousarray(wi, np.float32), np.ascontiguousarray(s3, np.float32)
    out = np.empty((len(wi), 26), np.float32)
    oracle_lib().orc_roughdielectric_sample(C.c_int(type_), C.c_float(alpha), C.c_float(eta), fp(wi), fp(s3), C.c_int(mask), fp(out),
                                            C.c_size_t(len(wi)))
    o = out.astype(np.float64)
    return dict(wo=o[:, 0:3], pdf=o[:, 3], eta=o[:, 4], type=o[:, 5].astype(int), w=o[:, 6], M=o[:, 7:23].reshape(-1, 4, 4))


def _eval(wi, wo, mask=3, type_=0, alpha=0.1, eta=ETA):
    wi, wo = np.ascontiguousarray(wi, np.float32), np.ascontiguousarray(wo, np.float32)
    out = np.empty((len(wi), 18), np.float32)
    oracle_lib().orc_roughdielectric_eval(C.c_int(type_), C.c_float(alpha), C.c_float(eta), fp(wi), fp(wo), C.c_int(mask), fp(out),
                                          C.c_size_t(len(wi)))
    o = out.astype(np.float64)
    return dict(v=o[:, 0], pdf=o[:, 1], M=o[:, 2:18].reshape(-1, 4, 4))


def _g1(type_, alpha, v, m):
    v, m = np.ascontiguousarray(v, np.float32), np.ascontiguousarray(m, np.float32)
    out = np.empty((len(v), 3), np.float32)
    oracle_lib().orc_microfacet_eval(C.c_int(type_), C.c_float(alpha), C.c_float(alpha), fp(v), fp(m), fp(out), C.c_size_t(len(v)))
    return out[:, 1].astype(np.float64)


def _inputs(n, seed, both_sides=True):
    rng = np.random.default_rng(seed)
    v = rng.normal(size=(n, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    v[:, 2] = np.abs(v[:, 2]) * 0.9 + 0.1                     # 6 .. 84 degrees from the normal
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    if both_sides:
        v[n // 2:, 2] *= -1                                   # the second half arrives from below the water
    return v.astype(np.float32), rng.uniform(0, 1, (n, 3)).astype(np.float32)


@pytest.mark.parametrize("type_,alpha", [(0, 0.1), (0, 0.4), (1, 0.2)])
def test_sample_eval_pdf_are_consistent(type_, alpha):
    wi, s3 = _inputs(6000, 1)
    s = _sample(wi, s3, type_=type_, alpha=alpha)
    ok = s['w'] > 0
    assert ok.mean() > 0.8
    refl = s['type'] == 0x8
    assert (refl & ok).any() and (~refl & ok).any()
    assert np.allclose(np.linalg.norm(s['wo'][ok], axis=1), 1, atol=3e-6)
    assert (np.sign(s['wo'][ok & refl, 2]) == np.sign(wi[ok & refl, 2])).all()
    assert (np.sign(s['wo'][ok & ~refl, 2]) == -np.sign(wi[ok & ~refl, 2])).all()
    assert np.allclose(s['eta'][ok], np.where(refl[ok], 1.0, np.where(wi[ok, 2] > 0, ETA, 1 / ETA)), rtol=1e-6)
    # pdf() of the sampled direction is the density sample() reports
    e = _eval(wi[ok], s['wo'][ok], type_=type_, alpha=alpha)
    assert np.allclose(e['pdf'], s['pdf'][ok], rtol=2e-3)
    # the fork's eval() carries G = 1 (roughdielectric.cpp:457) while sample() weights with smith_g1(wo, m): eval / pdf equals
    # 1 / smith_g1(wi, m) times the radiance scaling of the lobe (1 for reflection, eta_ti^2 for transmission)
    wo = s['wo'][ok]
    eta_s = np.where(wi[ok, 2] > 0, ETA, 1 / ETA)
    m = wi[ok] + wo * np.where(refl[ok], 1.0, eta_s)[:, None]
    m /= np.linalg.norm(m, axis=1, keepdims=True)
    m *= np.sign(m[:, 2])[:, None]
    up = wi[ok] * np.sign(wi[ok, 2])[:, None]
    g1_wi = _g1(type_, alpha, up, m)
    scale = np.where(refl[ok], 1.0, (1 / eta_s) ** 2)           # transmitted radiance: eta_ti^2
    good = (g1_wi > 1e-3) & (e['pdf'] > 1e-6)
    assert good.mean() > 0.9
    assert np.allclose((e['v'] * g1_wi / e['pdf'])[good], scale[good], rtol=5e-3)
    # and the unpolarized sampling weight is smith_g1(wo, m) times that scaling
    g1_wo = _g1(type_, alpha, wo * np.sign(wo[:, 2])[:, None], m)
    assert np.allclose(s['w'][ok][good], (g1_wo * scale)[good], rtol=2e-3, atol=1e-6)


def test_polarized_intensity_entries_follow_the_unpolarized_values():
    wi, s3 = _inputs(3000, 2)
    s = _sample(wi, s3)
    ok = s['w'] > 0
    # sample(): R00(dot(wo, m)) / F(dot(wi, m)) = 1 for reflection (same angle); for transmission the two Fresnel terms are
    # evaluated on opposite sides of the facet, which agree by reciprocity
    assert np.allclose(s['M'][ok, 0, 0], s['w'][ok], rtol=2e-3)
    e = _eval(wi[ok], s['wo'][ok])
    assert np.allclose(e['M'][:, 0, 0], e['v'], rtol=2e-3, atol=1e-7)
    assert np.isfinite(s['M']).all() and np.isfinite(e['M']).all()
    assert (np.abs(e['M'][:, 0, 1]) <= e['M'][:, 0, 0] * (1 + 1e-5) + 1e-9).all()


def test_lobe_masks_and_degenerate_inputs():
    wi, s3 = _inputs(1000, 3, both_sides=False)
    t = _sample(wi, s3, mask=2)
    ok = t['w'] > 0
    assert ok.mean() > 0.8 and (t['type'][ok] == 0x10).all() and (t['wo'][ok, 2] < 0).all()
    r = _sample(wi, s3, mask=1)
    ok = r['w'] > 0
    assert (r['type'][ok] == 0x8).all() and (r['wo'][ok, 2] > 0).all()
    none = _sample(wi, s3, mask=0)
    assert (none['pdf'] == 0).all() and (none['w'] == 0).all()
    # pdf() honours the mask: a reflected direction has no density when only transmission is enabled
    both = _sample(wi, s3)
    refl = (both['type'] == 0x8) & (both['w'] > 0)
    e = _eval(wi[refl], both['wo'][refl], mask=2)
    assert (e['pdf'] == 0).all() and (e['v'] == 0).all()
    # grazing wi (cos = 0): inactive
    g = _sample(np.array([[1, 0, 0]], np.float32), np.array([[0.3, 0.4, 0.5]], np.float32))
    assert g['w'][0] == 0 and (g['M'][0] == 0).all()


def test_energy_of_a_smooth_limit():
    """alpha -> 1e-4 (the distribution's floor): the rough BSDF samples the mirror / Snell directions of the smooth one."""
    wi = np.tile(np.array([[0.6, 0.0, 0.8]], np.float32), (2000, 1))
    s3 = np.random.default_rng(4).uniform(0, 1, (2000, 3)).astype(np.float32)
    s = _sample(wi, s3, alpha=1e-4)
    refl = s['type'] == 0x8
    assert np.allclose(s['wo'][refl], [-0.6, 0.0, 0.8], atol=2e-3)
    sin_t = 0.6 / ETA
    assert np.allclose(s['wo'][~refl], [-sin_t, 0.0, -np.sqrt(1 - sin_t ** 2)], atol=2e-3)
    F = ((0.8 - ETA * np.sqrt(1 - sin_t ** 2)) / (0.8 + ETA * np.sqrt(1 - sin_t ** 2))) ** 2 * 0.5 + \
        ((ETA * 0.8 - np.sqrt(1 - sin_t ** 2)) / (ETA * 0.8 + np.sqrt(1 - sin_t ** 2))) ** 2 * 0.5
    assert abs(refl.mean() - F) < 0.02

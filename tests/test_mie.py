"""Lorenz-Mie tabulation: oracle/omie.h against an independent float64 implementation, the
size distributions of mitsuba3_nasa_b200/mie.py, and (gpu) the CUDA tabulation against the oracle.

No reference test pins mie / sizedistr, so the pin is physics: Bohren & Huffman's Riccati-Bessel
form evaluated in float64 with scipy, and normalisation identities."""
import ctypes as C

import numpy as np
import pytest
from scipy import special as sp

from mitsuba3_nasa_b200 import mie as M
from orc import oracle_lib, fp


def mie64(wavelength, mu, radius, n_med, n_sph):
    m = n_sph / n_med
    x = 2 * np.pi * n_med * radius / wavelength
    nmax = int(x + 4.05 * x ** (1 / 3) + 8)
    n = np.arange(1, nmax + 1)
    jn = lambda z, d=False: sp.spherical_jn(n, z, derivative=d)
    yn = lambda z, d=False: sp.spherical_yn(n, z, derivative=d)
    psi = lambda z: z * jn(z)
    dpsi = lambda z: jn(z) + z * jn(z, True)
    xi = lambda z: z * (jn(z) + 1j * yn(z))
    dxi = lambda z: (jn(z) + 1j * yn(z)) + z * (jn(z, True) + 1j * yn(z, True))
    mx = m * x
    a = (m * psi(mx) * dpsi(x) - psi(x) * dpsi(mx)) / (m * psi(mx) * dxi(x) - xi(x) * dpsi(mx))
    b = (psi(mx) * dpsi(x) - m * psi(x) * dpsi(mx)) / (psi(mx) * dxi(x) - m * xi(x) * dpsi(mx))
    mu = np.atleast_1d(mu).astype(np.float64)
    pi = np.zeros((nmax + 1, mu.size))
    tau = np.zeros((nmax + 1, mu.size))
    pi[1] = 1
    tau[1] = mu
    for k in range(2, nmax + 1):
        pi[k] = (2 * k - 1) / (k - 1) * mu * pi[k - 1] - k / (k - 1) * pi[k - 2]
        tau[k] = k * mu * pi[k] - (k + 1) * pi[k - 1]
    kn = (2 * n + 1) / (n * (n + 1))
    S1 = ((kn * a)[:, None] * tau[1:] + (kn * b)[:, None] * pi[1:]).sum(0)      # the reference's S1 (mie.h:190)
    S2 = ((kn * a)[:, None] * pi[1:] + (kn * b)[:, None] * tau[1:]).sum(0)
    k = 2 * np.pi * n_med / wavelength
    Cs = 2 * np.pi / k ** 2 * ((2 * n + 1) * (abs(a) ** 2 + abs(b) ** 2)).sum()
    norm = k ** 2 * Cs
    return (0.5 * (abs(S1) ** 2 + abs(S2) ** 2) / norm, 0.5 * (abs(S1) ** 2 - abs(S2) ** 2) / norm,
            (S1 * np.conj(S2)).real / norm, (S1 * np.conj(S2)).imag / norm, Cs)


def oracle_tabulate(cost, wavelength, distr, ior_med=(1.0, 0.0), ior_sph=(1.33, 0.0), nmax=-1):
    cost = np.ascontiguousarray(cost, np.float32)
    out = [np.zeros(cost.size, np.float32) for _ in range(4)]
    if distr.is_monodisperse:
        radius, w, sdf, g = np.array([distr.min_radius], np.float32), None, None, 1
    else:
        radius, w, sdf = [np.ascontiguousarray(a, np.float32) for a in distr.eval_gauss_all()]
        g = radius.size
    p = lambda a: fp(a) if a is not None else None
    oracle_lib().orc_mie_tabulate(C.c_float(wavelength), C.c_float(ior_med[0]), C.c_float(ior_med[1]), C.c_float(ior_sph[0]),
                                  C.c_float(ior_sph[1]), C.c_int(nmax), p(radius), p(w), p(sdf), C.c_int(g), fp(cost),
                                  C.c_int(cost.size), *[fp(o) for o in out])
    return out


@pytest.mark.parametrize("radius", [0.05, 0.5, 3.0, 12.0])
def test_oracle_mie_matches_float64_series(radius):
    mu = np.linspace(-1, 1, 61).astype(np.float32)
    got = oracle_tabulate(mu, 0.55, M.MonodisperseSizeDistr({'radius': radius}))
    ref = mie64(0.55, mu, radius, 1.0, 1.33)
    scale = np.abs(ref[0]).max()
    tol = 2e-4 if radius > 5 else 3e-5          # float32 recurrences over ~nmax terms
    for q in range(4):
        assert np.abs(got[q] - ref[q]).max() / scale < tol, (q, np.abs(got[q] - ref[q]).max() / scale)
    assert np.abs(got[0] / ref[0] - 1).max() < 50 * tol


def test_oracle_mie_absorbing_sphere():
    mu = np.linspace(-1, 1, 41).astype(np.float32)
    got = oracle_tabulate(mu, 0.55, M.MonodisperseSizeDistr({'radius': 1.5}), ior_sph=(1.5, 0.05))
    ref = mie64(0.55, mu, 1.5, 1.0, 1.5 + 0.05j)
    assert np.abs(got[0] - ref[0]).max() / np.abs(ref[0]).max() < 1e-4


def test_phase_function_is_normalised():
    """mie.h:38-40: (|S1|^2 + |S2|^2) / (2 Ns) integrates to one over the sphere."""
    nodes, weights = np.polynomial.legendre.leggauss(400)
    m11 = oracle_tabulate(nodes.astype(np.float32), 0.55, M.MonodisperseSizeDistr({'radius': 2.0}))[0]
    assert abs(2 * np.pi * np.sum(weights * m11.astype(np.float64)) - 1) < 2e-4


def test_rayleigh_limit():
    """radius << wavelength: m11 -> 3/(16 pi) (1 + mu^2), m12/m11 -> -(1 - mu^2)/(1 + mu^2)."""
    mu = np.linspace(-0.95, 0.95, 21).astype(np.float32)
    m11, m12, m33, m34 = oracle_tabulate(mu, 0.55, M.MonodisperseSizeDistr({'radius': 0.002}))
    mu = mu.astype(np.float64)
    assert np.allclose(m11, 3 / (16 * np.pi) * (1 + mu ** 2), rtol=2e-3)
    assert np.allclose(m12 / m11, -(1 - mu ** 2) / (1 + mu ** 2), atol=2e-3)
    assert np.allclose(m33 / m11, 2 * mu / (1 + mu ** 2), atol=2e-3)


@pytest.mark.parametrize("cls,props", [
    (M.LogNormalSizeDistr, {'mean_radius': 10.0, 'std': 1.4, 'min_radius': 1.0, 'max_radius': 40.0, 'gauss_points': 100}),
    (M.GammaSizeDistr, {'a': 6.0, 'b': 0.25, 'min_radius': 0.5, 'max_radius': 40.0, 'gauss_points': 80}),
    (M.PowerLawSizeDistr, {'exponent': 3.0, 'min_radius': 0.5, 'max_radius': 5.0, 'gauss_points': 50})])
def test_size_distributions_are_normalised(cls, props):
    """sizedistr.cpp:27-45: the density integrates to one over [min_radius, max_radius] under its own quadrature."""
    d = cls(props)
    r, w, sdf = d.eval_gauss_all()
    assert r.dtype == np.float32 and r.size == props['gauss_points']
    assert r.min() > props['min_radius'] and r.max() < props['max_radius']
    assert abs(float(np.sum(w.astype(np.float64) * sdf)) - 1) < 1e-5
    assert abs(float(np.sum(w.astype(np.float64))) - (props['max_radius'] - props['min_radius'])) < 1e-3


def test_polydisperse_average_against_float64():
    d = M.LogNormalSizeDistr({'mean_radius': 1.0, 'std': 1.3, 'min_radius': 0.3, 'max_radius': 3.0, 'gauss_points': 24})
    mu = np.linspace(-1, 1, 31).astype(np.float32)
    got = oracle_tabulate(mu, 0.55, d)
    r, w, sdf = d.eval_gauss_all()
    num, den = np.zeros(mu.size), 0.0
    for ri, wi, si in zip(r, w, sdf):
        p11, _, _, _, cs = mie64(0.55, mu, float(ri), 1.0, 1.33)
        num += float(wi) * float(si) * cs * p11
        den += float(wi) * float(si) * cs
    ref = num / den
    assert np.abs(got[0] / ref - 1).max() < 5e-4


# ------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("case", ["mono_small", "mono_large", "lognormal", "absorbing"])
def test_device_tabulation_matches_oracle(case):
    cost = np.linspace(-1, 1, 257).astype(np.float32)
    kw = {}
    if case == "mono_small":
        d = M.MonodisperseSizeDistr({'radius': 0.3})
    elif case == "mono_large":
        d = M.MonodisperseSizeDistr({'radius': 20.0})
    elif case == "absorbing":
        d = M.MonodisperseSizeDistr({'radius': 2.0})
        kw = dict(ior_med=(1.0, 0.0), ior_sph=(1.45, 0.02))
    else:
        d = M.LogNormalSizeDistr({'mean_radius': 10.0, 'std': 1.4, 'min_radius': 1.0, 'max_radius': 40.0, 'gauss_points': 32})
    got = M.mie_tabulate(cost, 0.55, d, **kw)
    ref = oracle_tabulate(cost, 0.55, d, **kw)
    scale = np.abs(ref[0]).max()
    for q in range(4):
        err = np.abs(got[q].astype(np.float64) - ref[q]).max() / scale
        print(case, "entry", q, "max err / max m11 =", err, "bit-identical" if np.array_equal(got[q], ref[q]) else "")
        assert err < 1e-5


@pytest.mark.gpu
def test_mie_plugin_in_a_scene_matches_tabphase_polarized():
    """{'type': 'mie'} is tabulated at load time and behaves like tabphase_polarized with those tables."""
    from mitsuba3_nasa_b200 import scene as S, scenes, _lib
    from orc import OracleScene
    grid = np.random.default_rng(2).uniform(0.1, 1.0, (8, 8, 8, 1)).astype(np.float32)
    phase = {'type': 'mie', 'ior_sph': 1.33, 'wavelength': 0.55, 'n_cost': 181,
             'sizedistr': {'type': 'lognormal', 'mean_radius': 2.0, 'std': 1.3, 'min_radius': 0.5, 'max_radius': 6.0,
                           'gauss_points': 16}}
    d = scenes.c3_cloud(res=8, spp=8, max_depth=8, integrator='volpathaos', grid=grid, optical_depth=4.0, phase=phase)
    d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
    sc = S.load_dict(d, "llvm_ad_mono_polarized")
    ph = sc.desc.phases[sc.desc.media[0].phase]
    assert ph.type == 4 and ph.n_nodes == 181 and sc.buffers[ph.buf_m11].size == 181      # tabphase_polarized tables
    assert (sc.buffers[ph.buf_m11] > 0).all()
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    g = dev.sample_lanes(0, 8, 0, 512).astype(np.float64)
    o = orc.sample_lanes(0, 8, 0, 512).astype(np.float64)
    scale = np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)
    assert ((np.abs(g - o) / scale > 1e-4).any(axis=1)).mean() <= 2e-3
    with pytest.raises(S.PluginError, match="monochromatic"):
        S.load_dict(d, "llvm_ad_rgb")

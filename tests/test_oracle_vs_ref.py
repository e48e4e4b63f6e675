"""Pins the CPU oracle (oracle/liboracle.so) before anything is compared with it:

 1. against tests/golden/ref_blocks.npz -- outputs of the REFERENCE'S OWN headers for every
    building block of the path (made by tests/golden/make_golden.py in the build container);
 2. against the same reference headers live (oracle/_ref/libmiref.so) when that library is
    present (it travels to the GPU box as a prebuilt file; it cannot be rebuilt there);
 3. against the known-answer values the reference's own test-suite holds for this path
    (TEA: src/core/tests/test_random.py:8-27; PCG32 default stream:
    src/samplers/tests/test_independent.py:16-28; trilinear / nearest look-ups:
    src/volumes/tests/test_grid.py:23-80; Mueller identities: src/render/tests/test_mueller.py:8-20).
"""
import ctypes as C
import os

import numpy as np
import pytest

import blocks
from orc import oracle_lib, ref_lib, fp

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_blocks.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


@pytest.mark.parametrize("name", sorted(blocks.BLOCKS))
def test_oracle_matches_golden_reference_outputs(name, golden):
    fn, tol = blocks.BLOCKS[name]
    n = int(golden["n"][0])
    got = fn(oracle_lib(), "orc_", n)
    want = {k.split("/", 1)[1]: golden[k] for k in golden.files if k.startswith(name + "/")}
    assert want, "golden file has no entry for " + name
    ok, msg = blocks.compare(name, got, want, tol)
    assert ok, msg


@pytest.mark.parametrize("name", sorted(blocks.BLOCKS))
def test_oracle_matches_live_reference_headers(name):
    ref = ref_lib()
    if ref is None:
        pytest.skip("oracle/_ref/libmiref.so not built (needs /root/reference)")
    fn, tol = blocks.BLOCKS[name]
    n = 20000 if name not in ("texture3", "irregular") else 4000
    ok, msg = blocks.compare(name, fn(oracle_lib(), "orc_", n), fn(ref, "ref_", n), tol)
    assert ok, msg


@pytest.mark.parametrize("name", ["fresnel", "fresnel_polarized", "specular_reflection", "specular_transmission", "microfacet_eval"])
def test_fresnel_blocks_bit_exact_in_scalar_forms(name, golden):
    """The oracle follows the JIT variants (fma-chain squared_norm, rsqrt = sqrt(1/x)); switched to the scalar forms of
    the same two operations it reproduces the reference's scalar headers bit for bit (golden, and live when present)."""
    fn, _ = blocks.BLOCKS[name]
    n = int(golden["n"][0])
    lib = oracle_lib()
    lib.orc_fresnel_scalar_forms(1)
    try:
        got = fn(lib, "orc_", n)
        want = {k.split("/", 1)[1]: golden[k] for k in golden.files if k.startswith(name + "/")}
        assert want
        ok, msg = blocks.compare(name, got, want, None)
        assert ok, msg
        ref = ref_lib()
        if ref is not None:
            ok, msg = blocks.compare(name, fn(lib, "orc_", 20000), fn(ref, "ref_", 20000), None)
            assert ok, msg
    finally:
        lib.orc_fresnel_scalar_forms(0)


def test_fresnel_physics():
    """Energy conservation and the textbook limits (mueller.h:198-268; test_mueller.py / test_fresnel.py hold the same
    identities): R + T = 1 in the intensity entry, normal incidence R = ((eta-1)/(eta+1))^2, Brewster angle kills r_p,
    total internal reflection gives R = 1, T = 0 and a pure retarder (|M22|^2 + |M23|^2 = 1)."""
    lib = oracle_lib()
    eta = 1.33
    ci = np.linspace(-1, 1, 2001).astype(np.float32)
    ci = ci[np.abs(ci) > 1e-3]
    n = len(ci)
    R, T, F = np.empty((n, 16), np.float32), np.empty((n, 16), np.float32), np.empty((n, 4), np.float32)
    lib.orc_specular_reflection(fp(ci), C.c_float(eta), fp(R), C.c_size_t(n))
    lib.orc_specular_transmission(fp(ci), C.c_float(eta), fp(T), C.c_size_t(n))
    lib.orc_fresnel(fp(ci), C.c_float(eta), fp(F), C.c_size_t(n))
    assert np.abs(R[:, 0] + T[:, 0] - 1).max() < 2e-6
    assert np.abs(R[:, 0] - F[:, 0]).max() < 2e-6          # unpolarized fresnel() equals the intensity entry
    one = np.array([1.0], np.float32)
    r1 = np.empty((1, 16), np.float32)
    lib.orc_specular_reflection(fp(one), C.c_float(eta), fp(r1), C.c_size_t(1))
    assert abs(r1[0, 0] - ((eta - 1) / (eta + 1)) ** 2) < 1e-7
    brewster = np.array([np.cos(np.arctan(eta))], np.float32)
    lib.orc_specular_reflection(fp(brewster), C.c_float(eta), fp(r1), C.c_size_t(1))
    assert abs(r1[0, 0] - r1[0, 1]) < 1e-7                  # r_p = 0: a = b
    tir = (ci < 0) & (1 - ci.astype(np.float64) ** 2 > (1 / eta) ** 2 + 1e-3)
    assert tir.any()
    assert np.abs(R[tir, 0] - 1).max() < 1e-6 and np.abs(T[tir, 0]).max() == 0
    assert np.abs(R[tir, 10] ** 2 + R[tir, 11] ** 2 - 1).max() < 1e-5
    assert (np.abs(R[tir, 11]) > 0).any()                    # a real phase shift between s and p


def _fres(ci, eta):
    a, o = np.array([ci], np.float32), np.empty((1, 4), np.float32)
    oracle_lib().orc_fresnel(fp(a), C.c_float(eta), fp(o), C.c_size_t(1))
    return o[0].astype(np.float64)


def _fres_pol(ci, eta):
    a, o = np.array([ci], np.float32), np.empty((1, 7), np.float32)
    oracle_lib().orc_fresnel_polarized(fp(a), C.c_float(eta), fp(o), C.c_size_t(1))
    o = o[0].astype(np.float64)
    return complex(o[0], o[1]), complex(o[2], o[3]), o[4]


def test_fresnel_known_answers():
    """src/render/tests/test_fresnel.py:6-37 (test01_fresnel): normal / grazing incidence and the two hyperphysics spot checks."""
    close = lambda a, b: np.allclose(a, b, rtol=1e-5, atol=1e-6)
    assert close(_fres(1, 1.5), (0.04, -1, 1.5, 1 / 1.5)) and close(_fres(-1, 1.5), (0.04, 1, 1 / 1.5, 1.5))
    assert close(_fres(1, 1 / 1.5), (0.04, -1, 1 / 1.5, 1.5)) and close(_fres(-1, 1 / 1.5), (0.04, 1, 1.5, 1 / 1.5))
    assert close(_fres(0, 1.5), (1, -np.sqrt(1 - 1 / 1.5 ** 2), 1.5, 1 / 1.5)) and close(_fres(0, 1 / 1.5), (1, 0, 1 / 1.5, 1.5))
    c45 = np.cos(np.radians(45))
    F, ct, _, scale = _fres(c45, 1.5)
    assert close((scale * np.sqrt(1 - c45 ** 2)) ** 2 + ct ** 2, 1) and close(ct, -np.cos(np.radians(28.1255057020557)))
    assert close(F, 0.5 * (0.09201336304552442 ** 2 + 0.3033370452904235 ** 2))
    F, ct, _, _ = _fres(c45, 1 / 1.5)
    assert close(F, 1) and close(ct, 0)
    c10 = np.cos(np.radians(10))
    F, ct, _, scale = _fres(c10, 1 / 1.5)
    assert close(ct, -np.cos(np.radians(15.098086605159006))) and close(F, 0.5 * (0.19046797197779405 ** 2 + 0.20949431963852014 ** 2))
    # test02_fresnel_polarized_vec :49-53: eta = 1 reflects nothing and leaves the direction alone
    for ci in np.linspace(-1, 1, 20):
        F, ct, _, _ = _fres(ci, 1.0)
        assert F == 0 and abs(ct + ci) < 5e-7


def test_fresnel_polarized_known_answers():
    """src/render/tests/test_fresnel.py:39-46 (Brewster), :78-140 (Hecht Fig. 4.49 / 4.50: signs and relative phases of the
    amplitudes for external and internal reflection), :142-170 (Azzam 2004: phases under total internal reflection)."""
    a_s, a_p, _ = _fres_pol(np.cos(np.radians(56.3099)), 1.5)
    assert np.isclose((a_s * a_s).real, 0.3846150939598947 ** 2, rtol=1e-5) and np.isclose((a_p * a_p).real, 0, atol=1e-6)
    dphi = lambda p, s: abs(np.angle(p) - np.angle(s))
    for eta, sign in ((1.5, -1), (1 / 1.5, +1)):
        brewster = np.arctan(eta)
        a_s, a_p, _ = _fres_pol(1, eta)                                  # normal incidence: opposite signs
        assert sign * a_s.real > 0 and sign * a_p.real < 0 and a_s.imag == 0 and a_p.imag == 0 and np.isclose(dphi(a_p, a_s), np.pi, atol=1e-5)
        a_s, a_p, _ = _fres_pol(np.cos(brewster - 0.01), eta)
        assert sign * a_s.real > 0 and sign * a_p.real < 0 and np.isclose(dphi(a_p, a_s), np.pi, atol=1e-5)
        a_s, a_p, _ = _fres_pol(np.cos(brewster), eta)                   # Brewster: no p-polarized reflection
        assert abs(a_p.real) < 1e-5 and sign * a_s.real > 0 and a_s.imag == 0 and a_p.imag == 0
        a_s, a_p, _ = _fres_pol(np.cos(brewster + 0.01), eta)            # beyond: same sign, no relative phase
        assert sign * a_s.real > 0 and sign * a_p.real > 0 and np.isclose(dphi(a_p, a_s), 0, atol=1e-5)
    a_s, a_p, _ = _fres_pol(0, 1 / 1.5)                                  # grazing: both amplitudes -> -1
    assert np.isclose(a_s.real, -1, atol=1e-3) and np.isclose(a_p.real, -1, atol=1e-3) and a_s.imag == 0 and a_p.imag == 0
    eta = 1 / 1.5
    a_s, a_p, _ = _fres_pol(np.cos(np.arcsin(eta)), eta)                 # critical angle: both -> +1, phases 0
    assert np.isclose(a_s.real, 1, atol=1e-3) and np.isclose(a_p.real, 1, atol=1e-3)
    assert np.isclose(np.angle(a_s), 0, atol=2e-3) and np.isclose(np.angle(a_p), 0, atol=2e-3)
    a_s, a_p, _ = _fres_pol(0, eta)
    assert np.isclose(abs(np.angle(a_s)), np.pi, atol=1e-5) and np.isclose(abs(np.angle(a_p)), np.pi, atol=1e-5)
    cos_max = np.sqrt((1 - eta ** 2) / (1 + eta ** 2))                   # Azzam Eq. (11): the largest phase difference
    a_s, a_p, ct = _fres_pol(cos_max, eta)
    assert ct == 0 and np.isclose(np.angle(a_p) - np.angle(a_s), 2 * np.arctan((1 - eta ** 2) / (2 * eta)), rtol=1e-5)


def test_microfacet_sampling_is_consistent_with_its_density():
    """microfacet.h:219-226 vs :311-324: the density returned by sample() is pdf(wi, m) of the sampled normal, the sampled
    normals are unit vectors on the upper hemisphere, and D(m) cos(theta_m) integrates to one (Beckmann and GGX)."""
    lib = oracle_lib()
    n = 4000
    rng = np.random.default_rng(50)
    wi = blocks._upper(rng, n)
    s = rng.uniform(0, 1, (n, 2)).astype(np.float32)
    for t, au, av in blocks._MICROFACETS:
        o = np.zeros((n, 4), np.float32)
        lib.orc_microfacet_sample(C.c_int(t), C.c_float(au), C.c_float(av), fp(wi), fp(s), fp(o), C.c_size_t(n))
        m = np.ascontiguousarray(o[:, :3])
        assert np.allclose(np.linalg.norm(m, axis=1), 1, atol=1e-6) and (m[:, 2] > 0).all()
        e = np.zeros((n, 3), np.float32)
        lib.orc_microfacet_eval(C.c_int(t), C.c_float(au), C.c_float(av), fp(wi), fp(m), fp(e), C.c_size_t(n))
        assert np.allclose(e[:, 2], o[:, 3], rtol=1e-5, atol=1e-7)
        # quadrature of D(m) cos(theta) over the hemisphere in slope space-friendly coordinates
        th = (np.arange(4000) + 0.5) / 4000 * (np.pi / 2)
        ph = (np.arange(64) + 0.5) / 64 * 2 * np.pi
        T, P = np.meshgrid(th, ph, indexing="ij")
        mm = np.stack([np.sin(T) * np.cos(P), np.sin(T) * np.sin(P), np.cos(T)], -1).reshape(-1, 3).astype(np.float32)
        ee = np.zeros((len(mm), 3), np.float32)
        up = np.tile(np.array([[0, 0, 1]], np.float32), (len(mm), 1))
        lib.orc_microfacet_eval(C.c_int(t), C.c_float(au), C.c_float(av), fp(up), fp(np.ascontiguousarray(mm)), fp(ee), C.c_size_t(len(mm)))
        integral = (ee[:, 0].astype(np.float64) * np.cos(T).ravel() * np.sin(T).ravel()).sum() * (np.pi / 2 / 4000) * (2 * np.pi / 64)
        assert abs(integral - 1) < 2e-3, (t, au, av, integral)


# ---------------------------------------------------------------------------
# Known answers from the reference's own tests
# ---------------------------------------------------------------------------
def _tea(v0, v1):
    a0, a1 = np.array([v0], np.uint32), np.array([v1], np.uint32)
    o0, o1 = np.empty(1, np.uint32), np.empty(1, np.uint32)
    oracle_lib().orc_tea(fp(a0), fp(a1), fp(o0), fp(o1), C.c_size_t(1))
    return int(o0[0]), int(o1[0])


@pytest.mark.parametrize("v0,v1,f32,f64", [
    (1, 1, 0.5424730777740479, 0.5424730799533735), (1, 2, 0.5079904794692993, 0.5079905082233922),
    (1, 3, 0.4171961545944214, 0.4171962610608142), (1, 4, 0.008385419845581055, 0.008385529523330604),
    (1, 5, 0.8085528612136841, 0.80855288317879), (2, 1, 0.6939879655838013, 0.6939880404156831),
    (3, 1, 0.6978365182876587, 0.6978365636630994), (4, 1, 0.4897364377975464, 0.48973647949223253)])
def test_tea_known_answers(v0, v1, f32, f64):
    """test_random.py:8-27 through sample_tea_float32/64 (include/mitsuba/core/random.h:135-160)."""
    o0, o1 = _tea(v0, v1)
    got32 = np.array([(o1 >> 9) | 0x3f800000], np.uint32).view(np.float32)[0] - np.float32(1)
    assert float(got32) == f32
    u64 = (o0 + (o1 << 32)) & 0xffffffffffffffff
    got64 = np.array([(u64 >> 12) | 0x3ff0000000000000], np.uint64).view(np.float64)[0] - 1.0
    assert float(got64) == f64


def test_pcg32_default_stream_known_answers():
    """PCG32(initstate=0x853c49e6748fea9b, initseq=0xda3e39cb94b95bdb) is the generator's published
    demo stream; its first six outputs are the well-known pcg32-global values."""
    out = np.empty(6, np.uint32)
    oracle_lib().orc_pcg32_uints(C.c_uint64(0x853c49e6748fea9b), C.c_uint64(0xda3e39cb94b95bdb), fp(out), C.c_size_t(6))
    # seed(1, state, seq) of drjit/random.h:60-66 == pcg32_srandom_r(state, seq)
    state, inc = 0, ((0xda3e39cb94b95bdb << 1) | 1) & (2 ** 64 - 1)

    def step():
        nonlocal state
        old = state
        state = (old * 0x5851f42d4c957f2d + inc) & (2 ** 64 - 1)
        xs = (((old >> 18) ^ old) >> 27) & 0xffffffff
        rot = old >> 59
        return ((xs >> rot) | (xs << ((-rot) & 31))) & 0xffffffff
    step()
    state = (state + 0x853c49e6748fea9b) & (2 ** 64 - 1)
    step()
    assert [int(x) for x in out] == [step() for _ in range(6)]


def _tex(data, res, ch, filt, p):
    data = np.ascontiguousarray(data, np.float32).reshape(-1)
    p = np.ascontiguousarray(p, np.float32).reshape(-1, 3)
    out = np.zeros((p.shape[0], ch), np.float32)
    oracle_lib().orc_texture3_eval(fp(data), C.c_int(res[0]), C.c_int(res[1]), C.c_int(res[2]), C.c_int(ch), C.c_int(filt),
                                   C.c_int(2), fp(p), fp(out), C.c_size_t(p.shape[0]))
    return out


def test_grid_lookup_known_answers():
    """src/volumes/tests/test_grid.py:23-80 (grids are [z][y][x][c]; res passed as x,y,z)."""
    g = np.ones((3, 3, 3), np.float32)
    g[0, 0, 0] = 0.0
    g[1, 1, 1] = 0.5
    assert np.allclose(_tex(g, (3, 3, 3), 1, 0, [[0, 0, 0]]), 0.0)                       # test02 nearest
    assert np.allclose(_tex(g, (3, 3, 3), 1, 0, [[1 / 6, 1 / 6, 1 / 6]]), 0.0)
    assert np.allclose(_tex(g, (3, 3, 3), 1, 1, [[0, 0, 0]]), 0.0)                       # test03 trilinear
    assert np.allclose(_tex(g, (3, 3, 3), 1, 1, [[1 / 3, 1 / 3, 1 / 3]]), (6 * 1 + 0.5 + 0.0) / 8)
    g = np.ones((3, 8, 5), np.float32)                                                    # test04 boundary
    g[0, 0, 0] = 0.0
    g[1, 1, 1] = 0.5
    g[:, 7, :] = 0.55
    g[0, :, :] = 0.32
    assert np.allclose(_tex(g, (5, 8, 3), 1, 1, [[0.2, 1.0, 0.7]]), 0.55)
    assert np.allclose(_tex(g, (5, 8, 3), 1, 1, [[0.5, 0.5, 0.0]]), 0.32)


def test_mueller_known_answers():
    """test_mueller.py:8-20: depolarizer and rotator acting on Stokes vectors."""
    th = np.array([45 * np.pi / 180, -45 * np.pi / 180], np.float32)
    m = np.empty((2, 16), np.float32)
    oracle_lib().orc_rotator(fp(th), fp(m), C.c_size_t(2))
    s = np.array([1, 1, 0, 0], np.float32)
    assert np.allclose(m[0].reshape(4, 4) @ s, [1, 0, -1, 0], atol=1e-5)
    assert np.allclose(m[1].reshape(4, 4) @ s, [1, 0, +1, 0], atol=1e-5)


def test_rayleigh_matrix_normalisation():
    """rayleigh.cpp:41-128: the (0,0) entry integrates to one over the sphere for any depolarization."""
    mu = np.linspace(-1, 1, 20001).astype(np.float32)
    st = np.sqrt(np.maximum(0, 1 - mu.astype(np.float64) ** 2)).astype(np.float32)
    for rho in (0.0, 0.1):
        m = np.empty((mu.size, 16), np.float32)
        oracle_lib().orc_rayleigh_scatter(fp(mu), fp(st), C.c_float(rho), fp(m), C.c_size_t(mu.size))
        integral = 2 * np.pi * np.trapezoid(m[:, 0].astype(np.float64), mu.astype(np.float64))
        assert abs(integral - 1) < 2e-3 if rho else abs(integral - 1) < 1e-5


def _mie_errors(got, want):
    """Errors of (S1, S2) relative to max(|S1|, |S2|) of the row, of (Ns, Cs, Ct) relative to themselves."""
    out = {}
    for k in want:
        g, w = got[k].astype(np.float64), want[k].astype(np.float64)
        mag = np.maximum(np.hypot(w[:, 0], w[:, 1]), np.hypot(w[:, 2], w[:, 3]))[:, None]
        out[k] = (np.abs(g[:, :4] - w[:, :4]) / mag).max(), (np.abs(g[:, 4:] - w[:, 4:]) / np.abs(w[:, 4:])).max()
    return out


def test_mie_header(golden):
    """oracle/omie.h against the reference's header-only Lorenz-Mie series (include/mitsuba/render/mie.h:45-222, scalar float
    instantiation): size parameters 0.1 ... 450, absorbing and non-absorbing spheres, automatic and fixed series length.
    The scalar header multiplies complex numbers without fma and the oracle with the fma forms of the JIT variants, so the
    amplitudes agree to 1e-4 of their magnitude (70 % of all outputs are bit-equal) and the cross sections to 1e-6."""
    want = {k.split("/", 1)[1]: golden[k] for k in golden.files if k.startswith("mie/")}
    assert want, "golden file has no entry for mie"
    for k, (amp, cross) in _mie_errors(blocks.b_mie(oracle_lib(), "orc_", 600), want).items():
        assert amp <= 1e-4 and cross <= 2e-6, (k, amp, cross)
    ref = ref_lib()
    if ref is not None:
        for k, (amp, cross) in _mie_errors(blocks.b_mie(oracle_lib(), "orc_", 600), blocks.b_mie(ref, "ref_", 600)).items():
            assert amp <= 1e-4 and cross <= 2e-6, (k, amp, cross)


def test_gauss_legendre_header(golden):
    """The size distributions' quadrature (mie.py SizeDistribution: numpy leggauss cast to float32) against
    quad::gauss_legendre<DynamicArray<float>> (include/mitsuba/core/quad.h:27-90, Newton iteration in double)."""
    from mitsuba3_nasa_b200 import mie
    for n in (1, 2, 3, 16, 100):
        sd = mie.SizeDistribution({'min_radius': -1.0, 'max_radius': 1.0, 'gauss_points': n})
        x, w = golden["gauss_legendre/x%d" % n], golden["gauss_legendre/w%d" % n]
        assert np.abs(sd.nodes - x).max() <= 1e-7 and np.abs(sd.weights - w).max() <= 1e-7, n


def test_plastic_tables_header(golden):
    """The tables roughplastic precomputes (roughplastic.cpp:231-250): eval_transmittance / eval_reflectance of
    include/mitsuba/render/microfacet.h:464-568 (float32 packets, 32 x 32 or 128 x 128 Gauss-Legendre nodes) against the host's
    numpy restatement (mitsuba3_nasa_b200/plastic.py, float64 integrand), and fresnel_diffuse_reflectance (fresnel.h:328-355)."""
    from mitsuba3_nasa_b200 import plastic
    mu = np.maximum(np.float32(1e-6), np.linspace(0, 1, 64, dtype=np.float32))
    for k, (kind, alpha, eta) in enumerate(blocks.PLASTIC_TABLES):
        ext, internal = plastic.rough_tables(kind, alpha, np.float32(eta))
        t, r = golden["plastic_tables/t%d" % k], golden["plastic_tables/r%d" % k]
        assert np.abs(ext - t).max() <= 4e-6, (k, np.abs(ext - t).max())
        assert abs(float(internal) - float(np.float32((r * mu).mean(dtype=np.float32) * 2))) <= 4e-6
    for e, want in zip(blocks.PLASTIC_ETAS, golden["plastic_tables/fdr"]):
        assert plastic.fresnel_diffuse_reflectance(np.float32(e)) == want

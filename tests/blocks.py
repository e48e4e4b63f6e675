"""Building-block harness shared by tests/test_oracle_vs_ref.py and tests/golden/make_golden.py.

Every block is a function (lib, prefix, n) -> dict of output arrays; `prefix` is "orc_"
(oracle/liboracle.so, the restatement) or "ref_" (oracle/_ref/libmiref.so, the reference's
own headers compiled in place).  Inputs are regenerated from fixed seeds, so the golden
file only has to hold the reference outputs.

TOL: None = bit-exact.  A float = max abs error allowed; those blocks differ only in the
association order of 3-term dot products (the reference's scalar Array<float,3> sums
left to right, its JIT arrays and the oracle use fma chains; see oracle/omath.h).
"""
import ctypes as C

import numpy as np


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _unit(rng, n):
    v = rng.normal(size=(n, 3))
    return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)


def _unary(name, lo, hi, seed):
    def run(lib, prefix, n):
        x = np.random.default_rng(seed).uniform(lo, hi, n).astype(np.float32)
        y = np.empty_like(x)
        getattr(lib, prefix + name)(_p(x), _p(y), C.c_size_t(n))
        return {"y": y}
    return run


def b_sincos(lib, prefix, n):
    x = np.random.default_rng(1).uniform(-300, 300, n).astype(np.float32)
    s, c = np.empty_like(x), np.empty_like(x)
    getattr(lib, prefix + "sincos")(_p(x), _p(s), _p(c), C.c_size_t(n))
    return {"s": s, "c": c}


def b_tea(lib, prefix, n):
    rng = np.random.default_rng(2)
    v0 = rng.integers(0, 2 ** 32, n, dtype=np.uint32)
    v1 = rng.integers(0, 2 ** 32, n, dtype=np.uint32)
    v0[:8] = [1, 1, 1, 1, 1, 2, 3, 4]
    v1[:8] = [1, 2, 3, 4, 5, 1, 1, 1]
    o0, o1 = np.empty_like(v0), np.empty_like(v1)
    getattr(lib, prefix + "tea")(_p(v0), _p(v1), _p(o0), _p(o1), C.c_size_t(n))
    return {"o0": o0, "o1": o1}


def b_pcg32(lib, prefix, n):
    out = {}
    for k, (st, sq) in enumerate([(0x853c49e6748fea9b, 0xda3e39cb94b95bdb), (0, 1), (12345, 678910)]):
        f = np.empty(n, np.float32)
        u = np.empty(n, np.uint32)
        getattr(lib, prefix + "pcg32_floats")(C.c_uint64(st), C.c_uint64(sq), _p(f), C.c_size_t(n))
        getattr(lib, prefix + "pcg32_uints")(C.c_uint64(st), C.c_uint64(sq), _p(u), C.c_size_t(n))
        out["f%d" % k] = f
        out["u%d" % k] = u
    return out


def b_coordinate_system(lib, prefix, n):
    v = _unit(np.random.default_rng(3), n)
    s, t = np.empty_like(v), np.empty_like(v)
    getattr(lib, prefix + "coordinate_system")(_p(v), _p(s), _p(t), C.c_size_t(n))
    return {"s": s, "t": t}


def b_unit_angle(lib, prefix, n):
    rng = np.random.default_rng(4)
    a, b = _unit(rng, n), _unit(rng, n)
    o = np.empty(n, np.float32)
    getattr(lib, prefix + "unit_angle")(_p(a), _p(b), _p(o), C.c_size_t(n))
    return {"o": o}


def b_rotator(lib, prefix, n):
    th = np.random.default_rng(5).uniform(-7, 7, n).astype(np.float32)
    m = np.empty((n, 16), np.float32)
    getattr(lib, prefix + "rotator")(_p(th), _p(m), C.c_size_t(n))
    return {"m": m}


def b_rayleigh_scatter(lib, prefix, n):
    ct = np.random.default_rng(6).uniform(-1, 1, n).astype(np.float32)
    st = np.sqrt(np.maximum(0, 1 - ct.astype(np.float64) ** 2)).astype(np.float32)
    out = {}
    for k, rho in enumerate((0.0, 0.03, 0.2)):
        m = np.empty((n, 16), np.float32)
        getattr(lib, prefix + "rayleigh_scatter")(_p(ct), _p(st), C.c_float(rho), _p(m), C.c_size_t(n))
        out["m%d" % k] = m
    return out


def _ortho(rng, fwd):
    r = _unit(rng, len(fwd))
    b = np.cross(fwd, r)
    return (b / np.linalg.norm(b, axis=1, keepdims=True)).astype(np.float32)


def b_rotate_mueller_basis(lib, prefix, n):
    rng = np.random.default_rng(7)
    m_in = rng.uniform(-1, 1, (n, 16)).astype(np.float32)
    fi, fo = _unit(rng, n), _unit(rng, n)
    v = np.concatenate([fi, _ortho(rng, fi), _ortho(rng, fi), fo, _ortho(rng, fo), _ortho(rng, fo)], axis=1).astype(np.float32)
    v = np.ascontiguousarray(v)
    m = np.empty((n, 16), np.float32)
    getattr(lib, prefix + "rotate_mueller_basis")(_p(m_in), _p(v), _p(m), C.c_size_t(n))
    return {"m": m}


def b_mei_to_world_mueller(lib, prefix, n):
    rng = np.random.default_rng(8)
    m_in = rng.uniform(-1, 1, (n, 16)).astype(np.float32)
    d, i, o = _unit(rng, n), _unit(rng, n), _unit(rng, n)
    m = np.empty((n, 16), np.float32)
    getattr(lib, prefix + "mei_to_world_mueller")(_p(m_in), _p(d), _p(i), _p(o), _p(m), C.c_size_t(n))
    return {"m": m}


def _warp(name):
    def run(lib, prefix, n):
        s = np.random.default_rng(9).uniform(0, 1, (n, 2)).astype(np.float32)
        o = np.empty((n, 3), np.float32)
        getattr(lib, prefix + name)(_p(s), _p(o), C.c_size_t(n))
        return {"o": o}
    return run


def b_texture3(lib, prefix, n):
    rng = np.random.default_rng(10)
    out = {}
    k = 0
    for (rx, ry, rz, ch) in ((5, 4, 3, 1), (4, 4, 4, 3)):
        data = rng.uniform(0, 1, rx * ry * rz * ch).astype(np.float32)
        p = rng.uniform(-0.6, 1.6, (n, 3)).astype(np.float32)
        # Only clamp (the gridvolume default) can be pinned this way: the reference's *scalar* instantiation of
        # Texture::wrap (texture.h:1143-1172) aborts with an out-of-range access in repeat and mirror mode even
        # for interior points, so those two modes are restated from the source only.
        for filt in (0, 1):
            for wrap in (2,):
                o = np.zeros((n, ch), np.float32)
                getattr(lib, prefix + "texture3_eval")(_p(data), C.c_int(rx), C.c_int(ry), C.c_int(rz), C.c_int(ch), C.c_int(filt),
                                                        C.c_int(wrap), _p(p), _p(o), C.c_size_t(n))
                out["t%d" % k] = o
                k += 1
    return out


def b_irregular(lib, prefix, n):
    rng = np.random.default_rng(11)
    nodes = np.sort(rng.uniform(-1, 1, 64)).astype(np.float32)
    nodes[0], nodes[-1] = -1.0, 1.0
    pdf = rng.uniform(0.0, 2.0, 64).astype(np.float32)
    pdf[10:14] = 0.0
    x = rng.uniform(-1.1, 1.1, n).astype(np.float32)
    u = rng.uniform(0, 1, n).astype(np.float32)
    ev, evn, sm = np.empty(n, np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
    integral = C.c_float()
    getattr(lib, prefix + "irregular")(_p(nodes), _p(pdf), C.c_uint32(64), _p(x), _p(ev), _p(evn), _p(u), _p(sm), C.c_size_t(n),
                                        C.byref(integral))
    return {"eval": ev, "eval_n": evn, "sample": sm, "integral": np.array([integral.value], np.float32)}


def b_mie(lib, prefix, n):
    """mie() (include/mitsuba/render/mie.h:45-222): size parameters 0.1 ... 450 (the C3 distribution reaches x = 2 pi 40 / 0.55),
    water in air and an absorbing sphere in a denser host; one (mu, radius) pair per call."""
    n = min(n, 600)
    rng = np.random.default_rng(50)
    mu = rng.uniform(-1, 1, n).astype(np.float32)
    mu[:4] = [1.0, -1.0, 0.0, 0.5]
    radius = np.exp(rng.uniform(np.log(0.01), np.log(40.0), n)).astype(np.float32)
    out = {}
    for k, (med, sph, nmax) in enumerate((((1.0, 0.0), (1.33, 0.0), -1), ((1.1, 0.0), (1.5, 0.01), -1), ((1.0, 0.0), (1.33, 0.0), 20))):
        o = np.empty((n, 7), np.float32)
        getattr(lib, prefix + "mie")(C.c_float(0.55), _p(mu), _p(radius), C.c_float(med[0]), C.c_float(med[1]), C.c_float(sph[0]),
                                      C.c_float(sph[1]), C.c_int(nmax), _p(o), C.c_size_t(n))
        out["s%d" % k] = o
    return out


def b_gauss_legendre(lib, prefix, n):
    """quad::gauss_legendre (reference only: the product side is numpy in mitsuba3_nasa_b200/mie.py)."""
    out = {}
    for k in (1, 2, 3, 16, 100):
        x, w = np.empty(k, np.float32), np.empty(k, np.float32)
        lib.ref_gauss_legendre(C.c_int(k), _p(x), _p(w))
        out["x%d" % k] = x
        out["w%d" % k] = w
    return out


PLASTIC_TABLES = [(0, 0.1, 1.49 / 1.000277), (1, 0.3, 1.5), (0, 0.5, 1.33), (1, 0.05, 1.9)]       # (distribution, alpha, eta)
PLASTIC_ETAS = [1.49 / 1.000277, 1.000277 / 1.49, 1.33, 0.7, 2.5]


def b_plastic_tables(lib, prefix, n):
    """eval_transmittance / eval_reflectance over roughplastic's 64 cosines and fresnel_diffuse_reflectance (reference only: the
    product side is numpy in mitsuba3_nasa_b200/plastic.py)."""
    out = {}
    mu = np.maximum(np.float32(1e-6), np.linspace(0, 1, 64, dtype=np.float32))
    lib.ref_rough_transmittance.argtypes = [C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    for k, (kind, alpha, eta) in enumerate(PLASTIC_TABLES):
        t, r = np.zeros(64, np.float32), np.zeros(64, np.float32)
        lib.ref_rough_transmittance(kind, alpha, np.float32(eta), _p(mu), _p(t), _p(r), 64)
        out["t%d" % k] = t
        out["r%d" % k] = r
    lib.ref_fresnel_diffuse_reflectance.restype = C.c_float
    lib.ref_fresnel_diffuse_reflectance.argtypes = [C.c_float]
    out["fdr"] = np.array([lib.ref_fresnel_diffuse_reflectance(np.float32(e)) for e in PLASTIC_ETAS], np.float32)
    return out


def b_mie_scatter(lib, prefix, n):
    rng = np.random.default_rng(51)
    q = rng.uniform(-3, 3, (n, 7)).astype(np.float32)
    q[:, 4] = rng.uniform(0.1, 50, n)
    m = np.empty((n, 16), np.float32)
    getattr(lib, prefix + "mie_scatter")(_p(q), _p(m), C.c_size_t(n))
    return {"m": m}


def b_stokes_basis(lib, prefix, n):
    rng = np.random.default_rng(52)
    f = _unit(rng, n)
    f[:3] = [[0, 0, 1], [0, 0, -1], [1, 0, 0]]
    o = np.empty((n, 3), np.float32)
    getattr(lib, prefix + "stokes_basis")(_p(f), _p(o), C.c_size_t(n))
    return {"o": o}


def b_rotate_stokes_basis(lib, prefix, n):
    rng = np.random.default_rng(53)
    f = _unit(rng, n).astype(np.float64)
    a = np.cross(f, _unit(rng, n)); a /= np.linalg.norm(a, axis=1, keepdims=True)
    ang = rng.uniform(-np.pi, np.pi, (n, 1))
    b = np.cos(ang) * a + np.sin(ang) * np.cross(f, a)
    v = np.ascontiguousarray(np.concatenate([f, a, b], axis=1), dtype=np.float32)
    m = np.empty((n, 16), np.float32)
    getattr(lib, prefix + "rotate_stokes_basis")(_p(v), _p(m), C.c_size_t(n))
    return {"m": m}


def _xform(name):
    def run(lib, prefix, n):
        rng = np.random.default_rng(12)
        m = rng.uniform(-2, 2, 16).astype(np.float32)
        m[12:] = [0, 0, 0, 1]
        p = rng.uniform(-5, 5, (n, 3)).astype(np.float32)
        o = np.empty((n, 3), np.float32)
        getattr(lib, prefix + name)(_p(m), _p(p), _p(o), C.c_size_t(n))
        return {"o": o}
    return run


# Ocean-interface Fresnel blocks (oracle/ofresnel.h).  eta = 1.33 (water seen from the air side) and its
# reciprocal; cos_theta_i spans both hemispheres, grazing angles, normal incidence and total internal reflection.
_ETAS = (1.33, 1.0 / 1.33, 1.5, 1.0)


def _cos_theta(seed, n):
    c = np.random.default_rng(seed).uniform(-1, 1, n).astype(np.float32)
    c[:10] = [1.0, -1.0, 0.0, 1e-9, -1e-9, 1e-4, -1e-4, 0.5, -0.5, 0.6593]
    return c


def _fresnel_block(name, width, seed):
    def run(lib, prefix, n):
        ci = _cos_theta(seed, n)
        out = {}
        for k, eta in enumerate(_ETAS):
            o = np.zeros((n, width), np.float32)
            getattr(lib, prefix + name)(_p(ci), C.c_float(eta), _p(o), C.c_size_t(n))
            out["o%d" % k] = o
        return out
    return run


def b_refract(lib, prefix, n):
    rng = np.random.default_rng(33)
    wi, m = _unit(rng, n), _unit(rng, n)
    out = {}
    for k, eta in enumerate(_ETAS):
        o = np.zeros((n, 3), np.float32)
        getattr(lib, prefix + "refract")(_p(wi), _p(m), C.c_float(eta), _p(o), C.c_size_t(n))
        out["o%d" % k] = o
    return out


# Microfacet distribution of the rough ocean interface (oracle/omicrofacet.h): Beckmann (what the fork's tests use) and GGX,
# isotropic alpha = 0.1 / 0.5 and one anisotropic pair; directions on the upper hemisphere of wi, normals around +z.
_MICROFACETS = ((0, 0.1, 0.1), (0, 0.5, 0.5), (1, 0.1, 0.1), (1, 0.3, 0.6), (0, 0.2, 0.4))


def _upper(rng, n):
    v = _unit(rng, n)
    v[:, 2] = np.abs(v[:, 2]) + np.float32(1e-3)
    return np.ascontiguousarray((v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32))


def b_microfacet_eval(lib, prefix, n):
    rng = np.random.default_rng(40)
    wi = _upper(rng, n)
    m = _unit(rng, n) * np.float32(0.3) + np.array([0, 0, 1], np.float32)
    m = np.ascontiguousarray((m / np.linalg.norm(m, axis=1, keepdims=True)).astype(np.float32))
    out = {}
    for k, (t, au, av) in enumerate(_MICROFACETS):
        o = np.zeros((n, 3), np.float32)
        getattr(lib, prefix + "microfacet_eval")(C.c_int(t), C.c_float(au), C.c_float(av), _p(wi), _p(m), _p(o), C.c_size_t(n))
        out["o%d" % k] = o
    return out


def b_microfacet_sample(lib, prefix, n):
    rng = np.random.default_rng(41)
    wi = _upper(rng, n)
    s = rng.uniform(0, 1, (n, 2)).astype(np.float32)
    out = {}
    for k, (t, au, av) in enumerate(_MICROFACETS):
        o = np.zeros((n, 4), np.float32)
        getattr(lib, prefix + "microfacet_sample")(C.c_int(t), C.c_float(au), C.c_float(av), _p(wi), _p(s), _p(o), C.c_size_t(n))
        out["m%d" % k] = np.ascontiguousarray(o[:, :3])
        out["p%d" % k] = np.ascontiguousarray(o[:, 3] / (1 + np.abs(o[:, 3])))     # densities reach 1e2: compare them on a relative scale
    return out


# name -> (function, tolerance)
BLOCKS = {
    "exp": (_unary("exp", -90, 90, 21), None),
    "log": (_unary("log", 1e-30, 1e5, 22), None),
    "cbrt": (_unary("cbrt", -50, 50, 23), None),
    "asin": (_unary("asin", -1, 1, 24), None),
    "sincos": (b_sincos, None),
    "tea": (b_tea, None),
    "pcg32": (b_pcg32, None),
    "coordinate_system": (b_coordinate_system, None),
    "unit_angle": (b_unit_angle, 4e-7),
    "rotator": (b_rotator, None),
    "rayleigh_scatter": (b_rayleigh_scatter, None),
    "rotate_mueller_basis": (b_rotate_mueller_basis, 5e-6),
    "mei_to_world_mueller": (b_mei_to_world_mueller, 5e-6),
    "square_to_uniform_sphere": (_warp("square_to_uniform_sphere"), None),
    "square_to_cosine_hemisphere": (_warp("square_to_cosine_hemisphere"), 1e-5),   # z = sqrt(1 - squared_norm(p)): fma vs mul+add, amplified by 1/(2z) near the horizon
    "texture3": (b_texture3, None),
    "irregular": (b_irregular, None),
    # Lorenz-Mie series in fp32 as the reference evaluates it; the tolerance is relative to each quantity's scale (see
    # test_oracle_vs_ref.py test_mie_header): the scalar header multiplies complex numbers without fma
    "mie_scatter": (b_mie_scatter, 1e-5),
    "stokes_basis": (b_stokes_basis, None),
    "rotate_stokes_basis": (b_rotate_stokes_basis, 1e-5),
    "transform_point": (_xform("transform_point"), 4e-6),
    "transform_vector": (_xform("transform_vector"), 4e-6),
    "fresnel": (_fresnel_block("fresnel", 4, 30), None),
    # under total internal reflection the complex amplitudes go through squared_norm (fma chain vs left-to-right sum) and
    # rsqrt (sqrt(1/x) vs 1/sqrt(x)); with the oracle switched to the scalar forms these two blocks are bit-exact
    # (test_fresnel_blocks_bit_exact_in_scalar_forms)
    "fresnel_polarized": (_fresnel_block("fresnel_polarized", 7, 31), 5e-7),
    "specular_reflection": (_fresnel_block("specular_reflection", 16, 32), 1e-6),
    "specular_transmission": (_fresnel_block("specular_transmission", 16, 34), None),
    "erf": (_unary("erf", -6, 6, 42), None),
    "erfinv": (_unary("erfinv", -0.999999, 0.999999, 43), None),
    "exp2": (_unary("exp2", -130, 130, 44), None),
    # dot(wi, m) and rsqrt in their JIT forms (oracle) vs scalar forms (reference headers); bit-exact once the oracle is
    # switched to the scalar forms (test_fresnel_blocks_bit_exact_in_scalar_forms).  D reaches 1 / (pi alpha^2) = 32.
    "microfacet_eval": (b_microfacet_eval, 5e-5),
    # visible-normal sampling: the scalar headers' normalize() of the stretched direction differs from every fma / sum
    # order tried by an ulp in 4 % of the samples, and GGX amplifies that near grazing incidence
    "microfacet_sample": (b_microfacet_sample, 1e-4),
    "refract": (b_refract, 5e-5),   # dot(wi, m): fma chain vs left-to-right sum, amplified by sqrt near the critical angle
}


# blocks that are compared by their own tests (tests/test_oracle_vs_ref.py test_mie_header, test_gauss_legendre_header) but
# whose reference outputs live in the same golden file
EXTRA_GOLDEN = {"mie": b_mie, "gauss_legendre": b_gauss_legendre, "plastic_tables": b_plastic_tables}


def compare(name, got, want, tol):
    """Returns (ok, message)."""
    for key in want:
        a, b = np.asarray(got[key]), np.asarray(want[key])
        if a.shape != b.shape:
            return False, "%s.%s: shape %s vs %s" % (name, key, a.shape, b.shape)
        if tol is None:
            if a.dtype.kind == "f":
                same = (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b)) | ((a == 0) & (b == 0))
            else:
                same = a == b
            if not same.all():
                idx = np.argwhere(~same)[0]
                return False, "%s.%s: %d of %d entries differ, first at %s: %r vs %r" % (
                    name, key, (~same).sum(), same.size, tuple(idx), a[tuple(idx)], b[tuple(idx)])
        else:
            err = np.abs(a.astype(np.float64) - b.astype(np.float64))
            err[np.isnan(a) & np.isnan(b)] = 0
            if not (err <= tol).all():
                return False, "%s.%s: max abs err %.3g > %.3g" % (name, key, np.nanmax(err), tol)
    return True, ""

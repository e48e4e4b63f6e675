"""Host-side precomputation of the `plastic` / `roughplastic` BSDFs (src/bsdfs/plastic.cpp:186-200,
src/bsdfs/roughplastic.cpp:215-253): the diffuse Fresnel reflectance fits and, for the rough variant, the table of the
interface's transmittance over 64 cosines and its mean internal reflectance (include/mitsuba/render/microfacet.h:464-568:
a Gauss-Legendre product rule over the visible-normal sampling square).  Runs once per BSDF at scene load, on the host, in
numpy; the kernels only read the results (`mib_bsdf.ext_transmittance`, `.internal_reflectance`).

The reference integrates in float32 packets; here the integrand is evaluated in float64 and the result rounded, which agrees
with the reference's table to ~1e-6 (the CPU oracle's pin on the reference's scalar build holds the tolerance)."""
import numpy as np
from scipy.special import erf, erfinv

RES = 64                       # MI_ROUGH_TRANSMITTANCE_RES (microfacet.h:22)
BECKMANN, GGX = 0, 1


def fresnel_diffuse_reflectance(eta):
    """include/mitsuba/render/fresnel.h:328-355, float32 like the reference (fmadd chains)."""
    f = np.float32
    eta = f(eta)
    inv_eta = f(1.0) / eta
    if eta < 1.0:
        a = f(np.float64(eta) * np.float64(f(-1.4399)) + np.float64(f(0.7099)))
        b = f(np.float64(eta) * np.float64(a) + np.float64(f(0.6681)))
        return f(np.float64(f(0.0636)) * np.float64(inv_eta) + np.float64(b))
    acc = f(-1.36881)                                       # dr::horner: fmadd from the highest coefficient down
    for c in (4.98554, -7.80989, 6.75335, -3.4793, 0.919317):
        acc = f(np.float64(acc) * np.float64(inv_eta) + np.float64(f(c)))
    return acc


def fresnel(cos_theta_i, eta):
    """fresnel.h:41-99 -> (r, cos_theta_t, eta_it, eta_ti)"""
    outside = cos_theta_i >= 0.0
    rcp_eta = 1.0 / eta
    eta_it = np.where(outside, eta, rcp_eta)
    eta_ti = np.where(outside, rcp_eta, eta)
    cos_theta_t_sqr = 1.0 - eta_ti * eta_ti * (1.0 - cos_theta_i * cos_theta_i)
    cos_theta_i_abs = np.abs(cos_theta_i)
    cos_theta_t_abs = np.sqrt(np.maximum(cos_theta_t_sqr, 0.0))
    index_matched = eta == 1.0
    special = index_matched | (cos_theta_i_abs == 0.0)
    with np.errstate(divide='ignore', invalid='ignore'):
        a_s = (eta_it * cos_theta_t_abs - cos_theta_i_abs) / (eta_it * cos_theta_t_abs + cos_theta_i_abs)
        a_p = (eta_it * cos_theta_i_abs - cos_theta_t_abs) / (eta_it * cos_theta_i_abs + cos_theta_t_abs)
    r = 0.5 * (a_s * a_s + a_p * a_p)
    r = np.where(special, 0.0 if index_matched else 1.0, r)
    cos_theta_t = np.where(cos_theta_i >= 0.0, -cos_theta_t_abs, cos_theta_t_abs)
    return r, cos_theta_t, eta_it, eta_ti


def smith_g1(kind, alpha, v, m):
    """microfacet.h:341-366 (isotropic)"""
    xy_alpha_2 = (alpha * v[0]) ** 2 + (alpha * v[1]) ** 2
    with np.errstate(divide='ignore', invalid='ignore'):
        tan_theta_alpha_2 = xy_alpha_2 / (v[2] ** 2)
        if kind == BECKMANN:
            a = 1.0 / np.sqrt(tan_theta_alpha_2)
            a2 = a * a
            res = np.where(a >= 1.6, 1.0, (3.535 * a + 2.181 * a2) / (1.0 + 2.276 * a + 2.577 * a2))
        else:
            res = 2.0 / (1.0 + np.sqrt(1.0 + tan_theta_alpha_2))
    res = np.where(xy_alpha_2 == 0.0, 1.0, res)
    dot = v[0] * m[0] + v[1] * m[1] + v[2] * m[2]
    return np.where(dot * v[2] <= 0.0, 0.0, res)


def _sample_visible_11(kind, cos_theta_i, sx, sy):
    """microfacet.h:368-422"""
    if kind == BECKMANN:
        inv_sqrt_pi = 0.56418958354775628695
        with np.errstate(divide='ignore', invalid='ignore'):
            tan_theta_i = np.sqrt(np.maximum(1.0 - cos_theta_i * cos_theta_i, 0.0)) / cos_theta_i
            cot_theta_i = 1.0 / tan_theta_i
        maxval = erf(cot_theta_i)
        sx = np.clip(sx, 1e-6, 1.0 - 1e-6)
        sy = np.clip(sy, 1e-6, 1.0 - 1e-6)
        x = maxval - (maxval + 1.0) * erf(np.sqrt(-np.log(sx)))
        sx = sx * (1.0 + maxval + inv_sqrt_pi * tan_theta_i * np.exp(-cot_theta_i * cot_theta_i))
        for _ in range(3):
            slope = erfinv(x)
            value = 1.0 + x + inv_sqrt_pi * tan_theta_i * np.exp(-slope * slope) - sx
            derivative = 1.0 - slope * tan_theta_i
            x = x - value / derivative
        return erfinv(x), erfinv(2.0 * sy - 1.0)
    # GGX: concentric disk -> hemisphere of visible normals
    ux, uy = 2.0 * sx - 1.0, 2.0 * sy - 1.0                          # warp.h square_to_uniform_disk_concentric
    is_zero = (ux == 0.0) & (uy == 0.0)
    quadrant_1_or_3 = np.abs(ux) < np.abs(uy)
    r = np.where(quadrant_1_or_3, uy, ux)
    rp = np.where(quadrant_1_or_3, ux, uy)
    with np.errstate(divide='ignore', invalid='ignore'):
        phi = 0.25 * np.pi * rp / r
    phi = np.where(quadrant_1_or_3, 0.5 * np.pi - phi, phi)
    phi = np.where(is_zero, 0.0, phi)
    px, py = r * np.cos(phi), r * np.sin(phi)
    s = 0.5 * (1.0 + cos_theta_i)
    a = np.sqrt(np.maximum(1.0 - px * px, 0.0))
    py = a * (1.0 - s) + py * s
    z = np.sqrt(np.maximum(1.0 - px * px - py * py, 0.0))
    sin_theta_i = np.sqrt(np.maximum(1.0 - cos_theta_i * cos_theta_i, 0.0))
    norm = 1.0 / (sin_theta_i * py + cos_theta_i * z)
    return (cos_theta_i * py - sin_theta_i * z) * norm, px * norm


def sample_visible(kind, alpha, wi, sx, sy):
    """MicrofacetDistribution::sample, visible-normal branch (microfacet.h:244-326), isotropic.  wi: 3 arrays."""
    wp = np.array([alpha * wi[0], alpha * wi[1], wi[2]])
    wp = wp / np.sqrt((wp * wp).sum(axis=0))
    sin_theta_2 = wp[0] * wp[0] + wp[1] * wp[1]
    with np.errstate(divide='ignore', invalid='ignore'):
        inv = 1.0 / np.sqrt(sin_theta_2)
        cos_phi, sin_phi = np.clip(wp[0] * inv, -1.0, 1.0), np.clip(wp[1] * inv, -1.0, 1.0)
    tiny = np.abs(sin_theta_2) <= 4.0 * np.finfo(np.float32).eps / 2
    cos_phi, sin_phi = np.where(tiny, 1.0, cos_phi), np.where(tiny, 0.0, sin_phi)
    slx, sly = _sample_visible_11(kind, wp[2], sx, sy)
    rx = (cos_phi * slx - sin_phi * sly) * alpha
    ry = (sin_phi * slx + cos_phi * sly) * alpha
    m = np.array([-rx, -ry, np.ones_like(rx)])
    return m / np.sqrt((m * m).sum(axis=0))


def _quadrature(res):
    nodes, weights = np.polynomial.legendre.leggauss(res)
    nodes, weights = nodes.astype(np.float32).astype(np.float64), weights.astype(np.float32).astype(np.float64)
    nx, ny = np.meshgrid(nodes, nodes)
    wx, wy = np.meshgrid(weights, weights)
    return 0.5 * nx.reshape(-1) + 0.5, 0.5 * ny.reshape(-1) + 0.5, (wx * wy).reshape(-1) * 0.25


def eval_reflectance(kind, alpha, wi, eta):
    """microfacet.h:464-513.  wi: [3, n] -> [n]"""
    sx, sy, w = _quadrature(32 if eta > 1 else 128)
    out = np.zeros(wi.shape[1])
    for i in range(wi.shape[1]):
        v = wi[:, i:i + 1] * np.ones((1, sx.size))
        m = sample_visible(kind, alpha, v, sx, sy)
        d = (v * m).sum(axis=0)
        wo = 2.0 * d * m - v                                             # reflect(wi, m)
        f = fresnel(d, eta)[0]
        smith = smith_g1(kind, alpha, wo, m) * f
        smith = np.where((wo[2] <= 0.0) | (v[2] <= 0.0), 0.0, smith)
        out[i] = (np.nan_to_num(smith) * w).sum()
    return out


def eval_transmittance(kind, alpha, wi, eta):
    """microfacet.h:516-568"""
    sx, sy, w = _quadrature(32 if eta > 1 else 128)
    out = np.zeros(wi.shape[1])
    for i in range(wi.shape[1]):
        v = wi[:, i:i + 1] * np.ones((1, sx.size))
        m = sample_visible(kind, alpha, v, sx, sy)
        d = (v * m).sum(axis=0)
        f, cos_theta_t, _, eta_ti = fresnel(d, eta)
        wo = m * (d * eta_ti + cos_theta_t) - v * eta_ti                 # refract(wi, m, cos_theta_t, eta_ti), fresnel.h:277-281
        smith = smith_g1(kind, alpha, wo, m) * (1.0 - f)
        smith = np.where(wo[2] * v[2] >= 0.0, 0.0, smith)
        out[i] = (np.nan_to_num(smith) * w).sum()
    return out


def rough_tables(kind, alpha, eta):
    """roughplastic.cpp:231-250 -> (external transmittance [64] float32, mean internal reflectance float32)"""
    alpha = max(float(np.float32(alpha)), 1e-4)
    eta = float(np.float32(eta))
    mu = np.maximum(np.float32(1e-6), np.linspace(0.0, 1.0, RES, dtype=np.float32)).astype(np.float64)
    wi = np.array([np.sqrt(1.0 - mu * mu), np.zeros(RES), mu])
    ext = eval_transmittance(kind, alpha, wi, eta).astype(np.float32)
    refl = eval_reflectance(kind, alpha, wi, float(np.float32(1.0) / np.float32(eta)))
    internal = np.float32((refl * mu).mean() * 2.0)
    return ext, internal

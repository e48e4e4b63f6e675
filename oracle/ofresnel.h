// ofresnel.h -- TEST INFRASTRUCTURE (CPU oracle).  Fresnel building blocks of the
// atmosphere-ocean interface: groundwork for the `dielectric` BSDF and the
// refractive NEE branch of volpathaos (volpathaos.cpp:232-290,312-337).  Nothing on the
// product path calls this yet; the blocks are pinned against the reference headers first
// (tests/test_oracle_vs_ref.py, blocks "fresnel*", "specular_*", "refract").
//
// Restates, for real-valued eta:
//   fresnel                 include/mitsuba/render/fresnel.h:34-71
//   fresnel_polarized       include/mitsuba/render/fresnel.h:149-190
//   reflect / refract       include/mitsuba/render/fresnel.h:274-316
//   specular_reflection     include/mitsuba/render/mueller.h:198-224
//   specular_transmission   include/mitsuba/render/mueller.h:240-268
// with the complex arithmetic of ext/drjit/include/drjit/complex.h (operator* :80-87,
// rcp :101-103, sqrt :161-173, sincos_arg_diff :149-154).
#pragma once
#include "ospec.h"
#include "omie.h"   // Cx and its complex.h arithmetic

namespace orc {

static inline Cx cmul(Cx a, Cx b) { return a * b; }
static inline Cx cconj(Cx a) { return { a.re, -a.im }; }
// squared_norm(Array<T,2>) is an fma chain in the JIT variants (the parity target); the reference's *scalar* instantiation
// sums the products left to right.  Likewise rsqrt: sqrt(1/x) in the LLVM backend (omath.h), rcp(sqrt(x)) in the scalar
// headers (array_utils.h:216-218).  g_fresnel_scalar_forms (test-only, orc_fresnel_scalar_forms) switches to the scalar forms:
// with it the blocks equal the reference's scalar headers bit for bit, which proves that these two are the only differences.
static int g_fresnel_scalar_forms = 0;
static inline float csqnorm(Cx a) { return g_fresnel_scalar_forms ? a.re * a.re + a.im * a.im : cx_sqnorm(a); }
static inline float crsqrt(float a) { return g_fresnel_scalar_forms ? 1.f / std::sqrt(a) : rsqrt(a); }
static inline Cx cdiv(Cx a, Cx b) { float r = rcp(csqnorm(b)); return a * Cx{ b.re * r, -b.im * r }; }
// complex.h:161-173
static inline Cx csqrt(Cx z) {
    float n  = std::sqrt(csqnorm(z)),
          t1 = std::sqrt(.5f * (n + std::fabs(z.re))),
          t2 = .5f * z.im / t1;
    bool zero = n == 0.f, m = z.re >= 0.f;
    return { m ? t1 : std::fabs(t2), zero ? 0.f : (m ? t2 : (z.im >= 0.f ? t1 : -t1)) };
}

struct Fresnel { float r, cos_theta_t, eta_it, eta_ti; };

// fresnel.h:34-71
static inline Fresnel fresnel(float cos_theta_i, float eta) {
    bool outside = cos_theta_i >= 0.f;
    float rcp_eta = rcp(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fnmadd(fnmadd(cos_theta_i, cos_theta_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = std::fabs(cos_theta_i), ct = safe_sqrt(cos_theta_t_sqr);
    bool index_matched = eta == 1.f, special = index_matched || ci == 0.f;
    float a_s = fnmadd(eta_it, ct, ci) / fmadd(eta_it, ct, ci);
    float a_p = fnmadd(eta_it, ci, ct) / fmadd(eta_it, ci, ct);
    float r = 0.5f * (sqr(a_s) + sqr(a_p));
    if (special) r = index_matched ? 0.f : 1.f;
    return { r, mulsign_neg_ad(ct, cos_theta_i), eta_it, eta_ti };
}

struct FresnelPol { Cx a_s, a_p; float cos_theta_t, eta_it, eta_ti; };

// fresnel.h:149-190
static inline FresnelPol fresnel_polarized(float cos_theta_i, float eta) {
    bool outside = cos_theta_i >= 0.f;
    float rcp_eta = rcp(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fnmadd(fnmadd(cos_theta_i, cos_theta_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = std::fabs(cos_theta_i);
    Cx ct = csqrt(Cx{ cos_theta_t_sqr, 0.f });
    // sign of the root under total internal reflection; llvm_ad (is_diff) forms of mulsign / copysign: select(v2 >= 0, ...)
    ct = { mulsign_ad(ct.re, cos_theta_t_sqr), mulsign_ad(ct.im, cos_theta_t_sqr) };
    Cx e_ct = { eta_it * ct.re, eta_it * ct.im };
    Cx a_s = cdiv(Cx{ ci - e_ct.re, 0.f - e_ct.im }, Cx{ ci + e_ct.re, 0.f + e_ct.im });
    Cx a_p = cdiv(Cx{ eta_it * ci - ct.re, 0.f - ct.im }, Cx{ eta_it * ci + ct.re, 0.f + ct.im });
    if (eta == 1.f || eta == 0.f) { a_s = { 0.f, 0.f }; a_p = { 0.f, 0.f }; }
    float ct_signed = cos_theta_t_sqr >= 0.f ? mulsign_neg_ad(ct.re, cos_theta_i) : 0.f;
    return { a_s, a_p, ct_signed, eta_it, eta_ti };
}

// mueller.h:198-224
static inline Mueller specular_reflection(float cos_theta_i, float eta) {
    FresnelPol f = fresnel_polarized(cos_theta_i, eta);
    // sincos_arg_diff(a_p, a_s), complex.h:149-154
    float normalization = crsqrt(csqnorm(f.a_p) * csqnorm(f.a_s));
    Cx v = cmul(f.a_p, cconj(f.a_s));
    float sin_delta = v.im * normalization, cos_delta = v.re * normalization;
    Cx s2 = cmul(f.a_s, f.a_s), p2 = cmul(f.a_p, f.a_p);
    float r_s = std::sqrt(csqnorm(s2)), r_p = std::sqrt(csqnorm(p2)),
          a = 0.5f * (r_s + r_p), b = 0.5f * (r_s - r_p), c = std::sqrt(r_s * r_p);
    if (c == 0.f) { sin_delta = 0.f; cos_delta = 0.f; }
    Mueller M;
    M.m[0][0] = a; M.m[0][1] = b; M.m[1][0] = b; M.m[1][1] = a;
    M.m[2][2] = c * cos_delta; M.m[2][3] = -c * sin_delta;
    M.m[3][2] = c * sin_delta; M.m[3][3] = c * cos_delta;
    return M;
}

// mueller.h:240-268
static inline Mueller specular_transmission(float cos_theta_i, float eta) {
    FresnelPol f = fresnel_polarized(cos_theta_i, eta);
    float factor = -f.eta_it * (std::fabs(cos_theta_i) > 1e-8f ? f.cos_theta_t / cos_theta_i : 0.f);
    float a_s_r = 1.f + f.a_s.re, a_p_r = (1.f + f.a_p.re) * f.eta_ti;
    float t_s = sqr(a_s_r), t_p = sqr(a_p_r),
          a = 0.5f * factor * (t_s + t_p), b = 0.5f * factor * (t_s - t_p), c = factor * std::sqrt(t_s * t_p);
    Mueller M;
    M.m[0][0] = a; M.m[0][1] = b; M.m[1][0] = b; M.m[1][1] = a; M.m[2][2] = c; M.m[3][3] = c;
    return M;
}

// fresnel.h:274-278 (local frame)
static inline Vec3 reflect(Vec3 wi) { return { -wi.x, -wi.y, wi.z }; }
// fresnel.h:296-299 (local frame; cos_theta_t and eta_ti as returned by fresnel())
static inline Vec3 refract(Vec3 wi, float cos_theta_t, float eta_ti) { return { -eta_ti * wi.x, -eta_ti * wi.y, cos_theta_t }; }
// fresnel.h:313-316: refract wi with respect to a normal m
static inline Vec3 refract(Vec3 wi, Vec3 m, float cos_theta_t, float eta_ti) {
    float s = fmadd(dot(wi, m), eta_ti, cos_theta_t);
    return { fmsub(m.x, s, wi.x * eta_ti), fmsub(m.y, s, wi.y * eta_ti), fmsub(m.z, s, wi.z * eta_ti) };
}

// ---------------------------------------------------------------------------------------------------------------
// SmoothDielectric::sample, src/bsdfs/dielectric.cpp:247-371, TransportMode::Radiance, specular_reflectance /
// specular_transmittance as plain factors.  has_reflection / has_transmission mirror ctx.is_enabled (volpathaos' refractive
// NEE asks for transmission only, volpathaos.cpp:320-326).  pdf() of this BSDF is zero (:465-468); eval() is dielectric_eval below.
// ---------------------------------------------------------------------------------------------------------------
enum : uint32_t { BSDF_DeltaReflection = 0x20, BSDF_DeltaTransmission = 0x40 };

struct DielectricSample {
    Vec3 wo; float pdf = 0.f, eta = 1.f; uint32_t sampled_type = 0;
    float weight = 0.f;      // unpolarized variants
    Mueller weight_pol;      // polarized variants (rotated to the implicit Stokes bases of -wo_hat and wi_hat)
};

static inline DielectricSample dielectric_sample(Vec3 wi, float eta, float sample1, bool has_reflection, bool has_transmission,
                                                 float reflectance = 1.f, float transmittance = 1.f) {
    DielectricSample bs;
    Fresnel f = fresnel(wi.z, eta);
    float r_i = f.r, t_i = 1.f - r_i;
    bool selected_r;
    if (has_reflection && has_transmission) { selected_r = sample1 <= r_i; bs.pdf = selected_r ? r_i : t_i; }
    else if (has_reflection || has_transmission) { selected_r = has_reflection; bs.pdf = 1.f; }
    else return bs;
    bs.sampled_type = selected_r ? BSDF_DeltaReflection : BSDF_DeltaTransmission;
    bs.wo = selected_r ? reflect(wi) : refract(wi, f.cos_theta_t, f.eta_ti);
    bs.eta = selected_r ? 1.f : f.eta_it;

    // polarized weight (:296-339): light arrives along -wo_hat and leaves along +wi_hat
    Vec3 wo_hat = bs.wo, wi_hat = wi;
    Mueller R = specular_reflection(wo_hat.z, eta), T = specular_transmission(wo_hat.z, eta);
    Mueller W = selected_r ? R : T;
    if (has_reflection && has_transmission)
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) W.m[i][j] = W.m[i][j] / bs.pdf;
    Vec3 n(0.f, 0.f, 1.f);
    Vec3 s_axis_in = cross(n, -wo_hat), s_axis_out = cross(n, wi_hat);
    bool collinear = s_axis_in.x == 0.f && s_axis_in.y == 0.f && s_axis_in.z == 0.f;
    s_axis_in  = collinear ? Vec3(1.f, 0.f, 0.f) : normalize(s_axis_in);
    s_axis_out = collinear ? Vec3(1.f, 0.f, 0.f) : normalize(s_axis_out);
    W = rotate_mueller_basis(W, -wo_hat, s_axis_in, stokes_basis(-wo_hat), wi_hat, s_axis_out, stokes_basis(wi_hat));
    W = W * (selected_r ? reflectance : transmittance);

    // unpolarized weight (:341-362)
    float w = (has_reflection && has_transmission) ? 1.f : (has_reflection ? r_i : t_i);
    w *= selected_r ? reflectance : transmittance;

    if (!selected_r) {          // solid-angle compression across the interface (:364-369)
        float fac = sqr(f.eta_ti);
        w *= fac; W = W * fac;
    }
    bs.weight = w; bs.weight_pol = W;
    return bs;
}

// SmoothDielectric::eval, src/bsdfs/dielectric.cpp:374-463.  In the fork this is NOT zero: it returns the weight of the lobe
// that (wi, wo) falls into as if the delta function had been hit -- volpathaos' refractive NEE aims its shadow ray so that it
// is (volpathaos.cpp:312-337, :423-431).  Quirks kept: no Fresnel factor and no division by a pdf in the unpolarized form
// (`weight = 1` for every lane when both lobes are enabled, including lanes with cos_theta_i == 0); no special case for
// wo parallel to the normal in the polarized form (normalize(0) -> NaN).
static inline void dielectric_eval(Vec3 wi, Vec3 wo, float eta, float inv_eta, bool has_reflection, bool has_transmission, float *value, Mueller *value_pol) {
    float cos_theta_i = wi.z, cos_theta_o = wo.z;
    bool active = cos_theta_i != 0.f;
    bool reflect = cos_theta_i * cos_theta_o > 0.f;
    float inv_eta_side = cos_theta_i > 0.f ? inv_eta : eta;          // m_inv_eta = ext_ior / int_ior (:223)
    float F = fresnel(cos_theta_i, eta).r;
    bool eval_r = has_reflection && reflect && active, eval_t = has_transmission && !reflect && active;
    // unpolarized (:442-454)
    float w = 0.f;
    if (has_reflection && has_transmission) w = 1.f;
    else if (has_reflection || has_transmission) w = has_reflection ? F : 1.f - F;
    // polarized (:409-440)
    Vec3 wo_hat = wo, wi_hat = wi;
    Mueller W;                                                        // zero
    if (eval_r) W = specular_reflection(wo_hat.z, eta);
    if (eval_t) W = specular_transmission(wo_hat.z, eta);
    Vec3 n(0.f, 0.f, 1.f);
    Vec3 s_axis_in = normalize(cross(n, -wo_hat)), s_axis_out = normalize(cross(n, wi_hat));
    W = rotate_mueller_basis(W, -wo_hat, s_axis_in, stokes_basis(-wo_hat), wi_hat, s_axis_out, stokes_basis(wi_hat));
    if (eval_t) { float fac = sqr(inv_eta_side); w *= fac; W = W * fac; }
    *value = w; *value_pol = W;
}

} // namespace orc

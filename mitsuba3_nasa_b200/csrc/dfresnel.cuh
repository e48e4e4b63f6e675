// dfresnel.cuh -- Fresnel blocks of the atmosphere-ocean interface on the device (SURVEY 8f rank 2 groundwork).
//   fresnel                 include/mitsuba/render/fresnel.h:34-71
//   fresnel_polarized       include/mitsuba/render/fresnel.h:149-190      (real eta)
//   specular_reflection     include/mitsuba/render/mueller.h:198-224
//   specular_transmission   include/mitsuba/render/mueller.h:240-268
//   refract                 include/mitsuba/render/fresnel.h:313-316
// with the complex arithmetic of ext/drjit/include/drjit/complex.h in its JIT (llvm_ad) form: fma-chain squared_norm,
// rsqrt = sqrt(1/x), mulsign / copysign as selects.  No walk kernel calls these yet (there is no `dielectric` plugin);
// mib_debug_fresnel runs them on arrays so that tests can compare them with oracle/ofresnel.h value for value.
#pragma once
#include "dmath.cuh"

namespace mib {

struct Cx { float re, im; };
MIB_D Cx cmul_(Cx a, Cx b) { return { fms_(a.re, b.re, a.im * b.im), fma_(a.re, b.im, a.im * b.re) }; }      // complex.h:80-87
MIB_D float csqnorm_(Cx a) { return fma_(a.im, a.im, a.re * a.re); }
MIB_D Cx cdiv_(Cx a, Cx b) { float r = rcp_(csqnorm_(b)); return cmul_(a, Cx{ b.re * r, -b.im * r }); }    // complex.h:101-109
MIB_D Cx csqrt_(Cx z) {                                                                                     // complex.h:161-173
    float n = sqrt_(csqnorm_(z)), t1 = sqrt_(.5f * (n + fabsf(z.re))), t2 = div_(.5f * z.im, t1);
    bool zero = n == 0.f, m = z.re >= 0.f;
    return { m ? t1 : fabsf(t2), zero ? 0.f : (m ? t2 : (z.im >= 0.f ? t1 : -t1)) };
}

struct FresnelPol { Cx a_s, a_p; float cos_theta_t, eta_it, eta_ti; };

// out: (F, cos_theta_t, eta_it, eta_ti)
MIB_D float4 fresnel_(float cos_theta_i, float eta) {
    bool outside = cos_theta_i >= 0.f;
    float rcp_eta = rcp_(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fnma_(fnma_(cos_theta_i, cos_theta_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = fabsf(cos_theta_i), ct = safe_sqrt_(cos_theta_t_sqr);
    bool index_matched = eta == 1.f, special = index_matched || ci == 0.f;
    float a_s = div_(fnma_(eta_it, ct, ci), fma_(eta_it, ct, ci));
    float a_p = div_(fnma_(eta_it, ci, ct), fma_(eta_it, ci, ct));
    float r = 0.5f * (sqr_(a_s) + sqr_(a_p));
    if (special) r = index_matched ? 0.f : 1.f;
    return make_float4(r, cos_theta_i >= 0.f ? -ct : ct, eta_it, eta_ti);
}

MIB_D FresnelPol fresnel_polarized_(float cos_theta_i, float eta) {
    bool outside = cos_theta_i >= 0.f;
    float rcp_eta = rcp_(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fnma_(fnma_(cos_theta_i, cos_theta_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = fabsf(cos_theta_i);
    Cx ct = csqrt_(Cx{ cos_theta_t_sqr, 0.f });
    ct = { mulsign_ad(ct.re, cos_theta_t_sqr), mulsign_ad(ct.im, cos_theta_t_sqr) };
    Cx e_ct = { eta_it * ct.re, eta_it * ct.im };
    Cx a_s = cdiv_(Cx{ ci - e_ct.re, 0.f - e_ct.im }, Cx{ ci + e_ct.re, 0.f + e_ct.im });
    Cx a_p = cdiv_(Cx{ eta_it * ci - ct.re, 0.f - ct.im }, Cx{ eta_it * ci + ct.re, 0.f + ct.im });
    if (eta == 1.f || eta == 0.f) { a_s = { 0.f, 0.f }; a_p = { 0.f, 0.f }; }
    float ct_signed = cos_theta_t_sqr >= 0.f ? (cos_theta_i >= 0.f ? -ct.re : ct.re) : 0.f;
    return { a_s, a_p, ct_signed, eta_it, eta_ti };
}

// Both matrices have the block form [[a, b, 0, 0], [b, a, 0, 0], [0, 0, c, -s], [0, 0, s, c]]; out = (a, b, c, s).
MIB_D float4 specular_reflection_(float cos_theta_i, float eta) {
    FresnelPol f = fresnel_polarized_(cos_theta_i, eta);
    float normalization = rsqrt_(csqnorm_(f.a_p) * csqnorm_(f.a_s));       // sincos_arg_diff, complex.h:149-154
    Cx v = cmul_(f.a_p, Cx{ f.a_s.re, -f.a_s.im });
    float sin_delta = v.im * normalization, cos_delta = v.re * normalization;
    float r_s = sqrt_(csqnorm_(cmul_(f.a_s, f.a_s))), r_p = sqrt_(csqnorm_(cmul_(f.a_p, f.a_p)));
    float a = 0.5f * (r_s + r_p), b = 0.5f * (r_s - r_p), c = sqrt_(r_s * r_p);
    if (c == 0.f) { sin_delta = 0.f; cos_delta = 0.f; }
    return make_float4(a, b, c * cos_delta, c * sin_delta);
}

MIB_D float4 specular_transmission_(float cos_theta_i, float eta) {
    FresnelPol f = fresnel_polarized_(cos_theta_i, eta);
    float factor = -f.eta_it * (fabsf(cos_theta_i) > 1e-8f ? div_(f.cos_theta_t, cos_theta_i) : 0.f);
    float a_s_r = 1.f + f.a_s.re, a_p_r = (1.f + f.a_p.re) * f.eta_ti;
    float t_s = sqr_(a_s_r), t_p = sqr_(a_p_r);
    return make_float4(0.5f * factor * (t_s + t_p), 0.5f * factor * (t_s - t_p), factor * sqrt_(t_s * t_p), 0.f);
}

// refract wi about the normal m; cos_theta_t and eta_ti as returned by fresnel_(dot(wi, m), eta)
MIB_D float3 refract_(float3 wi, float3 m, float cos_theta_t, float eta_ti) {
    float d = fma_(wi.z, m.z, fma_(wi.y, m.y, wi.x * m.x));
    float s = fma_(d, eta_ti, cos_theta_t);
    return make_float3(fms_(m.x, s, wi.x * eta_ti), fms_(m.y, s, wi.y * eta_ti), fms_(m.z, s, wi.z * eta_ti));
}

} // namespace mib

// omicrofacet.h -- TEST INFRASTRUCTURE (CPU oracle).  Microfacet distribution of the rough ocean interface
// (`roughdielectric`, src/bsdfs/roughdielectric.cpp, is what the fork's atmosphere-ocean tests use,
// test_ad_volpathaos.py:234-338, Beckmann, alpha = 0.1).  Restates MicrofacetDistribution,
// include/mitsuba/render/microfacet.h:185-440 (eval, pdf, smith_g1, G, visible-normal sampling for Beckmann and GGX),
// with the Dr.Jit polynomials it needs: exp2 (drjit/math.h:756-820), erf (:1473-1527), erfinv (:1531-1552).
// Pinned bit for bit against the reference headers in tests/test_oracle_vs_ref.py (blocks "microfacet_*", "erf*").
// sample_visible = false (microfacet.h:246-300, :219-226): the normals are drawn from D * cos(theta_m), isotropic roughness only.
#pragma once
#include "omath.h"
#include "oplugins.h"   // square_to_uniform_disk_concentric (warp.h:54-90)
#include "ofresnel.h"   // g_fresnel_scalar_forms

namespace orc {

// drjit/math.h:756-820, single precision
static inline float exp2_(float x) {
    bool over = x > 127.f, under = x < -127.f;
    float n = std::floor(x), y = x - n;
    if (y > .5f) { n += 1.f; y -= 1.f; }
    const float c[6] = { 6.931472028550421e-1f, 2.402264791363012e-1f, 5.550332471162809e-2f,
                         9.618437357674640e-3f, 1.339887440266574e-3f, 1.535336188319500e-4f };
    float z = fmadd(y, estrin(y, c, 6), 1.f);
    return over ? Inf : (under ? 0.f : ldexp_(z, n));
}

// drjit/math.h:1473-1527, single precision
static inline float erf_(float x) {
    float xa = std::fabs(x), x2 = sqr(x);
    const float a[6] = { 0x1.20dd74p+0f, -0x1.812672p-2f, 0x1.ce0934p-4f, -0x1.b5a334p-6f, 0x1.4246b4p-8f, -0x1.273facp-11f };
    const float b[7] = { -0x1.a0d71ap+0f, -0x1.d51e3ap-1f, -0x1.3a904cp-3f, 0x1.1c395cp-5f, -0x1.6856bep-8f, 0x1.180f1ep-11f,
                         -0x1.8ca9f6p-16f };
    float c0 = estrin(x2, a, 6), c1 = estrin(xa, b, 7);
    float xb = 1.f - exp2_(c1 * xa);
    if (xa < 1.f) return x * c0;
    return copysign_bits(std::isfinite(xb) ? xb : 1.f, x);
}

// drjit/math.h:1531-1552
static inline float erfinv_(float x) {
    float w = -log_((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = std::sqrt(w) - 3.f;
    const float p[9] = { 1.50140941f, 0.246640727f, -0.00417768164f, -0.00125372503f, 0.00021858087f, -4.39150654e-06f,
                         -3.5233877e-06f, 3.43273939e-07f, 2.81022636e-08f };
    const float q[9] = { 2.83297682f, 1.00167406f, 0.00943887047f, -0.0076224613f, 0.00573950773f, -0.00367342844f,
                         0.00134934322f, 0.000100950558f, -0.000200214257f };
    return (w < 5.f ? estrin(w1, p, 9) : estrin(w2, q, 9)) * x;
}

enum { MICROFACET_BECKMANN = 0, MICROFACET_GGX = 1 };

// JIT forms (parity target) vs the scalar forms of the reference headers, as in ofresnel.h: dot is an fma chain vs a
// left-to-right sum, rsqrt is sqrt(1/x) vs 1/sqrt(x) (squared_norm of a 3-vector is an fma chain in both, of a 2-vector it follows dot).  With g_fresnel_scalar_forms set the blocks below equal the scalar
// reference headers bit for bit (tests/test_oracle_vs_ref.py).
static inline float md_dot(Vec3 a, Vec3 b) { return g_fresnel_scalar_forms ? a.x * b.x + a.y * b.y + a.z * b.z : dot(a, b); }
static inline float md_rsqrt(float a) { return g_fresnel_scalar_forms ? 1.f / std::sqrt(a) : rsqrt(a); }
static inline Vec3 md_normalize(Vec3 v) { return v * md_rsqrt(squared_norm(v)); }   // squared_norm: an fma chain in both forms (measured)
static inline float md_sqnorm2(float x, float y) { return g_fresnel_scalar_forms ? x * x + y * y : fmadd(y, y, sqr(x)); }   // Point2f: left-to-right in the scalar headers (measured)

struct Microfacet {
    int type; float alpha_u, alpha_v; bool sample_visible;
    // microfacet.h:76-99 + configure() :425-428
    Microfacet(int t, float au, float av, bool sv = true) : type(t), alpha_u(maximum(au, 1e-4f)), alpha_v(maximum(av, 1e-4f)), sample_visible(sv) {}

    // microfacet.h:185-207
    float eval(Vec3 m) const {
        float alpha_uv = alpha_u * alpha_v, cos_theta = m.z, cos_theta_2 = sqr(cos_theta), result;
        if (type == MICROFACET_BECKMANN)
            result = exp_(-(sqr(m.x / alpha_u) + sqr(m.y / alpha_v)) / cos_theta_2) / (Pi * alpha_uv * sqr(cos_theta_2));
        else
            result = rcp(Pi * alpha_uv * sqr(sqr(m.x / alpha_u) + sqr(m.y / alpha_v) + sqr(m.z)));
        return result * cos_theta > 1e-20f ? result : 0.f;
    }

    // microfacet.h:341-366
    float smith_g1(Vec3 v, Vec3 m) const {
        float xy_alpha_2 = sqr(alpha_u * v.x) + sqr(alpha_v * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqr(v.z), result;
        if (type == MICROFACET_BECKMANN) {
            float a = md_rsqrt(tan_theta_alpha_2), a_sqr = sqr(a);
            result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
        } else {
            result = 2.f / (1.f + std::sqrt(1.f + tan_theta_alpha_2));
        }
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (md_dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    float G(Vec3 wi, Vec3 wo, Vec3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }

    // microfacet.h:219-226
    float pdf(Vec3 wi, Vec3 m) const {
        if (!sample_visible) return eval(m) * m.z;
        return eval(m) * (smith_g1(wi, m) * std::fabs(md_dot(wi, m)) / wi.z);
    }

    // microfacet.h:368-422
    void sample_visible_11(float cos_theta_i, float sx, float sy, float *slope_x, float *slope_y) const {
        if (type == MICROFACET_BECKMANN) {
            const float InvSqrtPi = 0.56418958354775628695f;
            float tan_theta_i = safe_sqrt(fnmadd(cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i;
            float cot_theta_i = rcp(tan_theta_i);
            float maxval = erf_(cot_theta_i);
            sx = maximum(minimum(sx, 1.f - 1e-6f), 1e-6f);
            sy = maximum(minimum(sy, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * erf_(std::sqrt(-log_(sx)));
            sx *= 1.f + maxval + InvSqrtPi * tan_theta_i * exp_(-sqr(cot_theta_i));
            for (int i = 0; i < 3; ++i) {
                float slope = erfinv_(x),
                      value = 1.f + x + InvSqrtPi * tan_theta_i * exp_(-sqr(slope)) - sx,
                      derivative = 1.f - slope * tan_theta_i;
                x -= value / derivative;
            }
            *slope_x = erfinv_(x);
            *slope_y = erfinv_(fmsub(2.f, sy, 1.f));
        } else {
            float px, py;
            square_to_uniform_disk_concentric(sx, sy, &px, &py);
            float s = 0.5f * (1.f + cos_theta_i);
            float a = safe_sqrt(1.f - sqr(px));
            py = fmadd(py, s, fnmadd(a, s, a));               // dr::lerp(a, b, t) = fmadd(b, t, fnmadd(a, t, a))
            float x = px, y = py, z = safe_sqrt(1.f - md_sqnorm2(px, py));
            float sin_theta_i = safe_sqrt(1.f - sqr(cos_theta_i));
            float norm = rcp(fmadd(sin_theta_i, y, cos_theta_i * z));
            *slope_x = fmsub(cos_theta_i, y, sin_theta_i * z) * norm;
            *slope_y = x * norm;
        }
    }

    // microfacet.h:244-326, visible-normal branch.  Returns the microfacet normal; *pdf_out as the reference computes it.
    Vec3 sample(Vec3 wi, float sx, float sy, float *pdf_out) const {
        if (!sample_visible) {
            // :246-293 (isotropic branch; anisotropic roughness with sample_visible = false is refused by the host)
            float sin_phi, cos_phi;
            sincos((2.f * Pi) * sy, &sin_phi, &cos_phi);
            const float alpha_2 = alpha_u * alpha_u;
            float cos_theta, cos_theta_2, pdf;
            if (type == MICROFACET_BECKMANN) {
                cos_theta = rsqrt(fnmadd(alpha_2, log_(1.f - sx), 1.f));
                cos_theta_2 = sqr(cos_theta);
                float cos_theta_3 = maximum(cos_theta_2 * cos_theta, 1e-20f);
                pdf = (1.f - sx) / (Pi * alpha_u * alpha_v * cos_theta_3);
            } else {
                float tan_theta_m_2 = alpha_2 * sx / (1.f - sx);
                cos_theta = rsqrt(1.f + tan_theta_m_2);
                cos_theta_2 = sqr(cos_theta);
                float temp = 1.f + tan_theta_m_2 / alpha_2, cos_theta_3 = maximum(cos_theta_2 * cos_theta, 1e-20f);
                pdf = rcp(Pi * alpha_u * alpha_v * cos_theta_3 * sqr(temp));
            }
            float sin_theta = std::sqrt(1.f - cos_theta_2);
            *pdf_out = pdf;
            return Vec3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
        }
        Vec3 wi_p = md_normalize(Vec3(alpha_u * wi.x, alpha_v * wi.y, wi.z));
        // Frame::sincos_phi, include/mitsuba/core/frame.h:111-122
        float sin_theta_2 = fmadd(wi_p.x, wi_p.x, sqr(wi_p.y)), inv_sin_theta = md_rsqrt(sin_theta_2);
        float cos_phi = wi_p.x * inv_sin_theta, sin_phi = wi_p.y * inv_sin_theta;
        if (std::fabs(sin_theta_2) <= 4.f * EpsilonF) { cos_phi = 1.f; sin_phi = 0.f; }
        else { cos_phi = maximum(minimum(cos_phi, 1.f), -1.f); sin_phi = maximum(minimum(sin_phi, 1.f), -1.f); }
        float slx, sly;
        sample_visible_11(wi_p.z, sx, sy, &slx, &sly);
        float rx = fmsub(cos_phi, slx, sin_phi * sly) * alpha_u, ry = fmadd(sin_phi, slx, cos_phi * sly) * alpha_v;
        Vec3 m = md_normalize(Vec3(-rx, -ry, 1.f));
        *pdf_out = eval(m) * smith_g1(wi, m) * std::fabs(md_dot(wi, m)) / wi.z;
        return m;
    }
};

} // namespace orc

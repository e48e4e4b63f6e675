// dmicrofacet.cuh -- microfacet distribution of the rough ocean interface on the device (SURVEY 8f rank 2 groundwork).
// MicrofacetDistribution, include/mitsuba/render/microfacet.h:185-440 (Beckmann / GGX: eval, smith_g1, pdf, visible-normal
// sampling), with Dr.Jit's exp2 (drjit/math.h:756-820), erf (:1473-1527) and erfinv (:1531-1552); JIT forms (fma-chain dot,
// rsqrt = sqrt(1/x)).  Mirrors oracle/omicrofacet.h operation by operation.  No walk kernel includes this header yet;
// mib_debug_microfacet runs it on arrays for the comparison with the oracle.
#pragma once
#include "dmath.cuh"
#include "dplugins.cuh"   // square_to_uniform_disk_concentric

namespace mib {

// drjit/math.h:24-41: Estrin's scheme exactly as the recursive template evaluates it (same loop as oracle/omath.h estrin)
template <int N> MIB_D float estrin_(float x, const float (&coeff)[N]) {
    float buf[N], rec[N];
    for (int i = 0; i < N; ++i) buf[i] = coeff[i];
    int n = N;
    while (true) {
        int n_rec = (n - 1) / 2, n_fma = n / 2;
        for (int i = 0; i < n_fma; ++i) rec[i] = fma_(x, buf[2 * i + 1], buf[2 * i]);
        if (n_rec == n_fma) rec[n_rec] = buf[n - 1];
        if (n_rec == 0) return rec[0];
        x = x * x;
        n = n_rec + 1;
        for (int i = 0; i < n; ++i) buf[i] = rec[i];
    }
}

MIB_D float exp2_(float x) {
    bool over = x > 127.f, under = x < -127.f;
    float n = floorf(x), y = x - n;
    if (y > .5f) { n += 1.f; y -= 1.f; }
    const float c[6] = { 6.931472028550421e-1f, 2.402264791363012e-1f, 5.550332471162809e-2f,
                         9.618437357674640e-3f, 1.339887440266574e-3f, 1.535336188319500e-4f };
    float z = fma_(y, estrin_(y, c), 1.f);
    return over ? inf_f() : (under ? 0.f : ldexp_(z, n));
}

MIB_D float erf_(float x) {
    float xa = fabsf(x), x2 = x * x;
    const float a[6] = { 0x1.20dd74p+0f, -0x1.812672p-2f, 0x1.ce0934p-4f, -0x1.b5a334p-6f, 0x1.4246b4p-8f, -0x1.273facp-11f };
    const float b[7] = { -0x1.a0d71ap+0f, -0x1.d51e3ap-1f, -0x1.3a904cp-3f, 0x1.1c395cp-5f, -0x1.6856bep-8f, 0x1.180f1ep-11f,
                         -0x1.8ca9f6p-16f };
    float c0 = estrin_(x2, a), c1 = estrin_(xa, b);
    float xb = 1.f - exp2_(c1 * xa);
    if (xa < 1.f) return x * c0;
    float r = isfinite(xb) ? xb : 1.f;
    return __int_as_float((__float_as_int(r) & 0x7fffffff) | (__float_as_int(x) & 0x80000000));
}

MIB_D float erfinv_(float x) {
    float w = -log_((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = sqrt_(w) - 3.f;
    const float p[9] = { 1.50140941f, 0.246640727f, -0.00417768164f, -0.00125372503f, 0.00021858087f, -4.39150654e-06f,
                         -3.5233877e-06f, 3.43273939e-07f, 2.81022636e-08f };
    const float q[9] = { 2.83297682f, 1.00167406f, 0.00943887047f, -0.0076224613f, 0.00573950773f, -0.00367342844f,
                         0.00134934322f, 0.000100950558f, -0.000200214257f };
    return (w < 5.f ? estrin_(w1, p) : estrin_(w2, q)) * x;
}

struct Microfacet {
    int type; float alpha_u, alpha_v;     // type 0 = Beckmann, 1 = GGX
    bool sample_visible;                  // false: normals drawn from D * cos(theta_m) (microfacet.h:246-293, isotropic roughness)
    MIB_D Microfacet(int t, float au, float av, bool sv = true) : type(t), alpha_u(fmaxf(au, 1e-4f)), alpha_v(fmaxf(av, 1e-4f)), sample_visible(sv) {}

    MIB_D float eval(V3 m) const {
        float alpha_uv = alpha_u * alpha_v, cos_theta = m.z, cos_theta_2 = sqr_(cos_theta), result;
        if (type == 0)
            result = div_(exp_(div_(-(sqr_(div_(m.x, alpha_u)) + sqr_(div_(m.y, alpha_v))), cos_theta_2)), kPi * alpha_uv * sqr_(cos_theta_2));
        else
            result = rcp_(kPi * alpha_uv * sqr_(sqr_(div_(m.x, alpha_u)) + sqr_(div_(m.y, alpha_v)) + sqr_(m.z)));
        return result * cos_theta > 1e-20f ? result : 0.f;
    }

    MIB_D float smith_g1(V3 v, V3 m) const {
        float xy_alpha_2 = sqr_(alpha_u * v.x) + sqr_(alpha_v * v.y), tan_theta_alpha_2 = div_(xy_alpha_2, sqr_(v.z)), result;
        if (type == 0) {
            float a = rsqrt_(tan_theta_alpha_2), a_sqr = sqr_(a);
            result = a >= 1.6f ? 1.f : div_(3.535f * a + 2.181f * a_sqr, 1.f + 2.276f * a + 2.577f * a_sqr);
        } else {
            result = div_(2.f, 1.f + sqrt_(1.f + tan_theta_alpha_2));
        }
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }

    MIB_D float pdf(V3 wi, V3 m) const {
        if (!sample_visible) return eval(m) * m.z;
        return eval(m) * div_(smith_g1(wi, m) * fabsf(dot(wi, m)), wi.z);
    }

    MIB_D void sample_visible_11(float cos_theta_i, float sx, float sy, float &slope_x, float &slope_y) const {
        if (type == 0) {
            const float InvSqrtPi = 0.56418958354775628695f;
            float tan_theta_i = div_(safe_sqrt_(fnma_(cos_theta_i, cos_theta_i, 1.f)), cos_theta_i);
            float cot_theta_i = rcp_(tan_theta_i);
            float maxval = erf_(cot_theta_i);
            sx = fmaxf(fminf(sx, 1.f - 1e-6f), 1e-6f);
            sy = fmaxf(fminf(sy, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * erf_(sqrt_(-log_(sx)));
            sx *= 1.f + maxval + InvSqrtPi * tan_theta_i * exp_(-sqr_(cot_theta_i));
            for (int i = 0; i < 3; ++i) {
                float slope = erfinv_(x),
                      value = 1.f + x + InvSqrtPi * tan_theta_i * exp_(-sqr_(slope)) - sx,
                      derivative = 1.f - slope * tan_theta_i;
                x -= div_(value, derivative);
            }
            slope_x = erfinv_(x);
            slope_y = erfinv_(fms_(2.f, sy, 1.f));
        } else {
            float px, py;
            square_to_uniform_disk_concentric(sx, sy, px, py);
            float s = 0.5f * (1.f + cos_theta_i);
            float a = safe_sqrt_(1.f - sqr_(px));
            py = fma_(py, s, fnma_(a, s, a));
            float z = safe_sqrt_(1.f - fma_(py, py, sqr_(px)));
            float sin_theta_i = safe_sqrt_(1.f - sqr_(cos_theta_i));
            float nrm = rcp_(fma_(sin_theta_i, py, cos_theta_i * z));
            slope_x = fms_(cos_theta_i, py, sin_theta_i * z) * nrm;
            slope_y = px * nrm;
        }
    }

    // microfacet.h:244-326 (both sample_visible settings); returns the normal, pdf_out as the reference computes it
    MIB_D V3 sample(V3 wi, float sx, float sy, float &pdf_out) const {
        if (!sample_visible) {
            float sin_phi, cos_phi;
            sincos_((2.f * kPi) * sy, sin_phi, cos_phi);
            const float alpha_2 = alpha_u * alpha_u;
            float cos_theta, cos_theta_2;
            if (type == 0) {
                cos_theta = rsqrt_(fnma_(alpha_2, log_(1.f - sx), 1.f));
                cos_theta_2 = sqr_(cos_theta);
                const float cos_theta_3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdf_out = div_(1.f - sx, kPi * alpha_u * alpha_v * cos_theta_3);
            } else {
                const float tan_theta_m_2 = div_(alpha_2 * sx, 1.f - sx);
                cos_theta = rsqrt_(1.f + tan_theta_m_2);
                cos_theta_2 = sqr_(cos_theta);
                const float temp = 1.f + div_(tan_theta_m_2, alpha_2), cos_theta_3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdf_out = rcp_(kPi * alpha_u * alpha_v * cos_theta_3 * sqr_(temp));
            }
            const float sin_theta = sqrt_(1.f - cos_theta_2);
            return v3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
        }
        V3 wi_p = normalize(v3(alpha_u * wi.x, alpha_v * wi.y, wi.z));
        float sin_theta_2 = fma_(wi_p.x, wi_p.x, sqr_(wi_p.y)), inv_sin_theta = rsqrt_(sin_theta_2);
        float cos_phi = wi_p.x * inv_sin_theta, sin_phi = wi_p.y * inv_sin_theta;
        if (fabsf(sin_theta_2) <= 4.f * kEps) { cos_phi = 1.f; sin_phi = 0.f; }
        else { cos_phi = fmaxf(fminf(cos_phi, 1.f), -1.f); sin_phi = fmaxf(fminf(sin_phi, 1.f), -1.f); }
        float slx, sly;
        sample_visible_11(wi_p.z, sx, sy, slx, sly);
        float rx = fms_(cos_phi, slx, sin_phi * sly) * alpha_u, ry = fma_(sin_phi, slx, cos_phi * sly) * alpha_v;
        V3 m = normalize(v3(-rx, -ry, 1.f));
        pdf_out = div_(eval(m) * smith_g1(wi, m) * fabsf(dot(wi, m)), wi.z);
        return m;
    }
};

} // namespace mib

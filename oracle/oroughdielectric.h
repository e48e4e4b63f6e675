// oroughdielectric.h -- TEST INFRASTRUCTURE (CPU oracle).  RoughDielectric of the fork, src/bsdfs/roughdielectric.cpp:
// sample :248-411, eval :413-519, pdf :521-603, TransportMode::Radiance, both sample_visible settings, specular_reflectance /
// specular_transmittance = 1.  JIT forms throughout (fma-chain dot, rsqrt = sqrt(1/x)).  Built on the pinned blocks of
// ofresnel.h and omicrofacet.h.  Reference quirks kept on purpose:
//   * eval() uses G = 1 (:457), while sample() weights with smith_g1(wo, m) (:399-400): eval / pdf is therefore
//     1 / smith_g1(wi, m), not the sampling weight;
//   * the polarized sampling weight divides R by F and T by (1 - F) of the *unpolarized* Fresnel term evaluated at
//     dot(wi, m), while R, T are evaluated at dot(wo_hat, m) (:345-352).
// Nothing on the product path uses this yet.
#pragma once
#include "omicrofacet.h"

namespace orc {

enum : uint32_t { BSDF_GlossyReflection = 0x8, BSDF_GlossyTransmission = 0x10 };

struct RoughDielectric {
    int type; float alpha_u, alpha_v, eta; bool sample_visible = true;

    static Mueller zero_nan(Mueller M) { zero_nans(M); return M; }

    // Mueller weight of one lobe in the implicit Stokes bases of -wo_hat and wi_hat (:337-367, :466-490)
    static Mueller rotated(const Mueller &W, Vec3 m, Vec3 wo_hat, Vec3 wi_hat) {
        Vec3 s_axis_in = normalize(cross(m, -wo_hat)), s_axis_out = normalize(cross(m, wi_hat));
        return zero_nan(rotate_mueller_basis(W, -wo_hat, s_axis_in, stokes_basis(-wo_hat), wi_hat, s_axis_out, stokes_basis(wi_hat)));
    }

    struct Sample { Vec3 wo; float pdf = 0.f, eta = 1.f; uint32_t sampled_type = 0; float weight = 0.f; Mueller weight_pol; };

    Sample sample(Vec3 wi, float sample1, float s2x, float s2y, bool has_reflection, bool has_transmission) const {
        Sample bs;
        float cos_theta_i = wi.z;
        bool active = cos_theta_i != 0.f;
        Microfacet distr(type, alpha_u, alpha_v, sample_visible);
        Vec3 wi_up(mulsign_ad(wi.x, cos_theta_i), mulsign_ad(wi.y, cos_theta_i), mulsign_ad(wi.z, cos_theta_i));
        Vec3 m = distr.sample(wi_up, s2x, s2y, &bs.pdf);
        active = active && bs.pdf != 0.f;
        Fresnel f = fresnel(dot(wi, m), eta);
        float F = f.r;
        bool selected_r;
        if (has_reflection && has_transmission) { selected_r = sample1 <= F && active; bs.pdf *= selected_r ? F : 1.f - F; }
        else if (has_reflection || has_transmission) selected_r = has_reflection && active;
        else { bs.pdf = 0.f; return bs; }
        bool selected_t = !selected_r && active;
        bs.eta = selected_r ? 1.f : f.eta_it;
        bs.sampled_type = selected_r ? BSDF_GlossyReflection : BSDF_GlossyTransmission;
        if (selected_r) { float d2 = 2.f * dot(wi, m); bs.wo = Vec3(fmsub(m.x, d2, wi.x), fmsub(m.y, d2, wi.y), fmsub(m.z, d2, wi.z)); }
        else bs.wo = refract(wi, m, f.cos_theta_t, f.eta_ti);
        active = active && dot(wi, m) * wi.z > 0.f && dot(bs.wo, m) * bs.wo.z > 0.f;

        // polarized weight
        Vec3 wo_hat = bs.wo, wi_hat = wi;
        float cos_theta_o_hat = dot(wo_hat, m);
        Mueller R = specular_reflection(cos_theta_o_hat, eta), T = specular_transmission(cos_theta_o_hat, eta), W;
        if (has_reflection && has_transmission) {
            float div = selected_r ? F : 1.f - F;
            W = selected_r ? R : T;
            for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) W.m[i][j] = W.m[i][j] / div;
        } else W = has_reflection ? R : T;
        W = rotated(W, m, wo_hat, wi_hat);
        // unpolarized weight
        float w = (has_reflection && has_transmission) ? 1.f : (has_reflection ? F : 1.f - F);

        float dwh_dwo = 0.f;
        if (selected_r) dwh_dwo = rcp(4.f * dot(bs.wo, m));
        if (selected_t) {
            float factor = sqr(f.eta_ti);
            w *= factor; W = W * factor;
            dwh_dwo = (sqr(bs.eta) * dot(bs.wo, m)) / sqr(dot(wi, m) + bs.eta * dot(bs.wo, m));
        }
        // :402-408: the visible-normal weight, or (the fork's G = 1 form of) the weight of sampling D cos(theta_m)
        float g1 = sample_visible ? distr.smith_g1(bs.wo, m) : dot(wi, m) / (cos_theta_i * m.z);
        w *= g1; W = W * g1;
        bs.pdf *= std::fabs(dwh_dwo);
        if (!active) { w = 0.f; W = Mueller(); }
        bs.weight = w; bs.weight_pol = W;
        return bs;
    }

    // half-vector of (wi, wo) on the side of the macro normal (:431-436, :536-537)
    Vec3 half_vector(Vec3 wi, Vec3 wo, bool reflect, float eta_side) const {
        float k = reflect ? 1.f : eta_side;
        Vec3 m = normalize(Vec3(wi.x + wo.x * k, wi.y + wo.y * k, wi.z + wo.z * k));
        return Vec3(mulsign_ad(m.x, m.z), mulsign_ad(m.y, m.z), mulsign_ad(m.z, m.z));
    }

    // eval(): *value = unpolarized BSDF * |cos theta_o|, *value_pol = its Mueller form
    void eval(Vec3 wi, Vec3 wo, bool has_reflection, bool has_transmission, float *value, Mueller *value_pol) const {
        *value = 0.f; *value_pol = Mueller();
        float cos_theta_i = wi.z, cos_theta_o = wo.z;
        bool active = cos_theta_i != 0.f;
        bool reflect = cos_theta_i * cos_theta_o > 0.f;
        float inv_eta_c = rcp(eta);
        float eta_s = cos_theta_i > 0.f ? eta : inv_eta_c, inv_eta_s = cos_theta_i > 0.f ? inv_eta_c : eta;
        Vec3 m = half_vector(wi, wo, reflect, eta_s);
        active = active && dot(wi, m) * wi.z > 0.f && dot(wo, m) * wo.z > 0.f;
        Microfacet distr(type, alpha_u, alpha_v, sample_visible);
        float D = distr.eval(m), F = fresnel(dot(wi, m), eta).r, G = 1.f;
        bool eval_r = has_reflection && reflect && active, eval_t = has_transmission && !reflect && active;
        if (!eval_r && !eval_t) return;
        float cos_theta_o_hat = dot(wo, m);
        Mueller W = rotated(eval_r ? specular_reflection(cos_theta_o_hat, eta) : specular_transmission(cos_theta_o_hat, eta), m, wo, wi);
        float w = eval_r ? F : 1.f - F;
        if (eval_r) {
            float k = D * G / (4.f * std::fabs(cos_theta_i));
            *value = w * D * G / (4.f * std::fabs(cos_theta_i));
            *value_pol = W * k;
        } else {
            float scale = sqr(inv_eta_s);
            float k = (scale * D * G * eta_s * eta_s * dot(wi, m) * dot(wo, m)) / (cos_theta_i * sqr(dot(wi, m) + eta_s * dot(wo, m)));
            *value = std::fabs(w * k);
            Mueller V = W * k;
            for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) V.m[i][j] = std::fabs(V.m[i][j]);     // dr::abs on the matrix (:507-509)
            *value_pol = V;
        }
    }

    float pdf(Vec3 wi, Vec3 wo, bool has_reflection, bool has_transmission) const {
        float cos_theta_i = wi.z, cos_theta_o = wo.z;
        bool active = cos_theta_i != 0.f;
        bool reflect = cos_theta_i * cos_theta_o > 0.f;
        active = active && ((has_reflection && reflect) || (has_transmission && !reflect));
        float eta_s = cos_theta_i > 0.f ? eta : rcp(eta);
        Vec3 m = half_vector(wi, wo, reflect, eta_s);
        active = active && dot(wi, m) * wi.z > 0.f && dot(wo, m) * wo.z > 0.f;
        float dwh_dwo = reflect ? rcp(4.f * dot(wo, m)) : (eta_s * eta_s * dot(wo, m)) / sqr(dot(wi, m) + eta_s * dot(wo, m));
        Microfacet distr(type, alpha_u, alpha_v, sample_visible);
        Vec3 wi_up(mulsign_ad(wi.x, cos_theta_i), mulsign_ad(wi.y, cos_theta_i), mulsign_ad(wi.z, cos_theta_i));
        float prob = distr.pdf(wi_up, m);
        if (has_transmission && has_reflection) {
            float F = fresnel(dot(wi, m), eta).r;
            prob *= reflect ? F : 1.f - F;
        }
        return active ? prob * std::fabs(dwh_dwo) : 0.f;
    }
};

} // namespace orc

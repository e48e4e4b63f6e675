// obsdf.h -- TEST INFRASTRUCTURE (CPU oracle).  BSDFs of the path:
//   null             src/bsdfs/null.cpp:36-86
//   diffuse          src/bsdfs/diffuse.cpp:101-160
//   dielectric       src/bsdfs/dielectric.cpp:247-468      (ofresnel.h dielectric_sample / dielectric_eval; pdf is zero)
//   roughdielectric  src/bsdfs/roughdielectric.cpp:248-603 (oroughdielectric.h; the fork's G = 1 in eval)
//   thindielectric   src/bsdfs/thindielectric.cpp:137-224  (reflection lobe + a NULL transmission lobe of weight 1 - r')
//   plastic          src/bsdfs/plastic.cpp:202-352         (delta reflection + diffuse base with internal scattering)
//   roughplastic     src/bsdfs/roughplastic.cpp:256-515    (microfacet reflection + diffuse base; tables filled by the host)
// The two refractive ones are the ocean interface of the fork's atmosphere-ocean scenes (volpathaos.cpp:232-290,312-337).
// ctx_mask: bit 0 = reflection lobes enabled, bit 1 = transmission lobes enabled (BSDFContext::type_mask; volpathaos'
// refractive NEE asks for transmission only).
#pragma once
#include "oroughdielectric.h"

namespace orc {

enum : uint32_t { BSDF_Null = 0x1, BSDF_DiffuseReflection = 0x2, BSDF_Smooth = 0x2 | 0x4 | 0x8 | 0x10,
                  BSDF_Delta = 0x1 | 0x20 | 0x40, BSDF_Transmission = 0x10 | 0x40 };
static inline uint32_t bsdf_flags(const mib_bsdf &b) {
    switch (b.type) {
        case MIB_BSDF_NULL: return BSDF_Null;
        case MIB_BSDF_DIELECTRIC: return BSDF_DeltaReflection | BSDF_DeltaTransmission;
        case MIB_BSDF_ROUGHDIELECTRIC: return BSDF_GlossyReflection | BSDF_GlossyTransmission;
        case MIB_BSDF_THINDIELECTRIC: return BSDF_DeltaReflection | BSDF_Null;
        case MIB_BSDF_PLASTIC: return BSDF_DeltaReflection | BSDF_DiffuseReflection;
        case MIB_BSDF_ROUGHPLASTIC: return BSDF_GlossyReflection | BSDF_DiffuseReflection;
        default: return BSDF_DiffuseReflection;
    }
}

template <typename Spec> struct BSDFSample { Vec3 wo; float pdf = 0.f, eta = 1.f; uint32_t sampled_type = 0; Spec value; };

template <int SPEC> struct BSDF {
    using Tr = SpecTraits<SPEC>;
    using Spec = typename Tr::Spec;
    using USpec = typename Tr::USpec;
    // diffuse reflectance: a constant, or the checkerboard texture (src/textures/checkerboard.cpp:66-84) at si.uv
    static USpec reflectance(const mib_bsdf &b, const SurfaceInteraction &si) {
        const float *src = b.reflectance;
        if (b.texture == MIB_TEXTURE_CHECKERBOARD) {
            // Transform3f::transform_affine(Point2f): column 2, then fmadd of columns 0 and 1
            float u = fmadd(b.to_uv[1], si.uv[1], fmadd(b.to_uv[0], si.uv[0], b.to_uv[2])),
                  v = fmadd(b.to_uv[4], si.uv[1], fmadd(b.to_uv[3], si.uv[0], b.to_uv[5]));
            bool mx = u - std::floor(u) > .5f, my = v - std::floor(v) > .5f;
            src = mx == my ? b.color0 : b.color1;
        }
        USpec r; for (int c = 0; c < Tr::NC; ++c) r.v[c] = src[Tr::NC == 3 ? c : 0]; return r;
    }
    static Spec pick(float w, const Mueller &W) { if constexpr (Tr::POL) return W; else { (void) W; return Spec(w); } }
    static RoughDielectric rough(const mib_bsdf &b) { return RoughDielectric{ b.distribution, b.alpha_u, b.alpha_v, b.eta, b.sample_visible != 0 }; }

    static USpec uspec3(const float *src) { USpec r; for (int c = 0; c < Tr::NC; ++c) r.v[c] = src[Tr::NC == 3 ? c : 0]; return r; }
    // roughplastic.cpp:392-402
    static float lerp_gather(const float *data, float x) {
        x *= (float) (MIB_ROUGH_TRANSMITTANCE_RES - 1);
        uint32_t index = std::min((uint32_t) x, (uint32_t) (MIB_ROUGH_TRANSMITTANCE_RES - 2));
        float v0 = data[index], v1 = data[index + 1], t = x - (float) index;
        return fmadd(v1, t, fnmadd(v0, t, v0));                                      // dr::lerp
    }
    // diffuse base seen through the interface: diff / (1 - internal reflections) (plastic.cpp:255-256,285-286; roughplastic.cpp:382-385)
    static USpec plastic_base(const mib_bsdf &b) {
        USpec diff = uspec3(b.reflectance);
        USpec den = b.nonlinear ? USpec(1.f) - diff * b.internal_reflectance : USpec(1.f - b.internal_reflectance);
        return diff / den;
    }
    static float rough_prob_specular(const mib_bsdf &b, float cos_theta_i) {
        float t_i = lerp_gather(b.ext_transmittance, cos_theta_i);
        float prob_specular = (1.f - t_i) * b.specular_sampling_weight, prob_diffuse = t_i * (1.f - b.specular_sampling_weight);
        return prob_specular / (prob_specular + prob_diffuse);
    }
    static USpec roughplastic_eval(const mib_bsdf &b, const SurfaceInteraction &si, Vec3 wo) {
        float cos_theta_i = si.wi.z, cos_theta_o = wo.z;
        if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return USpec(0.f);
        Microfacet distr(b.distribution, b.alpha_u, b.alpha_v);
        Vec3 H = md_normalize(wo + si.wi);
        float D = distr.eval(H), F = fresnel(md_dot(si.wi, H), b.eta).r, G = distr.G(si.wi, wo, H);
        USpec value(F * D * G / (4.f * cos_theta_i));
        if (b.has_specular_reflectance) value = value * uspec3(b.specular_reflectance);
        float t_i = lerp_gather(b.ext_transmittance, cos_theta_i), t_o = lerp_gather(b.ext_transmittance, cos_theta_o);
        return value + plastic_base(b) * (InvPi * b.inv_eta_2 * cos_theta_o * t_i * t_o);
    }
    static float roughplastic_pdf(const mib_bsdf &b, const SurfaceInteraction &si, Vec3 wo) {
        float cos_theta_i = si.wi.z, cos_theta_o = wo.z;
        if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return 0.f;
        float prob_specular = rough_prob_specular(b, cos_theta_i), prob_diffuse = 1.f - prob_specular;
        Microfacet distr(b.distribution, b.alpha_u, b.alpha_v);
        Vec3 H = md_normalize(wo + si.wi);
        float result = distr.eval(H) * distr.smith_g1(si.wi, H) / (4.f * cos_theta_i);
        result *= prob_specular;
        result += prob_diffuse * (InvPi * wo.z);
        return result;
    }
    static USpec plastic_eval(const mib_bsdf &b, const SurfaceInteraction &si, Vec3 wo) {
        float cos_theta_i = si.wi.z, cos_theta_o = wo.z;
        if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return USpec(0.f);
        float f_i = fresnel(cos_theta_i, b.eta).r, f_o = fresnel(cos_theta_o, b.eta).r;
        return plastic_base(b) * ((InvPi * cos_theta_o) * b.inv_eta_2 * (1.f - f_i) * (1.f - f_o));
    }
    static float plastic_pdf(const mib_bsdf &b, const SurfaceInteraction &si, Vec3 wo) {
        float cos_theta_i = si.wi.z, cos_theta_o = wo.z;
        if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return 0.f;
        float f_i = fresnel(cos_theta_i, b.eta).r, prob_specular = f_i * b.specular_sampling_weight;
        float prob_diffuse = (1.f - f_i) * (1.f - b.specular_sampling_weight);
        prob_diffuse = prob_diffuse / (prob_specular + prob_diffuse);
        return (InvPi * cos_theta_o) * prob_diffuse;
    }

    static BSDFSample<Spec> sample(const mib_bsdf &b, const SurfaceInteraction &si, float s1, float s2x, float s2y, int ctx_mask = 3) {
        BSDFSample<Spec> bs;
        bs.value = Spec(0.f);
        if constexpr (Tr::POL) bs.value = Mueller();
        if (b.type == MIB_BSDF_PLASTIC || b.type == MIB_BSDF_ROUGHPLASTIC) {
            // plastic.cpp:202-264, roughplastic.cpp:256-311; both lobes are reflections (ctx bit 0).  A hit on the back side
            // returns the zeroed sample (pdf 0, value 0): in JIT variants pdf() is then evaluated for wo = 0 and not masked, a
            // value nothing observable depends on (the sample's weight is 0 either way).
            const float cos_theta_i = si.wi.z;
            if (!(ctx_mask & 1) || !(cos_theta_i > 0.f)) return bs;
            const bool rough = b.type == MIB_BSDF_ROUGHPLASTIC;
            float f_i = 0.f, prob_specular;
            if (rough) prob_specular = rough_prob_specular(b, cos_theta_i);
            else {
                f_i = fresnel(cos_theta_i, b.eta).r;
                prob_specular = f_i * b.specular_sampling_weight;
                float prob_diffuse = (1.f - f_i) * (1.f - b.specular_sampling_weight);
                prob_specular = prob_specular / (prob_specular + prob_diffuse);
            }
            const float prob_diffuse = 1.f - prob_specular;
            const bool sample_specular = s1 < prob_specular;
            bs.eta = 1.f;
            USpec value(0.f);
            if (rough) {
                if (sample_specular) {
                    Microfacet distr(b.distribution, b.alpha_u, b.alpha_v);
                    float mpdf;
                    Vec3 m = distr.sample(si.wi, s2x, s2y, &mpdf);
                    float d2 = 2.f * md_dot(si.wi, m);
                    bs.wo = Vec3(fmsub(m.x, d2, si.wi.x), fmsub(m.y, d2, si.wi.y), fmsub(m.z, d2, si.wi.z));
                    bs.sampled_type = BSDF_GlossyReflection;
                } else {
                    bs.wo = square_to_cosine_hemisphere(s2x, s2y);
                    bs.sampled_type = BSDF_DiffuseReflection;
                }
                bs.pdf = roughplastic_pdf(b, si, bs.wo);
                if (bs.pdf > 0.f) value = div_scalar(roughplastic_eval(b, si, bs.wo), bs.pdf);
            } else if (sample_specular) {
                bs.wo = reflect(si.wi);
                bs.pdf = prob_specular;
                bs.sampled_type = BSDF_DeltaReflection;
                value = USpec(f_i / bs.pdf);
                if (b.has_specular_reflectance) value = value * uspec3(b.specular_reflectance);
            } else {
                bs.wo = square_to_cosine_hemisphere(s2x, s2y);
                bs.pdf = prob_diffuse * (InvPi * bs.wo.z);
                bs.sampled_type = BSDF_DiffuseReflection;
                float f_o = fresnel(bs.wo.z, b.eta).r;
                value = plastic_base(b) * (b.inv_eta_2 * (1.f - f_i) * (1.f - f_o) / prob_diffuse);
            }
            bs.value = Depol<Spec, Tr::NC>::make(value);
            return bs;
        }
        if (b.type == MIB_BSDF_NULL) {
            bs.wo = -si.wi; bs.sampled_type = BSDF_Null; bs.eta = 1.f; bs.pdf = 1.f;
            bs.value = Spec(1.f); // identity in polarized modes (null.cpp:57-63)
            return bs;
        }
        if (b.type == MIB_BSDF_THINDIELECTRIC) {
            // thindielectric.cpp:137-190; r' = r + trt + tr^3t + ... accounts for the internal reflections
            bool has_r = (ctx_mask & 1) != 0, has_t = (ctx_mask & 2) != 0;       // (the transmission lobe is BSDFFlags::Null)
            float r = fresnel(std::fabs(si.wi.z), b.eta).r;
            r *= 2.f / (1.f + r);
            float t = 1.f - r, weight;
            bool selected_r;
            if (has_r && has_t) { selected_r = s1 <= r; weight = 1.f; bs.pdf = selected_r ? r : t; }
            else if (has_r || has_t) { selected_r = has_r; weight = has_r ? r : t; bs.pdf = 1.f; }
            else return bs;
            bs.wo = selected_r ? reflect(si.wi) : -si.wi;
            bs.eta = 1.f;
            bs.sampled_type = selected_r ? (uint32_t) BSDF_DeltaReflection : (uint32_t) BSDF_Null;
            bs.value = Depol<Spec, Tr::NC>::make(USpec(weight));
            return bs;
        }
        if (b.type == MIB_BSDF_DIELECTRIC) {
            DielectricSample d = dielectric_sample(si.wi, b.eta, s1, (ctx_mask & 1) != 0, (ctx_mask & 2) != 0);
            bs.wo = d.wo; bs.pdf = d.pdf; bs.eta = d.eta; bs.sampled_type = d.sampled_type; bs.value = pick(d.weight, d.weight_pol);
            return bs;
        }
        if (b.type == MIB_BSDF_ROUGHDIELECTRIC) {
            RoughDielectric::Sample d = rough(b).sample(si.wi, s1, s2x, s2y, (ctx_mask & 1) != 0, (ctx_mask & 2) != 0);
            bs.wo = d.wo; bs.pdf = d.pdf; bs.eta = d.eta; bs.sampled_type = d.sampled_type; bs.value = pick(d.weight, d.weight_pol);
            return bs;
        }
        // diffuse.cpp:108-111: nothing is sampled when the context disables the diffuse reflection lobe -- volpathaos asks any
        // BSDF it took for "the refractive one" for its transmission lobes only (volpathaos.cpp:262-285,318-326: a diffuse
        // bounce off a surface that is not horizontal counts as a refraction there)
        if (!(ctx_mask & 1)) return bs;
        // JIT variants have no early exit: all fields are written for every lane and
        // only the value is masked by (cos_theta_i > 0 && pdf > 0) (diffuse.cpp:101-126)
        float cos_theta_i = si.wi.z;
        if (g_scalar_shapes && !(cos_theta_i > 0.f)) return bs;      // scalar variants' early exit (test-only switch, oscene.h)
        bs.wo = square_to_cosine_hemisphere(s2x, s2y);
        bs.pdf = InvPi * bs.wo.z;
        bs.eta = 1.f;
        bs.sampled_type = BSDF_DiffuseReflection;
        bs.value = Depol<Spec, Tr::NC>::make((cos_theta_i > 0.f && bs.pdf > 0.f) ? reflectance(b, si) : USpec(0.f));
        return bs;
    }
    static Spec eval(const mib_bsdf &b, const SurfaceInteraction &si, Vec3 wo) {
        Spec zero = Depol<Spec, Tr::NC>::make(USpec(0.f));
        if (b.type == MIB_BSDF_NULL || b.type == MIB_BSDF_THINDIELECTRIC) return zero;
        if (b.type == MIB_BSDF_DIELECTRIC) {
            float v; Mueller V;
            dielectric_eval(si.wi, wo, b.eta, b.inv_eta, true, true, &v, &V);
            return pick(v, V);
        }
        if (b.type == MIB_BSDF_ROUGHDIELECTRIC) {
            float v; Mueller V;
            rough(b).eval(si.wi, wo, true, true, &v, &V);
            return pick(v, V);
        }
        if (b.type == MIB_BSDF_PLASTIC) return Depol<Spec, Tr::NC>::make(plastic_eval(b, si, wo));
        if (b.type == MIB_BSDF_ROUGHPLASTIC) return Depol<Spec, Tr::NC>::make(roughplastic_eval(b, si, wo));
        float cos_theta_i = si.wi.z, cos_theta_o = wo.z;
        if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return zero;
        USpec value = reflectance(b, si) * InvPi * cos_theta_o;
        return Depol<Spec, Tr::NC>::make(value);
    }
    static float pdf(const mib_bsdf &b, const SurfaceInteraction &si, Vec3 wo) {
        if (b.type == MIB_BSDF_NULL || b.type == MIB_BSDF_DIELECTRIC || b.type == MIB_BSDF_THINDIELECTRIC) return 0.f;
        if (b.type == MIB_BSDF_ROUGHDIELECTRIC) return rough(b).pdf(si.wi, wo, true, true);
        if (b.type == MIB_BSDF_PLASTIC) return plastic_pdf(b, si, wo);
        if (b.type == MIB_BSDF_ROUGHPLASTIC) return roughplastic_pdf(b, si, wo);
        float cos_theta_i = si.wi.z, cos_theta_o = wo.z;
        return (cos_theta_i > 0.f && cos_theta_o > 0.f) ? InvPi * wo.z : 0.f;
    }
    static Spec eval_null_transmission(const mib_bsdf &b, const SurfaceInteraction &si) {
        if (b.type == MIB_BSDF_NULL) return Spec(1.f);
        if (b.type == MIB_BSDF_THINDIELECTRIC) {                     // thindielectric.cpp:202-215
            float r = fresnel(std::fabs(si.wi.z), b.eta).r;
            r *= 2.f / (1.f + r);
            return Depol<Spec, Tr::NC>::make(USpec(1.f - r));
        }
        return Depol<Spec, Tr::NC>::make(USpec(0.f)); // bsdf.h default: 0
    }
};

} // namespace orc

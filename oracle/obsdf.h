// obsdf.h -- TEST INFRASTRUCTURE (CPU oracle).  BSDFs of the path:
//   null             src/bsdfs/null.cpp:36-86
//   diffuse          src/bsdfs/diffuse.cpp:101-160
//   dielectric       src/bsdfs/dielectric.cpp:247-468      (ofresnel.h dielectric_sample / dielectric_eval; pdf is zero)
//   roughdielectric  src/bsdfs/roughdielectric.cpp:248-603 (oroughdielectric.h; the fork's G = 1 in eval)
//   thindielectric   src/bsdfs/thindielectric.cpp:137-224  (reflection lobe + a NULL transmission lobe of weight 1 - r')
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
    static RoughDielectric rough(const mib_bsdf &b) { return RoughDielectric{ b.distribution, b.alpha_u, b.alpha_v, b.eta }; }

    static BSDFSample<Spec> sample(const mib_bsdf &b, const SurfaceInteraction &si, float s1, float s2x, float s2y, int ctx_mask = 3) {
        BSDFSample<Spec> bs;
        bs.value = Spec(0.f);
        if constexpr (Tr::POL) bs.value = Mueller();
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
        float cos_theta_i = si.wi.z, cos_theta_o = wo.z;
        if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return zero;
        USpec value = reflectance(b, si) * InvPi * cos_theta_o;
        return Depol<Spec, Tr::NC>::make(value);
    }
    static float pdf(const mib_bsdf &b, const SurfaceInteraction &si, Vec3 wo) {
        if (b.type == MIB_BSDF_NULL || b.type == MIB_BSDF_DIELECTRIC || b.type == MIB_BSDF_THINDIELECTRIC) return 0.f;
        if (b.type == MIB_BSDF_ROUGHDIELECTRIC) return rough(b).pdf(si.wi, wo, true, true);
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

// dbsdf.cuh -- BSDFs of the path on the device beyond null / diffuse:
//   SmoothDielectric   src/bsdfs/dielectric.cpp:247-468         (sample; the fork's non-zero eval; pdf = 0)
//   RoughDielectric    src/bsdfs/roughdielectric.cpp:248-603    (sample / eval with G = 1 / pdf)
//   SmoothPlastic      src/bsdfs/plastic.cpp:202-352            (delta reflection + diffuse base with internal scattering)
//   RoughPlastic       src/bsdfs/roughplastic.cpp:256-515       (microfacet reflection + diffuse base; the transmittance table
//                                                                and internal reflectance come from the host, plastic.py)
//   diffuse reflectance through a checkerboard texture          src/textures/checkerboard.cpp:66-84
// and the dispatch over all BSDF kinds (bsdf_sample / bsdf_eval / bsdf_pdf) in the layout of the throughput.  The reference's
// quirks are kept (SmoothDielectric::eval without Fresnel factor, eval with G = 1, the polarized sampling weight's mixed
// angles); oracle/obsdf.h, oracle/oroughdielectric.h and oracle/ofresnel.h list them.  Used by the one-thread-one-path
// kernels (dwalk.cuh volpath, docean.cuh volpathaos).
#pragma once
#include "dplugins.cuh"
#include "dfresnel.cuh"
#include "dmicrofacet.cuh"

namespace mib {

// Diffuse reflectance of channel c: a constant, or the checkerboard texture at the hit's uv (src/textures/checkerboard.cpp:66-84)
MIB_D float bsdf_reflectance(const mib_bsdf &b, const SurfIt &si, int c) {
    if (MIB_FT(FT_MESH) && b.texture == MIB_TEXTURE_CHECKERBOARD) {
        const float u = fma_(b.to_uv[1], si.tv, fma_(b.to_uv[0], si.tu, b.to_uv[2])),
                    v = fma_(b.to_uv[4], si.tv, fma_(b.to_uv[3], si.tu, b.to_uv[5]));
        const bool mx = u - floorf(u) > .5f, my = v - floorf(v) > .5f;
        return mx == my ? b.color0[c] : b.color1[c];
    }
    return b.reflectance[c];
}

constexpr uint32_t kBsdfGlossyRefl = 0x8, kBsdfGlossyTrans = 0x10, kBsdfDeltaRefl = 0x20, kBsdfDeltaTrans = 0x40,
                   kBsdfTransmission = 0x10 | 0x40;

MIB_D uint32_t bsdf_flags(const mib_bsdf &b) {
    switch (b.type) {
        case MIB_BSDF_NULL: return kBsdfNull;
        case MIB_BSDF_DIELECTRIC: return kBsdfDeltaRefl | kBsdfDeltaTrans;
        case MIB_BSDF_ROUGHDIELECTRIC: return kBsdfGlossyRefl | kBsdfGlossyTrans;
        case MIB_BSDF_THINDIELECTRIC: return kBsdfDeltaRefl | kBsdfNull;
        case MIB_BSDF_PLASTIC: return kBsdfDeltaRefl | kBsdfDiffuseRefl;
        case MIB_BSDF_ROUGHPLASTIC: return kBsdfGlossyRefl | kBsdfDiffuseRefl;
        default: return kBsdfDiffuseRefl;
    }
}

struct BsdfS { V3 wo; float pdf, eta; uint32_t sampled; };

MIB_D void m4_zero(float *M) {
#pragma unroll
    for (int i = 0; i < 16; ++i) M[i] = 0.f;
}
// [[a, b, 0, 0], [b, a, 0, 0], [0, 0, c, -s], [0, 0, s, c]] from (a, b, c, s) (dfresnel.cuh)
MIB_D void m4_from_block(float4 k, float *M) {
    m4_zero(M);
    M[0] = k.x; M[1] = k.y; M[4] = k.y; M[5] = k.x;
    M[10] = k.z; M[11] = -k.w; M[14] = k.w; M[15] = k.z;
}
MIB_D void m4_scale(float *M, float f) {
#pragma unroll
    for (int i = 0; i < 16; ++i) M[i] *= f;
}
MIB_D void m4_zero_nans(float *M) {
#pragma unroll
    for (int i = 0; i < 16; ++i) if (M[i] != M[i]) M[i] = 0.f;
}
// rotate_mueller_basis (mueller.h:421-434): R_out * M * R_in^T
MIB_D void m4_rotate_basis(float *M, V3 in_f, V3 in_cur, V3 in_tgt, V3 out_f, V3 out_cur, V3 out_tgt) {
    float ci, s_i, co, so;
    stokes_rotation(in_f, in_cur, in_tgt, ci, s_i);
    stokes_rotation(out_f, out_cur, out_tgt, co, so);
    m4_rot_left(M, co, so);
    m4_rot_right_t(M, ci, s_i);
}

// ---- SmoothDielectric ---------------------------------------------------------------------------------------------------
// sample (dielectric.cpp:247-371): w = unpolarized weight, W = Mueller weight (polarized variants only)
template <bool POL>
MIB_D void dielectric_sample(V3 wi, float eta, float s1, bool has_r, bool has_t, BsdfS &bs, float &w, float *W) {
    bs.wo = v3(0.f); bs.pdf = 0.f; bs.eta = 1.f; bs.sampled = 0; w = 0.f;
    if constexpr (POL) m4_zero(W);
    const float4 f = fresnel_(wi.z, eta);                   // (F, cos_theta_t, eta_it, eta_ti)
    const float r_i = f.x, t_i = 1.f - r_i;
    bool sel_r;
    if (has_r && has_t) { sel_r = s1 <= r_i; bs.pdf = sel_r ? r_i : t_i; }
    else if (has_r || has_t) { sel_r = has_r; bs.pdf = 1.f; }
    else return;
    bs.sampled = sel_r ? kBsdfDeltaRefl : kBsdfDeltaTrans;
    bs.wo = sel_r ? v3(-wi.x, -wi.y, wi.z) : v3(-f.w * wi.x, -f.w * wi.y, f.y);
    bs.eta = sel_r ? 1.f : f.z;
    if constexpr (POL) {
        const V3 wo_hat = bs.wo, wi_hat = wi;
        m4_from_block(sel_r ? specular_reflection_(wo_hat.z, eta) : specular_transmission_(wo_hat.z, eta), W);
        if (has_r && has_t) {
#pragma unroll
            for (int i = 0; i < 16; ++i) W[i] = div_(W[i], bs.pdf);
        }
        const V3 n = v3(0.f, 0.f, 1.f);
        V3 s_in = cross(n, -wo_hat), s_out = cross(n, wi_hat);
        const bool collinear = s_in.x == 0.f && s_in.y == 0.f && s_in.z == 0.f;
        s_in = collinear ? v3(1.f, 0.f, 0.f) : normalize(s_in);
        s_out = collinear ? v3(1.f, 0.f, 0.f) : normalize(s_out);
        m4_rotate_basis(W, -wo_hat, s_in, stokes_basis(-wo_hat), wi_hat, s_out, stokes_basis(wi_hat));
    }
    w = (has_r && has_t) ? 1.f : (has_r ? r_i : t_i);
    if (!sel_r) {
        const float fac = sqr_(f.w);
        w *= fac;
        if constexpr (POL) m4_scale(W, fac);
    }
}
// eval (dielectric.cpp:374-463, the fork's)
template <bool POL>
MIB_D void dielectric_eval(V3 wi, V3 wo, float eta, float inv_eta, float &w, float *W) {
    const float cos_theta_i = wi.z, cos_theta_o = wo.z;
    const bool active = cos_theta_i != 0.f, reflect = cos_theta_i * cos_theta_o > 0.f;
    const float inv_eta_side = cos_theta_i > 0.f ? inv_eta : eta;
    const bool eval_r = reflect && active, eval_t = !reflect && active;
    w = 1.f;
    if constexpr (POL) {
        const V3 wo_hat = wo, wi_hat = wi;
        m4_zero(W);
        if (eval_r) m4_from_block(specular_reflection_(wo_hat.z, eta), W);
        if (eval_t) m4_from_block(specular_transmission_(wo_hat.z, eta), W);
        const V3 n = v3(0.f, 0.f, 1.f);
        const V3 s_in = normalize(cross(n, -wo_hat)), s_out = normalize(cross(n, wi_hat));
        m4_rotate_basis(W, -wo_hat, s_in, stokes_basis(-wo_hat), wi_hat, s_out, stokes_basis(wi_hat));
    }
    if (eval_t) {
        const float fac = sqr_(inv_eta_side);
        w *= fac;
        if constexpr (POL) m4_scale(W, fac);
    }
}

// ---- RoughDielectric ----------------------------------------------------------------------------------------------------
MIB_D void rough_rotated(float *W, V3 m, V3 wo_hat, V3 wi_hat) {           // roughdielectric.cpp:337-367, :466-490
    const V3 s_in = normalize(cross(m, -wo_hat)), s_out = normalize(cross(m, wi_hat));
    m4_rotate_basis(W, -wo_hat, s_in, stokes_basis(-wo_hat), wi_hat, s_out, stokes_basis(wi_hat));
    m4_zero_nans(W);
}
MIB_D V3 rough_half_vector(V3 wi, V3 wo, bool reflect, float eta_side) {    // :431-436, :536-537
    const float k = reflect ? 1.f : eta_side;
    const V3 m = normalize(v3(wi.x + wo.x * k, wi.y + wo.y * k, wi.z + wo.z * k));
    return v3(mulsign_ad(m.x, m.z), mulsign_ad(m.y, m.z), mulsign_ad(m.z, m.z));
}
// sample (roughdielectric.cpp:248-411)
template <bool POL>
MIB_D void rough_sample(const mib_bsdf &b, V3 wi, float s1, float s2x, float s2y, bool has_r, bool has_t, BsdfS &bs, float &w, float *W) {
    bs.wo = v3(0.f); bs.pdf = 0.f; bs.eta = 1.f; bs.sampled = 0; w = 0.f;
    if constexpr (POL) m4_zero(W);
    const float eta = b.eta, cos_theta_i = wi.z;
    bool active = cos_theta_i != 0.f;
    const Microfacet distr(b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
    const V3 wi_up = v3(mulsign_ad(wi.x, cos_theta_i), mulsign_ad(wi.y, cos_theta_i), mulsign_ad(wi.z, cos_theta_i));
    const V3 m = distr.sample(wi_up, s2x, s2y, bs.pdf);
    active = active && bs.pdf != 0.f;
    const float wi_m = dot(wi, m);
    const float4 f = fresnel_(wi_m, eta);
    const float F = f.x;
    bool sel_r;
    if (has_r && has_t) { sel_r = s1 <= F && active; bs.pdf *= sel_r ? F : 1.f - F; }
    else if (has_r || has_t) sel_r = has_r && active;
    else { bs.pdf = 0.f; return; }
    const bool sel_t = !sel_r && active;
    bs.eta = sel_r ? 1.f : f.z;
    bs.sampled = sel_r ? kBsdfGlossyRefl : kBsdfGlossyTrans;
    if (sel_r) { const float d2 = 2.f * wi_m; bs.wo = v3(fms_(m.x, d2, wi.x), fms_(m.y, d2, wi.y), fms_(m.z, d2, wi.z)); }
    else {
        const float3 r = refract_(make_float3(wi.x, wi.y, wi.z), make_float3(m.x, m.y, m.z), f.y, f.w);
        bs.wo = v3(r.x, r.y, r.z);
    }
    const float wo_m = dot(bs.wo, m);
    active = active && wi_m * wi.z > 0.f && wo_m * bs.wo.z > 0.f;
    if constexpr (POL) {
        const float4 blk = ((has_r && has_t) ? sel_r : has_r) ? specular_reflection_(wo_m, eta) : specular_transmission_(wo_m, eta);
        m4_from_block(blk, W);
        if (has_r && has_t) {
            const float dv = sel_r ? F : 1.f - F;
#pragma unroll
            for (int i = 0; i < 16; ++i) W[i] = div_(W[i], dv);
        }
        rough_rotated(W, m, bs.wo, wi);
    }
    w = (has_r && has_t) ? 1.f : (has_r ? F : 1.f - F);
    float dwh_dwo = 0.f;
    if (sel_r) dwh_dwo = rcp_(4.f * wo_m);
    if (sel_t) {
        const float factor = sqr_(f.w);
        w *= factor;
        if constexpr (POL) m4_scale(W, factor);
        dwh_dwo = div_(sqr_(bs.eta) * wo_m, sqr_(wi_m + bs.eta * wo_m));
    }
    // roughdielectric.cpp:402-408: the visible-normal weight, or (the fork's G = 1 form of) the weight of sampling D cos(theta_m)
    const float g1 = distr.sample_visible ? distr.smith_g1(bs.wo, m) : div_(wi_m, cos_theta_i * m.z);
    w *= g1;
    if constexpr (POL) m4_scale(W, g1);
    bs.pdf *= fabsf(dwh_dwo);
    if (!active) { w = 0.f; if constexpr (POL) m4_zero(W); }
}
// eval (roughdielectric.cpp:413-519)
template <bool POL>
MIB_D void rough_eval(const mib_bsdf &b, V3 wi, V3 wo, float &w_out, float *W) {
    w_out = 0.f;
    if constexpr (POL) m4_zero(W);
    const float eta = b.eta, cos_theta_i = wi.z, cos_theta_o = wo.z;
    bool active = cos_theta_i != 0.f;
    const bool reflect = cos_theta_i * cos_theta_o > 0.f;
    const float inv_eta_c = rcp_(eta);
    const float eta_s = cos_theta_i > 0.f ? eta : inv_eta_c, inv_eta_s = cos_theta_i > 0.f ? inv_eta_c : eta;
    const V3 m = rough_half_vector(wi, wo, reflect, eta_s);
    const float wi_m = dot(wi, m), wo_m = dot(wo, m);
    active = active && wi_m * wi.z > 0.f && wo_m * wo.z > 0.f;
    const Microfacet distr(b.distribution, b.alpha_u, b.alpha_v);
    const float D = distr.eval(m), F = fresnel_(wi_m, eta).x, G = 1.f;
    const bool eval_r = reflect && active, eval_t = !reflect && active;
    if (!eval_r && !eval_t) return;
    if constexpr (POL) {
        m4_from_block(eval_r ? specular_reflection_(wo_m, eta) : specular_transmission_(wo_m, eta), W);
        rough_rotated(W, m, wo, wi);
    }
    const float w = eval_r ? F : 1.f - F;
    if (eval_r) {
        const float k = div_(D * G, 4.f * fabsf(cos_theta_i));
        w_out = div_(w * D * G, 4.f * fabsf(cos_theta_i));
        if constexpr (POL) m4_scale(W, k);
    } else {
        const float scale = sqr_(inv_eta_s);
        const float k = div_(scale * D * G * eta_s * eta_s * wi_m * wo_m, cos_theta_i * sqr_(wi_m + eta_s * wo_m));
        w_out = fabsf(w * k);
        if constexpr (POL) {
#pragma unroll
            for (int i = 0; i < 16; ++i) W[i] = fabsf(W[i] * k);
        }
    }
}

// ---- ThinDielectric (thindielectric.cpp:137-224): reflection + a NULL transmission lobe ------------------------------
MIB_D float thin_reflectance(const mib_bsdf &b, const SurfIt &si) {
    float r = fresnel_(fabsf(si.wi.z), b.eta).x;
    r *= div_(2.f, 1.f + r);                       // r' = r + trt + tr^3t + ...
    return r;
}
// eval_null_transmission of any BSDF, unpolarized value (null: 1, thindielectric: 1 - r', others: 0)
MIB_D float bsdf_null_transmission(const mib_bsdf &b, const SurfIt &si) {
    if (b.type == MIB_BSDF_NULL) return 1.f;
    if (b.type == MIB_BSDF_THINDIELECTRIC) return 1.f - thin_reflectance(b, si);
    return 0.f;
}

// ---- plastic / roughplastic (unpolarized values; polarized variants wrap them in a depolarizer) -------------------------
MIB_D float plastic_lerp_gather(const float *data, float x) {                          // roughplastic.cpp:392-402
    x *= (float) (MIB_ROUGH_TRANSMITTANCE_RES - 1);
    const uint32_t index = min((uint32_t) x, (uint32_t) (MIB_ROUGH_TRANSMITTANCE_RES - 2));
    return lerp_(data[index], data[index + 1], x - (float) index);
}
// diffuse base seen through the interface: diff / (1 - internal reflections)
MIB_D float plastic_base(const mib_bsdf &b, int c) {
    const float diff = b.reflectance[c];
    return div_(diff, b.nonlinear ? 1.f - diff * b.internal_reflectance : 1.f - b.internal_reflectance);
}
MIB_D float rough_prob_specular(const mib_bsdf &b, float cos_theta_i) {
    const float t_i = plastic_lerp_gather(b.ext_transmittance, cos_theta_i);
    const float ps = (1.f - t_i) * b.specular_sampling_weight, pd = t_i * (1.f - b.specular_sampling_weight);
    return div_(ps, ps + pd);
}
template <int NC> MIB_D void roughplastic_eval(const mib_bsdf &b, V3 wi, V3 wo, float *v) {
    const float cos_theta_i = wi.z, cos_theta_o = wo.z;
    if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) { for (int c = 0; c < NC; ++c) v[c] = 0.f; return; }
    const Microfacet distr(b.distribution, b.alpha_u, b.alpha_v);
    const V3 H = normalize(wo + wi);
    const float D = distr.eval(H), F = fresnel_(dot(wi, H), b.eta).x, G = distr.smith_g1(wi, H) * distr.smith_g1(wo, H);
    const float spec = div_(F * D * G, 4.f * cos_theta_i);
    const float t_i = plastic_lerp_gather(b.ext_transmittance, cos_theta_i), t_o = plastic_lerp_gather(b.ext_transmittance, cos_theta_o);
    const float k = kInvPi * b.inv_eta_2 * cos_theta_o * t_i * t_o;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        const int cc = NC == 3 ? c : 0;
        v[c] = (b.has_specular_reflectance ? spec * b.specular_reflectance[cc] : spec) + plastic_base(b, cc) * k;
    }
}
MIB_D float roughplastic_pdf(const mib_bsdf &b, V3 wi, V3 wo) {
    const float cos_theta_i = wi.z, cos_theta_o = wo.z;
    if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return 0.f;
    const float prob_specular = rough_prob_specular(b, cos_theta_i), prob_diffuse = 1.f - prob_specular;
    const Microfacet distr(b.distribution, b.alpha_u, b.alpha_v);
    const V3 H = normalize(wo + wi);
    float result = div_(distr.eval(H) * distr.smith_g1(wi, H), 4.f * cos_theta_i);
    result *= prob_specular;
    result += prob_diffuse * (kInvPi * wo.z);
    return result;
}
template <int NC> MIB_D void plastic_eval(const mib_bsdf &b, V3 wi, V3 wo, float *v) {
    const float cos_theta_i = wi.z, cos_theta_o = wo.z;
    if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) { for (int c = 0; c < NC; ++c) v[c] = 0.f; return; }
    const float f_i = fresnel_(cos_theta_i, b.eta).x, f_o = fresnel_(cos_theta_o, b.eta).x;
    const float k = (kInvPi * cos_theta_o) * b.inv_eta_2 * (1.f - f_i) * (1.f - f_o);
#pragma unroll
    for (int c = 0; c < NC; ++c) v[c] = plastic_base(b, NC == 3 ? c : 0) * k;
}
MIB_D float plastic_pdf(const mib_bsdf &b, V3 wi, V3 wo) {
    const float cos_theta_i = wi.z, cos_theta_o = wo.z;
    if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return 0.f;
    const float f_i = fresnel_(cos_theta_i, b.eta).x, ps = f_i * b.specular_sampling_weight;
    float pd = (1.f - f_i) * (1.f - b.specular_sampling_weight);
    pd = div_(pd, ps + pd);
    return (kInvPi * cos_theta_o) * pd;
}
// sample of both (plastic.cpp:202-264, roughplastic.cpp:256-311); a hit on the back side returns the zeroed sample
template <int NC> MIB_D void plastic_sample(const mib_bsdf &b, V3 wi, float s1, float s2x, float s2y, bool has_r, BsdfS &bs, float *v) {
    bs.wo = v3(0.f); bs.pdf = 0.f; bs.eta = 1.f; bs.sampled = 0;
#pragma unroll
    for (int c = 0; c < NC; ++c) v[c] = 0.f;
    const float cos_theta_i = wi.z;
    if (!has_r || !(cos_theta_i > 0.f)) return;
    const bool rough = b.type == MIB_BSDF_ROUGHPLASTIC;
    float f_i = 0.f, prob_specular;
    if (rough) prob_specular = rough_prob_specular(b, cos_theta_i);
    else {
        f_i = fresnel_(cos_theta_i, b.eta).x;
        const float ps = f_i * b.specular_sampling_weight, pd = (1.f - f_i) * (1.f - b.specular_sampling_weight);
        prob_specular = div_(ps, ps + pd);
    }
    const float prob_diffuse = 1.f - prob_specular;
    const bool sample_specular = s1 < prob_specular;
    if (rough) {
        if (sample_specular) {
            const Microfacet distr(b.distribution, b.alpha_u, b.alpha_v);
            float mpdf;
            const V3 m = distr.sample(wi, s2x, s2y, mpdf);
            const float d2 = 2.f * dot(wi, m);
            bs.wo = v3(fms_(m.x, d2, wi.x), fms_(m.y, d2, wi.y), fms_(m.z, d2, wi.z));
            bs.sampled = kBsdfGlossyRefl;
        } else {
            bs.wo = square_to_cosine_hemisphere(s2x, s2y);
            bs.sampled = kBsdfDiffuseRefl;
        }
        bs.pdf = roughplastic_pdf(b, wi, bs.wo);
        if (bs.pdf > 0.f) {
            roughplastic_eval<NC>(b, wi, bs.wo, v);
            const float ip = rcp_(bs.pdf);
#pragma unroll
            for (int c = 0; c < NC; ++c) v[c] *= ip;
        }
    } else if (sample_specular) {
        bs.wo = v3(-wi.x, -wi.y, wi.z);
        bs.pdf = prob_specular;
        bs.sampled = kBsdfDeltaRefl;
        const float w = div_(f_i, bs.pdf);
#pragma unroll
        for (int c = 0; c < NC; ++c) v[c] = b.has_specular_reflectance ? w * b.specular_reflectance[NC == 3 ? c : 0] : w;
    } else {
        bs.wo = square_to_cosine_hemisphere(s2x, s2y);
        bs.pdf = prob_diffuse * (kInvPi * bs.wo.z);
        bs.sampled = kBsdfDiffuseRefl;
        const float f_o = fresnel_(bs.wo.z, b.eta).x;
        const float k = div_(b.inv_eta_2 * (1.f - f_i) * (1.f - f_o), prob_diffuse);
#pragma unroll
        for (int c = 0; c < NC; ++c) v[c] = plastic_base(b, NC == 3 ? c : 0) * k;
    }
}

// ---- BSDF dispatch: V = value in the layout of the throughput (NC channels, or a 4x4 Mueller matrix in the local frame) -
template <int SPEC>
MIB_D void bsdf_sample(const mib_bsdf &b, const SurfIt &si, float s1, float s2x, float s2y, bool has_r, bool has_t, BsdfS &bs, float *V) {
    constexpr int NC = ST<SPEC>::NC;
    constexpr bool POL = ST<SPEC>::POL;
    if (b.type == MIB_BSDF_NULL) {
        bs.wo = -si.wi; bs.sampled = kBsdfNull; bs.eta = 1.f; bs.pdf = 1.f;
        if constexpr (POL) m4_identity(V, 1.f);
        else { for (int c = 0; c < NC; ++c) V[c] = 1.f; }
        return;
    }
    if (b.type == MIB_BSDF_THINDIELECTRIC) {
        const float r = thin_reflectance(b, si), t = 1.f - r;
        float weight;
        bool sel_r;
        bs.wo = v3(0.f); bs.pdf = 0.f; bs.eta = 1.f; bs.sampled = 0;
        if (has_r && has_t) { sel_r = s1 <= r; weight = 1.f; bs.pdf = sel_r ? r : t; }
        else if (has_r || has_t) { sel_r = has_r; weight = has_r ? r : t; bs.pdf = 1.f; }
        else { sel_r = false; weight = 0.f; }
        if (has_r || has_t) {
            bs.wo = sel_r ? v3(-si.wi.x, -si.wi.y, si.wi.z) : -si.wi;
            bs.sampled = sel_r ? kBsdfDeltaRefl : kBsdfNull;
        }
        if constexpr (POL) { m4_zero(V); V[0] = weight; }
        else { for (int c = 0; c < NC; ++c) V[c] = weight; }
        return;
    }
    if (b.type == MIB_BSDF_PLASTIC || b.type == MIB_BSDF_ROUGHPLASTIC) {
        float v[NC];
        plastic_sample<NC>(b, si.wi, s1, s2x, s2y, has_r, bs, v);
        if constexpr (POL) { m4_zero(V); V[0] = v[0]; }
        else { for (int c = 0; c < NC; ++c) V[c] = v[c]; }
        return;
    }
    if (b.type == MIB_BSDF_DIELECTRIC || b.type == MIB_BSDF_ROUGHDIELECTRIC) {
        float w;
        if (b.type == MIB_BSDF_DIELECTRIC) dielectric_sample<POL>(si.wi, b.eta, s1, has_r, has_t, bs, w, V);
        else rough_sample<POL>(b, si.wi, s1, s2x, s2y, has_r, has_t, bs, w, V);
        if constexpr (!POL) { for (int c = 0; c < NC; ++c) V[c] = w; }
        return;
    }
    if (!has_r) {                                                      // diffuse.cpp:108-111: the diffuse lobe is disabled
        bs.wo = v3(0.f); bs.pdf = 0.f; bs.eta = 0.f; bs.sampled = 0;
        if constexpr (POL) m4_zero(V);
        else { for (int c = 0; c < NC; ++c) V[c] = 0.f; }
        return;
    }
    bs.wo = square_to_cosine_hemisphere(s2x, s2y);                     // diffuse.cpp:101-126
    bs.pdf = kInvPi * bs.wo.z;
    bs.eta = 1.f;
    bs.sampled = kBsdfDiffuseRefl;
    const bool ok = si.wi.z > 0.f && bs.pdf > 0.f;
    if constexpr (POL) { m4_zero(V); V[0] = ok ? bsdf_reflectance(b, si, 0) : 0.f; }
    else { for (int c = 0; c < NC; ++c) V[c] = ok ? bsdf_reflectance(b, si, NC == 3 ? c : 0) : 0.f; }
}
template <int SPEC>
MIB_D void bsdf_eval(const mib_bsdf &b, const SurfIt &si, V3 wo, float *V) {
    constexpr int NC = ST<SPEC>::NC;
    constexpr bool POL = ST<SPEC>::POL;
    if (b.type == MIB_BSDF_DIELECTRIC || b.type == MIB_BSDF_ROUGHDIELECTRIC) {
        float w;
        if (b.type == MIB_BSDF_DIELECTRIC) dielectric_eval<POL>(si.wi, wo, b.eta, b.inv_eta, w, V);
        else rough_eval<POL>(b, si.wi, wo, w, V);
        if constexpr (!POL) { for (int c = 0; c < NC; ++c) V[c] = w; }
        return;
    }
    if (b.type == MIB_BSDF_PLASTIC || b.type == MIB_BSDF_ROUGHPLASTIC) {
        float v[NC];
        if (b.type == MIB_BSDF_PLASTIC) plastic_eval<NC>(b, si.wi, wo, v);
        else roughplastic_eval<NC>(b, si.wi, wo, v);
        if constexpr (POL) { m4_zero(V); V[0] = v[0]; }
        else { for (int c = 0; c < NC; ++c) V[c] = v[c]; }
        return;
    }
    const bool ok = b.type == MIB_BSDF_DIFFUSE && si.wi.z > 0.f && wo.z > 0.f;  // diffuse.cpp:128-160; null, thindielectric: 0
    if constexpr (POL) { m4_zero(V); V[0] = ok ? bsdf_reflectance(b, si, 0) * kInvPi * wo.z : 0.f; }
    else { for (int c = 0; c < NC; ++c) V[c] = ok ? bsdf_reflectance(b, si, NC == 3 ? c : 0) * kInvPi * wo.z : 0.f; }
}

// pdf (roughdielectric.cpp:521-603)
MIB_D float rough_pdf(const mib_bsdf &b, V3 wi, V3 wo) {
    const float eta = b.eta, cos_theta_i = wi.z, cos_theta_o = wo.z;
    bool active = cos_theta_i != 0.f;
    const bool reflect = cos_theta_i * cos_theta_o > 0.f;
    const float eta_s = cos_theta_i > 0.f ? eta : rcp_(eta);
    const V3 m = rough_half_vector(wi, wo, reflect, eta_s);
    const float wi_m = dot(wi, m), wo_m = dot(wo, m);
    active = active && wi_m * wi.z > 0.f && wo_m * wo.z > 0.f;
    const float dwh_dwo = reflect ? rcp_(4.f * wo_m) : div_(eta_s * eta_s * wo_m, sqr_(wi_m + eta_s * wo_m));
    const Microfacet distr(b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
    const V3 wi_up = v3(mulsign_ad(wi.x, cos_theta_i), mulsign_ad(wi.y, cos_theta_i), mulsign_ad(wi.z, cos_theta_i));
    float prob = distr.pdf(wi_up, m);
    const float F = fresnel_(wi_m, eta).x;
    prob *= reflect ? F : 1.f - F;
    return active ? prob * fabsf(dwh_dwo) : 0.f;
}
MIB_D float bsdf_pdf(const mib_bsdf &b, const SurfIt &si, V3 wo) {
    if (b.type == MIB_BSDF_NULL || b.type == MIB_BSDF_DIELECTRIC || b.type == MIB_BSDF_THINDIELECTRIC) return 0.f;
    if (b.type == MIB_BSDF_ROUGHDIELECTRIC) return rough_pdf(b, si.wi, wo);
    if (b.type == MIB_BSDF_PLASTIC) return plastic_pdf(b, si.wi, wo);
    if (b.type == MIB_BSDF_ROUGHPLASTIC) return roughplastic_pdf(b, si.wi, wo);
    return (si.wi.z > 0.f && wo.z > 0.f) ? kInvPi * wo.z : 0.f;
}

} // namespace mib

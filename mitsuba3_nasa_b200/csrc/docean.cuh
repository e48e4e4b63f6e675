// docean.cuh -- the atmosphere-ocean interface of volpathaos on the device.
//   SmoothDielectric   src/bsdfs/dielectric.cpp:247-468         (sample; the fork's non-zero eval; pdf = 0)
//   RoughDielectric    src/bsdfs/roughdielectric.cpp:248-603    (sample / eval with G = 1 / pdf)
//   refractive NEE     src/integrators/volpathaos.cpp:232-290 (main loop), :299-455 (sample_emitter)
// One thread walks one path (k_forward<SPEC, 4>).  Differences from the walk of dwalk.cuh, which stays the tuned path for
// scenes without an interface:
//   * the shadow walk carries its transmittance as a full Mueller matrix in polarized mode (a dielectric between the
//     scattering point and the emitter multiplies it by a non-diagonal matrix), and returns the emitted Stokes 4-vector;
//   * the path remembers the BSDF of the interface while it is below it (`refractive_bsdf`, volpathaos.cpp:80,282-285) and
//     the shadow ray then leaves in a direction that the interface refracts toward the emitter (:312-337).
// The reference's quirks are kept (SmoothDielectric::eval without Fresnel factor, eval with G = 1, the polarized sampling
// weight's mixed angles); oracle/obsdf.h, oracle/oroughdielectric.h and oracle/ofresnel.h list them.
#pragma once
#include "dwalk.cuh"
#include "dfresnel.cuh"
#include "dmicrofacet.cuh"

namespace mib {

constexpr uint32_t kBsdfGlossyRefl = 0x8, kBsdfGlossyTrans = 0x10, kBsdfDeltaRefl = 0x20, kBsdfDeltaTrans = 0x40,
                   kBsdfTransmission = 0x10 | 0x40;

MIB_D uint32_t bsdf_flags(const mib_bsdf &b) {
    switch (b.type) {
        case MIB_BSDF_NULL: return kBsdfNull;
        case MIB_BSDF_DIELECTRIC: return kBsdfDeltaRefl | kBsdfDeltaTrans;
        case MIB_BSDF_ROUGHDIELECTRIC: return kBsdfGlossyRefl | kBsdfGlossyTrans;
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
    const Microfacet distr(b.distribution, b.alpha_u, b.alpha_v);
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
    const float g1 = distr.smith_g1(bs.wo, m);
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
    const bool ok = b.type != MIB_BSDF_NULL && si.wi.z > 0.f && wo.z > 0.f;     // diffuse.cpp:128-160; null: 0
    if constexpr (POL) { m4_zero(V); V[0] = ok ? bsdf_reflectance(b, si, 0) * kInvPi * wo.z : 0.f; }
    else { for (int c = 0; c < NC; ++c) V[c] = ok ? bsdf_reflectance(b, si, NC == 3 ? c : 0) * kInvPi * wo.z : 0.f; }
}

// R += first column of (T * V) * Em, Em = the emitted Stokes vector as a matrix whose only non-zero column is the first
// (result + throughput * value * emitted, volpathaos.cpp:203 / :246); unpolarized: per channel
template <int SPEC>
MIB_D void add_contribution(float *R, const float *T, const float *V, const float *em) {
    if constexpr (ST<SPEC>::POL) {
        float X[16];
        m4_mul(T, V, X);
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            float acc = X[4 * r] * em[0];
            acc = fma_(X[4 * r + 1], em[1], acc);
            acc = fma_(X[4 * r + 2], em[2], acc);
            acc = fma_(X[4 * r + 3], em[3], acc);
            R[r] += acc;
        }
    } else {
#pragma unroll
        for (int c = 0; c < ST<SPEC>::NC; ++c) R[c] += (T[c] * V[c]) * em[c];
    }
}

// ---------------------------------------------------------------------------------------------------------------------------
// Shadow walk of volpathaos with an interface (volpathaos.cpp:299-455).  em: NR floats (Stokes 4-vector when polarized).
// ---------------------------------------------------------------------------------------------------------------------------
template <int SPEC>
MIB_D void sample_emitter_ocean(Pcg32 &rng, Counters &cnt, V3 ref_p, V3 ref_n, int medium, uint32_t channel, int refractive_bsdf,
                                DirSample<ST<SPEC>::NC> &ds, float *em) {
    constexpr int NC = ST<SPEC>::NC, NT = ST<SPEC>::NT, NR = ST<SPEC>::NR;
    constexpr bool POL = ST<SPEC>::POL;
    const bool ltr = sc.d.integrator.rng_arg_order == MIB_RNG_ARGS_LEFT_TO_RIGHT;
    float tr[NT], ev[NC];
    T_init<SPEC>(tr);
    const float sx = rng.next(), sy = rng.next();
    sample_emitter_direction<NC>(ref_p, sx, sy, ds, ev);
    const V3 emitter_d = normalize(ds.d);
    bool active = true;
    if (ds.pdf == 0.f) { for (int c = 0; c < NC; ++c) ev[c] = 0.f; active = false; }
    if (active && refractive_bsdf >= 0) {
        // a direction that the interface (normal (0, 0, 1) in world space) refracts into emitter_d (:312-337)
        SurfIt rsi;
        rsi.wi = emitter_d;
        float s1, s2x, s2y;
        if (ltr) { s1 = rng.next(); s2x = rng.next(); s2y = rng.next(); }
        else { s2x = rng.next(); s2y = rng.next(); s1 = rng.next(); }
        BsdfS bs;
        float V[NT];
        bsdf_sample<SPEC>(sc.d.bsdfs[refractive_bsdf], rsi, s1, s2x, s2y, false, true, bs, V);
        if (!(bs.pdf > kEps)) { for (int c = 0; c < NC; ++c) ev[c] = 0.f; }
        else {
            ds.d = -bs.wo;
            const float ip = rcp_(bs.pdf);
#pragma unroll
            for (int c = 0; c < NC; ++c) ev[c] *= ip;
        }
    }
    if (!active) {
        if constexpr (POL) { em[0] = ev[0]; em[1] = 0.f; em[2] = 0.f; em[3] = 0.f; }
        else { for (int c = 0; c < NC; ++c) em[c] = ev[c]; }
        return;
    }
    Ray ray = spawn_ray(ref_p, ref_n, ds.d);
    const float max_dist = ds.dist * (1.f - kShadowEps);
    float total_dist = 0.f;
    SurfIt si;
    si.t = inf_f(); si.shape = -1;
    bool needs_intersection = true;
    while (active) {
        const float remaining = max_dist - total_dist;
        ray.maxt = remaining;
        active = active && remaining > 0.f;
        if (!active) break;
        bool escaped_medium = false, active_medium = medium >= 0, active_surface = !active_medium;
        if (active_medium) {
            const mib_medium &md = sc.d.media[medium];
            cnt.n_nee_steps++;
            MedIt<NC> mei;
            sample_interaction<NC>(medium, ray, rng.next(), channel, mei, cnt);
            if (md.type == MIB_MEDIUM_HOMOGENEOUS && valid(mei)) ray.maxt = fminf(mei.t, remaining);
            if (needs_intersection) ray_intersect(ray, si);
            if (si.t < mei.t) mei.t = inf_f();
            needs_intersection = needs_intersection && !valid(si);
            const bool is_spectral = md.has_spectral_extinction != 0;
            if (is_spectral) {
                const float t = fminf(remaining, fminf(mei.t, si.t)) - mei.mint;
                float trv[NC], ffp[NC], u[NC];
                const bool surf = si.t < mei.t || mei.t > remaining;
#pragma unroll
                for (int c = 0; c < NC; ++c) { trv[c] = exp_(-t * mei.maj[c]); ffp[c] = surf ? trv[c] : trv[c] * mei.maj[c]; }
                const float tr_pdf = ch<NC>(ffp, channel), ip = rcp_(tr_pdf);
#pragma unroll
                for (int c = 0; c < NC; ++c) u[c] = tr_pdf > 0.f ? trv[c] * ip : 0.f;
                T_mul_u<SPEC>(tr, u);
            }
            if ((mei.t > remaining) && valid(mei)) total_dist = ds.dist;
            if (mei.t > remaining) mei.t = inf_f();
            escaped_medium = !valid(mei);
            active_medium = valid(mei);
            if (active_medium) {
                total_dist += mei.t;
                ray.o = mei.p;
                si.t = si.t - mei.t;
                float u[NC];
#pragma unroll
                for (int c = 0; c < NC; ++c) u[c] = is_spectral ? mei.sigma_n[c] : div_(mei.sigma_n[c], mei.maj[c]);
                T_mul_u<SPEC>(tr, u);
            }
        }
        const bool intersect = active_surface && needs_intersection;
        if (intersect) ray_intersect(ray, si);
        needs_intersection = needs_intersection && !intersect;
        active_surface = active_surface || escaped_medium;
        if (active_surface) total_dist += si.t;
        active_surface = active_surface && valid(si) && !active_medium;
        if (active_surface) {
            const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
            if (bsdf.type == MIB_BSDF_NULL) {
                if constexpr (POL) {                                   // eval_null_transmission in world space (:417-421)
                    float b11, b12, b21, b22;
                    null_block(si, b11, b12, b21, b22);
                    m4_mul_right_block(tr, b11, b12, b21, b22);
                }
            } else {                                                   // bsdf->eval(ctx, si, to_local(emitter_d)) (:423-431)
                const V3 wo = to_local(si.sh, emitter_d);
                float V[NT];
                bsdf_eval<SPEC>(bsdf, si, wo, V);
                if constexpr (POL) m4_to_world(si.sh, V, -wo, si.wi);
                T_mul_W<SPEC>(tr, V);
                ray.d = emitter_d;
            }
            ray = spawn_ray(si.p, si.n, ray.d);
        }
        ray.maxt = remaining;
        needs_intersection = needs_intersection || active_surface;
        active = active && (active_medium || active_surface) && T_any_nonzero<SPEC>(tr);
        if (active_surface && is_medium_transition(si)) medium = target_medium(si, ray.d);
    }
    if constexpr (POL) {
#pragma unroll
        for (int r = 0; r < NR; ++r) em[r] = tr[4 * r] * ev[0];
    } else {
#pragma unroll
        for (int c = 0; c < NC; ++c) em[c] = tr[c] * ev[c];
    }
}

// ---------------------------------------------------------------------------------------------------------------------------
// One iteration of volpathaos' loop (volpathaos.cpp:85-294) for scenes with an interface.  State: PathState (dwalk.cuh),
// whose refractive_bsdf is the BSDF of the interface while the path is below it.
// ---------------------------------------------------------------------------------------------------------------------------
template <int SPEC>
MIB_D void ocean_step(Pcg32 &rng, Counters &cnt, PathState<SPEC> &st) {
    constexpr int NC = ST<SPEC>::NC, NT = ST<SPEC>::NT, NR = ST<SPEC>::NR;
    constexpr bool POL = ST<SPEC>::POL;
    const mib_integrator &I = sc.d.integrator;
    float *T = st.T, *R = st.R;
    Ray &ray = st.ray;
    SurfIt &si = st.si;
    int &medium = st.medium;
    uint32_t &depth = st.depth;
    const uint32_t channel = st.channel;
    bool &active = st.active, &needs_intersection = st.needs_intersection;
    const uint32_t maxd = (uint32_t) I.max_depth, rrd = (uint32_t) I.rr_depth;
    const bool ltr = I.rng_arg_order == MIB_RNG_ARGS_LEFT_TO_RIGHT;
    MedIt<NC> mei;

    active = active && T_any_nonzero<SPEC>(T);
    const float q = fminf(T_max<SPEC>(T) * sqr_(st.eta), .95f);
    const bool perform_rr = depth > rrd;
    if (active) { const float u = rng.next(); active = (u < q) || !perform_rr; }
    if (perform_rr) T_mul_f<SPEC>(T, rcp_(q));
    active = active && depth < maxd;
    if (!active) return;

    bool active_medium = medium >= 0, active_surface = !active_medium;
    bool act_null = false, act_scatter = false, escaped = false, is_spectral = false;
    if (active_medium) {
        const mib_medium &md = sc.d.media[medium];
        is_spectral = md.has_spectral_extinction != 0;
        sample_interaction<NC>(medium, ray, rng.next(), channel, mei, cnt);
        if (md.type == MIB_MEDIUM_HOMOGENEOUS && valid(mei)) ray.maxt = mei.t;
        if (needs_intersection) ray_intersect(ray, si);
        needs_intersection = needs_intersection && !valid(si);
        if (si.t < mei.t) mei.t = inf_f();
        if (is_spectral) {                                               // medium.cpp:99-111
            const float t = fminf(mei.t, si.t) - mei.mint;
            float trv[NC], ffp[NC], u[NC];
            const bool surf = si.t < mei.t;
#pragma unroll
            for (int c = 0; c < NC; ++c) { trv[c] = exp_(-t * mei.maj[c]); ffp[c] = surf ? trv[c] : trv[c] * mei.maj[c]; }
            const float tr_pdf = ch<NC>(ffp, channel), ip = rcp_(tr_pdf);
#pragma unroll
            for (int c = 0; c < NC; ++c) u[c] = tr_pdf > 0.f ? trv[c] * ip : 0.f;
            T_mul_u<SPEC>(T, u);
        }
        escaped = !valid(mei);
        active_medium = valid(mei);
        if (active_medium) {
            const float u = rng.next();
            act_null = u >= div_(ch<NC>(mei.sigma_t, channel), ch<NC>(mei.maj, channel));
            act_scatter = !act_null;
            if (is_spectral && act_null) {
                const float mjc = ch<NC>(mei.maj, channel), isn = rcp_(ch<NC>(mei.sigma_n, channel));
                float u2[NC];
#pragma unroll
                for (int c = 0; c < NC; ++c) u2[c] = (mei.sigma_n[c] * mjc) * isn;
                T_mul_u<SPEC>(T, u2);
            }
            if (act_scatter) depth += 1;
        }
    }
    active = active && depth < maxd;
    act_scatter = act_scatter && active;
    if (act_null) { cnt.n_null++; ray.o = mei.p; si.t = si.t - mei.t; }

    if (act_scatter) {
        cnt.n_real++;
        const mib_medium &md = sc.d.media[mei.medium];
        float u[NC];
        if (is_spectral) {
            const float mjc = ch<NC>(mei.maj, channel), ist = rcp_(ch<NC>(mei.sigma_t, channel));
#pragma unroll
            for (int c = 0; c < NC; ++c) u[c] = (mei.sigma_s[c] * mjc) * ist;
        } else {
#pragma unroll
            for (int c = 0; c < NC; ++c) u[c] = div_(mei.sigma_s[c], mei.sigma_t[c]);
        }
        T_mul_u<SPEC>(T, u);
        const int phase = md.phase;
        const Frame3 mei_sh = make_frame(ray.d);
        const bool sample_emitters = md.sample_emitters != 0;
        st.specular_chain = !sample_emitters;
        float W[NT];
        if (sample_emitters) {
            DirSample<NC> ds;
            float em[NR];
            sample_emitter_ocean<SPEC>(rng, cnt, mei.p, v3(0.f), medium, channel, st.refractive_bsdf, ds, em);
            const V3 wo = to_local(mei_sh, ds.d);
            phase_eval<SPEC>(phase, wo, W);
            if constexpr (POL) m4_to_world(mei_sh, W, -wo, v3(0.f, 0.f, -1.f));
            add_contribution<SPEC>(R, T, W, em);
        }
        float s1, s2x, s2y;
        if (ltr) { s1 = rng.next(); s2x = rng.next(); s2y = rng.next(); }
        else { s2x = rng.next(); s2y = rng.next(); s1 = rng.next(); }
        (void) s1;
        V3 wo;
        const float ppdf = phase_sample<SPEC>(phase, s2x, s2y, wo, W);
        if constexpr (POL) m4_to_world(mei_sh, W, -wo, v3(0.f, 0.f, -1.f));
        act_scatter = act_scatter && ppdf > 0.f;
        if (act_scatter) {
            T_mul_W<SPEC>(T, W);
            ray = spawn_ray(mei.p, v3(0.f), to_world(mei_sh, wo));
            needs_intersection = true;
        }
    }

    active_surface = active_surface || escaped;
    if (active_surface && needs_intersection) ray_intersect(ray, si);
    active_surface = active_surface && valid(si);
    if (active_surface) {
        cnt.n_surface++;
        const int bsdf_index = sc.d.shapes[si.shape].bsdf;
        const mib_bsdf &bsdf = sc.d.bsdfs[bsdf_index];
        const uint32_t flags = bsdf_flags(bsdf);
        float V[NT];
        if ((flags & kBsdfSmooth) && (depth + 1 < maxd)) {
            const bool diffuse = (flags & kBsdfDiffuseRefl) != 0, refractive = (flags & kBsdfTransmission) != 0;     // :234-236
            if (diffuse || (refractive && si.wi.z > 0.f)) {
                DirSample<NC> ds;
                float em[NR];
                sample_emitter_ocean<SPEC>(rng, cnt, si.p, si.n, medium, channel, st.refractive_bsdf, ds, em);
                const V3 wo = to_local(si.sh, ds.d);
                bsdf_eval<SPEC>(bsdf, si, wo, V);
                if constexpr (POL) m4_to_world(si.sh, V, -wo, si.wi);
                add_contribution<SPEC>(R, T, V, em);
            }
        }
        float s1, s2x, s2y;
        if (ltr) { s1 = rng.next(); s2x = rng.next(); s2y = rng.next(); }
        else { s2x = rng.next(); s2y = rng.next(); s1 = rng.next(); }
        BsdfS bs;
        bsdf_sample<SPEC>(bsdf, si, s1, s2x, s2y, true, true, bs, V);
        if constexpr (POL) m4_to_world(si.sh, V, -bs.wo, si.wi);
        T_mul_W<SPEC>(T, V);
        st.eta *= bs.eta;
        const bool non_null = !(bs.sampled & kBsdfNull);
        if (non_null) depth += 1;
        const Ray bsdf_ray = spawn_ray(si.p, si.n, to_world(si.sh, bs.wo));
        // "Refraction occurred, intentionally or not.  Assumes surface is aligned with x-y plane in world space" (:262-267)
        const bool refracted = non_null && (ray.d.z * bsdf_ray.d.z > 0.f);
        ray = bsdf_ray;
        needs_intersection = true;
        if (refracted) st.refractive_bsdf = st.refractive_bsdf < 0 ? bsdf_index : -1;          // :282-285
        st.specular_chain = st.specular_chain || (non_null && (bs.sampled & kBsdfDelta));
        st.specular_chain = st.specular_chain && !(bs.sampled & kBsdfSmooth);
        if (is_medium_transition(si)) medium = target_medium(si, ray.d);
    }
    active = active && (active_surface || active_medium);
}

} // namespace mib

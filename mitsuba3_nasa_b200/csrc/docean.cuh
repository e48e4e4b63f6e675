// docean.cuh -- the atmosphere-ocean interface of volpathaos on the device.
//   refractive NEE     src/integrators/volpathaos.cpp:232-290 (main loop), :299-455 (sample_emitter)
//   BSDFs              dbsdf.cuh (dielectric.cpp:247-468, roughdielectric.cpp:248-603)
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

namespace mib {

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
            if (bsdf_flags(bsdf) & kBsdfNull) {
                if (bsdf.type == MIB_BSDF_NULL) {
                    if constexpr (POL) {                               // eval_null_transmission in world space (:417-421)
                        float b11, b12, b21, b22;
                        null_block(si, b11, b12, b21, b22);
                        m4_mul_right_block(tr, b11, b12, b21, b22);
                    }
                } else {                                               // thindielectric: depolarizer(1 - r') (thindielectric.cpp:202-215)
                    float v[NC];
#pragma unroll
                    for (int c = 0; c < NC; ++c) v[c] = bsdf_null_transmission(bsdf, si);
                    T_mul_depol<SPEC>(tr, v);
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

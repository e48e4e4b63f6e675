// dmis.cuh -- volpathmis, the spectral-MIS volumetric path tracer (src/integrators/volpathmis.cpp:126-598 with
// use_spectral_mis, the plugin's default), as straight-line device code: one thread walks one path.
// Instead of a throughput the path carries two NC x NC weight matrices, p_over_f[i][j] = (pdf of the sampling strategy of
// channel j) / (contribution in channel i), one for the unidirectional strategy and one for next-event estimation; the
// balance heuristic over channels and strategies is NC / (row sum) (mis_weight).  Unpolarized spectrum kinds only (the
// reference throws in polarized variants, :134-136).  Every decision consumes the lane's PCG32 stream in the reference's order.
#pragma once
#include "dwalk.cuh"

namespace mib {

// volpathmis.cpp:506-527.  SpectralMis: W[i][j] *= p[j] / f[i], non-finite ratios and NaN products become 0; otherwise
// (mib_integrator.spectral_mis = 0) the weights are one spectrum, kept in row 0: W[j] *= p[channel] / f[j], non-finite -> 0
template <int NC> MIB_D void mis_update(float *W, const float *p, const float *f, uint32_t channel) {
    if (!sc.d.integrator.spectral_mis) {
        const float pdf = ch<NC>(p, channel);
#pragma unroll
        for (int j = 0; j < NC; ++j) {
            const float ratio = W[j] * div_(pdf, f[j]);
            W[j] = isfinite(ratio) ? ratio : 0.f;
        }
        return;
    }
#pragma unroll
    for (int i = 0; i < NC; ++i) {
        const float inv_f = rcp_(f[i]);
#pragma unroll
        for (int j = 0; j < NC; ++j) {
            float ratio = p[j] * inv_f;
            ratio = isfinite(ratio) ? ratio : 0.f;
            ratio *= W[i * NC + j];
            W[i * NC + j] = isnan(ratio) ? 0.f : ratio;
        }
    }
}
template <int NC> MIB_D void mis_update(float *W, float p, const float *f, uint32_t channel) {
    float pv[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) pv[c] = p;
    mis_update<NC>(W, pv, f, channel);
}
template <int NC> MIB_D void mis_update(float *W, float p, float f, uint32_t channel) {
    float pv[NC], fv[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) { pv[c] = p; fv[c] = f; }
    mis_update<NC>(W, pv, fv, channel);
}
// :530-543 / :546-563
template <int NC> MIB_D void mis_weights(const float *W, float *w) {
    if (!sc.d.integrator.spectral_mis) {
        float mn = fabsf(W[0]);
#pragma unroll
        for (int j = 1; j < NC; ++j) mn = fminf(mn, fabsf(W[j]));
#pragma unroll
        for (int j = 0; j < NC; ++j) w[j] = mn == 0.f ? 0.f : div_(1.f, W[j]);
        return;
    }
#pragma unroll
    for (int i = 0; i < NC; ++i) {
        float s = W[i * NC];
#pragma unroll
        for (int j = 1; j < NC; ++j) s += W[i * NC + j];
        w[i] = s == 0.f ? 0.f : div_((float) NC, s);
    }
}
template <int NC> MIB_D void mis_weights(const float *A, const float *B, float *w) {
    if (!sc.d.integrator.spectral_mis) {
        float sum[NC], mn = 0.f;
#pragma unroll
        for (int j = 0; j < NC; ++j) { sum[j] = A[j] + B[j]; mn = j == 0 ? fabsf(sum[0]) : fminf(mn, fabsf(sum[j])); }
#pragma unroll
        for (int j = 0; j < NC; ++j) w[j] = mn == 0.f ? 0.f : div_(1.f, sum[j]);
        return;
    }
#pragma unroll
    for (int i = 0; i < NC; ++i) {
        float s = A[i * NC] + B[i * NC];
#pragma unroll
        for (int j = 1; j < NC; ++j) s += A[i * NC + j] + B[i * NC + j];
        w[i] = s == 0.f ? 0.f : div_((float) NC, s);
    }
}
template <int NC> MIB_D bool mis_any(const float *W) {
    float w[NC];
    mis_weights<NC>(W, w);
    bool any = false;
#pragma unroll
    for (int c = 0; c < NC; ++c) any = any || w[c] != 0.f;
    return any;
}

template <int SPEC> struct MisState {
    static constexpr int NC = ST<SPEC>::NC;
    float W[NC * NC], Wnee[NC * NC];          // p_over_f, p_over_f_nee
    float R[NC];
    Ray ray;
    SurfIt si;
    float eta;
    V3 last_p;
    int medium;
    uint32_t depth, channel;
    bool active, specular_chain, needs_intersection, last_event_was_null;
};

template <int SPEC>
MIB_D void mis_init(Pcg32 &rng, const Ray &ray, MisState<SPEC> &st) {
    constexpr int NC = ST<SPEC>::NC;
#pragma unroll
    for (int i = 0; i < NC * NC; ++i) { st.W[i] = 1.f; st.Wnee[i] = 1.f; }
#pragma unroll
    for (int i = 0; i < NC; ++i) st.R[i] = 0.f;
    st.ray = ray;
    st.eta = 1.f;
    st.medium = sc.d.sensor.medium;
    st.active = true;
    st.specular_chain = !sc.d.integrator.hide_emitters;
    st.needs_intersection = true;
    st.last_event_was_null = false;
    st.depth = 0; st.channel = 0;
    if (SPEC == MIB_SPEC_RGB) st.channel = (uint32_t) fminf(rng.next() * 3.f, 2.f);
    st.si.t = inf_f(); st.si.shape = -1; st.si.wi = v3(0.f);
    st.last_p = v3(0.f);
}

// volpathmis.cpp:378-503: next-event estimation with ratio tracking; both weight sets are carried to the emitter
template <int SPEC>
MIB_D void sample_emitter_mis(Pcg32 &rng, Counters &cnt, V3 ref_p, V3 ref_n, const SurfIt *ref_si, int medium, uint32_t channel,
                              const float *W0, float *Wn, float *Wu, DirSample<ST<SPEC>::NC> &ds, float *emitted) {
    constexpr int NC = ST<SPEC>::NC;
#pragma unroll
    for (int i = 0; i < NC * NC; ++i) { Wn[i] = W0[i]; Wu[i] = W0[i]; }
    float sx = rng.next(), sy = rng.next();
    sample_emitter_direction<NC>(ref_p, sx, sy, ds, emitted);
#pragma unroll
    for (int c = 0; c < NC; ++c) emitted[c] = ds.pdf == 0.f ? 0.f : emitted[c] * ds.pdf;
    bool active = ds.pdf != 0.f;
    if (!active) return;
    mis_update<NC>(Wn, ds.pdf, 1.f, channel);

    Ray ray = spawn_ray_to(ref_p, ref_n, ds.p);
    const float max_dist = ray.maxt;
    if (ref_si && is_medium_transition(*ref_si)) medium = target_medium(*ref_si, ray.d);
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
            needs_intersection = false;
            const bool is_spectral = md.has_spectral_extinction != 0;
            if (is_spectral) {
                const float t = fminf(remaining, fminf(mei.t, si.t)) - mei.mint;
                const bool surf = si.t < mei.t || mei.t > remaining;
                float trv[NC], ffp[NC];
#pragma unroll
                for (int c = 0; c < NC; ++c) { trv[c] = exp_(-t * mei.maj[c]); ffp[c] = surf ? trv[c] : trv[c] * mei.maj[c]; }
                mis_update<NC>(Wn, ffp, trv, channel);
                mis_update<NC>(Wu, ffp, trv, channel);
            }
            if ((mei.t > remaining) && valid(mei)) total_dist = ds.dist;
            if (mei.t > remaining) mei.t = inf_f();
            escaped_medium = !valid(mei);
            active_medium = valid(mei);
            if (active_medium) {
                total_dist += mei.t;
                ray.o = mei.p;
                si.t = si.t - mei.t;
                float r[NC];
#pragma unroll
                for (int c = 0; c < NC; ++c) r[c] = div_(mei.sigma_n[c], mei.maj[c]);
                if (is_spectral) { mis_update<NC>(Wn, 1.f, mei.sigma_n, channel); mis_update<NC>(Wu, r, mei.sigma_n, channel); }
                else { mis_update<NC>(Wn, 1.f, r, channel); mis_update<NC>(Wu, mei.sigma_n, mei.sigma_n, channel); }
            }
        }
        if (active_surface && needs_intersection) ray_intersect(ray, si);
        active_surface = active_surface || escaped_medium;
        if (active_surface) total_dist += si.t;
        active_surface = active_surface && valid(si) && !active_medium;
        if (active_surface) {
            const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
            const float nt = bsdf_null_transmission(bsdf, si);
            mis_update<NC>(Wn, 1.f, nt, channel);
            mis_update<NC>(Wu, 1.f, nt, channel);
            ray = spawn_ray(si.p, si.n, ray.d);
        }
        ray.maxt = remaining;
        needs_intersection = needs_intersection || active_surface;
        bool go;
        if (sc.d.integrator.spectral_mis) go = mis_any<NC>(Wu);
        else {
            go = false;
#pragma unroll
            for (int j = 0; j < NC; ++j) go = go || Wu[j] != 0.f || Wn[j] != 0.f;
        }
        active = active && (active_medium || active_surface) && go;
        if (active_surface && is_medium_transition(si)) medium = target_medium(si, ray.d);
    }
}

// One iteration of the reference's while loop (volpathmis.cpp:175-374)
template <int SPEC>
MIB_D void mis_step(Pcg32 &rng, Counters &cnt, MisState<SPEC> &st) {
    constexpr int NC = ST<SPEC>::NC;
    static_assert(!ST<SPEC>::POL, "volpathmis has no polarized mode");
    const mib_integrator &I = sc.d.integrator;
    float *W = st.W, *Wnee = st.Wnee, *R = st.R;
    Ray &ray = st.ray;
    SurfIt &si = st.si;
    int &medium = st.medium;
    uint32_t &depth = st.depth;
    const uint32_t channel = st.channel;
    bool &active = st.active, &specular_chain = st.specular_chain, &needs_intersection = st.needs_intersection;
    const uint32_t maxd = (uint32_t) I.max_depth, rrd = (uint32_t) I.rr_depth;
    const bool ltr = I.rng_arg_order == MIB_RNG_ARGS_LEFT_TO_RIGHT;
    MedIt<NC> mei;

    {
        float w[NC];
        mis_weights<NC>(W, w);
        float wm = w[0];
#pragma unroll
        for (int c = 1; c < NC; ++c) wm = fmaxf(wm, w[c]);
        const float q = fminf(wm * sqr_(st.eta), .95f);
        const bool perform_rr = !st.last_event_was_null && depth > rrd;
        const float u = rng.next();
        active = !(u >= q && perform_rr);
        if (perform_rr) mis_update<NC>(W, q, 1.f, channel);
        st.last_event_was_null = false;
        active = active && depth < maxd;
        active = active && mis_any<NC>(W);
        if (!active) return;
    }

    bool active_medium = medium >= 0, active_surface = !active_medium;
    bool escaped = false, is_spectral = false;
    if (active_medium) {
        const mib_medium &md = sc.d.media[medium];
        is_spectral = md.has_spectral_extinction != 0;
        sample_interaction<NC>(medium, ray, rng.next(), channel, mei, cnt);
        if (md.type == MIB_MEDIUM_HOMOGENEOUS && valid(mei)) ray.maxt = mei.t;
        if (needs_intersection) ray_intersect(ray, si);
        needs_intersection = false;
        if (si.t < mei.t) mei.t = inf_f();
        if (is_spectral) {                                                   // medium.cpp:99-111
            const float t = fminf(mei.t, si.t) - mei.mint;
            const bool surf = si.t < mei.t;
            float trv[NC], ffp[NC];
#pragma unroll
            for (int c = 0; c < NC; ++c) { trv[c] = exp_(-t * mei.maj[c]); ffp[c] = surf ? trv[c] : trv[c] * mei.maj[c]; }
            mis_update<NC>(W, ffp, trv, channel);
            mis_update<NC>(Wnee, ffp, trv, channel);
        }
        escaped = !valid(mei);
        active_medium = valid(mei);
    }
    if (active_medium) {
        const mib_medium &md = sc.d.media[mei.medium];
        const float u = rng.next();
        const bool act_null = u >= div_(ch<NC>(mei.sigma_t, channel), ch<NC>(mei.maj, channel));
        bool act_scatter = !act_null;
        st.last_event_was_null = act_null;
        if (act_scatter) { depth += 1; st.last_p = mei.p; }
        const bool sample_emitters = md.sample_emitters != 0;
        active = active && depth < maxd;
        act_scatter = act_scatter && active;
        specular_chain = specular_chain && !act_scatter;
        specular_chain = specular_chain || (act_scatter && !sample_emitters);
        if (act_null) {
            cnt.n_null++;
            float r[NC];
#pragma unroll
            for (int c = 0; c < NC; ++c) r[c] = div_(mei.sigma_n[c], mei.maj[c]);
            if (is_spectral) { mis_update<NC>(W, r, mei.sigma_n, channel); mis_update<NC>(Wnee, 1.f, mei.sigma_n, channel); }
            else { mis_update<NC>(W, mei.sigma_n, mei.sigma_n, channel); mis_update<NC>(Wnee, 1.f, r, channel); }
            ray.o = mei.p;
            si.t = si.t - mei.t;
        }
        if (act_scatter) {
            cnt.n_real++;
            if (is_spectral) {
                float r[NC];
#pragma unroll
                for (int c = 0; c < NC; ++c) r[c] = div_(mei.sigma_t[c], mei.maj[c]);
                mis_update<NC>(W, r, mei.sigma_s, channel);
            } else mis_update<NC>(W, mei.sigma_t, mei.sigma_s, channel);
            const int phase = md.phase;
            const Frame3 mei_sh = make_frame(ray.d);
            float P[NC];
            if (sample_emitters) {
                DirSample<NC> ds;
                float em[NC], Wn[NC * NC], Wu[NC * NC], w[NC];
                sample_emitter_mis<SPEC>(rng, cnt, mei.p, v3(0.f), nullptr, medium, channel, W, Wn, Wu, ds, em);
                const V3 wo = to_local(mei_sh, ds.d);
                const float ppdf = phase_eval<SPEC>(phase, wo, P);
                mis_update<NC>(Wn, 1.f, P, channel);
                mis_update<NC>(Wu, ds.delta ? 0.f : ppdf, P, channel);
                mis_weights<NC>(Wn, Wu, w);
#pragma unroll
                for (int c = 0; c < NC; ++c) R[c] += w[c] * em[c];
            }
#pragma unroll
            for (int i = 0; i < NC * NC; ++i) Wnee[i] = W[i];
            float s1, s2x, s2y;
            if (ltr) { s1 = rng.next(); s2x = rng.next(); s2y = rng.next(); }
            else { s2x = rng.next(); s2y = rng.next(); s1 = rng.next(); }
            (void) s1;
            V3 wo;
            const float ppdf = phase_sample<SPEC>(phase, s2x, s2y, wo, P);
            ray = spawn_ray(mei.p, v3(0.f), to_world(mei_sh, wo));
            needs_intersection = true;
#pragma unroll
            for (int c = 0; c < NC; ++c) P[c] *= ppdf;
            mis_update<NC>(W, ppdf, P, channel);
            mis_update<NC>(Wnee, 1.f, P, channel);
        }
    }

    active_surface = active_surface || escaped;
    if (active_surface && needs_intersection) ray_intersect(ray, si);
    if (active_surface) {
        const bool count_direct = depth == 0u || specular_chain;
        const int emitter = si_emitter(si);
        if (emitter >= 0 && !(depth == 0u && I.hide_emitters)) {
            if (!count_direct) mis_update<NC>(Wnee, pdf_emitter_direction(emitter, st.last_p, si.p, si.sh.n, -si.wi), 1.f, channel);
            const mib_emitter &e = sc.d.emitters[emitter];
            float w[NC];
            if (count_direct) mis_weights<NC>(W, w);
            else mis_weights<NC>(W, Wnee, w);
#pragma unroll
            for (int c = 0; c < NC; ++c) R[c] += w[c] * emitter_eval<NC>(e, c, si.wi.z, -si.wi);
        }
    }
    active_surface = active_surface && valid(si);
    if (active_surface) {
        cnt.n_surface++;
        const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
        float V[NC];
        if ((bsdf_flags(bsdf) & kBsdfSmooth) && (depth + 1 < maxd)) {
            DirSample<NC> ds;
            float em[NC], Wn[NC * NC], Wu[NC * NC], w[NC];
            sample_emitter_mis<SPEC>(rng, cnt, si.p, si.n, &si, medium, channel, W, Wn, Wu, ds, em);
            const V3 wo = to_local(si.sh, ds.d);
            bsdf_eval<SPEC>(bsdf, si, wo, V);
            const float bpdf = bsdf_pdf(bsdf, si, wo);
            mis_update<NC>(Wn, 1.f, V, channel);
            mis_update<NC>(Wu, ds.delta ? 0.f : bpdf, V, channel);
            mis_weights<NC>(Wn, Wu, w);
#pragma unroll
            for (int c = 0; c < NC; ++c) R[c] += w[c] * em[c];
        }
        float s1, s2x, s2y;
        if (ltr) { s1 = rng.next(); s2x = rng.next(); s2y = rng.next(); }
        else { s2x = rng.next(); s2y = rng.next(); s1 = rng.next(); }
        BsdfS bs;
        bsdf_sample<SPEC>(bsdf, si, s1, s2x, s2y, true, true, bs, V);
        active_surface = bs.pdf > 0.f;
        if (active_surface) {
            st.eta *= bs.eta;
            ray = spawn_ray(si.p, si.n, to_world(si.sh, bs.wo));
            needs_intersection = true;
        }
        const bool non_null = active_surface && !(bs.sampled & kBsdfNull);
        specular_chain = specular_chain || (non_null && (bs.sampled & kBsdfDelta));
        specular_chain = specular_chain && !(active_surface && (bs.sampled & kBsdfSmooth));
        if (non_null) {
            depth += 1; st.last_p = si.p;
#pragma unroll
            for (int i = 0; i < NC * NC; ++i) Wnee[i] = W[i];
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) V[c] *= bs.pdf;
        if (active_surface) mis_update<NC>(W, bs.pdf, V, channel);
        if (non_null) mis_update<NC>(Wnee, 1.f, V, channel);
        if (active_surface && is_medium_transition(si)) medium = target_medium(si, ray.d);
    }
    active = active && (active_surface || active_medium);
}

} // namespace mib

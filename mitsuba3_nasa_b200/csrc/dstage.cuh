// dstage.cuh -- the forward walks (volpath.cpp:95-450, volpathaos.cpp:61-455) cut into STAGES.
//
// Why: in the reference both the main walk and the shadow (NEE) walk are data-dependent loops,
// the shadow loop is nested in the main one, and every loop iteration branches into one of
// {null collision, real scatter, null surface, diffuse surface, path end}.  One thread per
// path therefore leaves 6-10 of 32 lanes active per warp instruction (profiles/r01_*).
// The pool kernel (dpool.cuh) keeps the paths of a CTA in shared memory and lets every warp
// execute ONE stage for 32 paths that all wait for that stage:
//
//   START --> MAIN --real scatter--> SCATTER --> NEE --> PHASE --> MAIN
//               |  --diffuse hit--------------> NEE --> BSDF  --> MAIN
//               |  --null surface--------------------> BSDF  --> MAIN
//               '--path ended--> START (film accumulate, next lane)
//
// The functions below are those stages on a register copy of the path state.  Arithmetic,
// RNG draw order and every decision are exactly the reference's (dwalk.cuh is the same walk
// as one straight-line loop; the GPU parity tests compare both against the oracle lane by lane).
#pragma once
#include "dwalk.cuh"

namespace mib {

// Shadow walk in flight + the contribution waiting for its result.
template <int SPEC> struct NeeState {
    static constexpr int NC = ST<SPEC>::NC, NR = ST<SPEC>::NR;
    Ray ray;
    SurfIt si;
    V3 emitter_d;
    float tr[NC], ev[NC];
    float total_dist, max_dist, ds_dist;
    int medium;
    bool active, needs_intersection;
    // pending contribution: R += a * em   or   R += (a * em) * w
    float a[NR], w;
    bool mul_w_after;
};

template <int SPEC> struct WalkState {
    static constexpr int NC = ST<SPEC>::NC;
    float T[ST<SPEC>::NT];
    float R[ST<SPEC>::NR];
    Ray ray;
    SurfIt si;
    float eta, last_pdf;
    V3 scatter_p;                 // mei.p of the real scatter being processed
    V3 last_p;                    // position of the last scattering event (MIS pdf of an area light that is hit)
    int medium;
    uint32_t depth, channel;
    bool specular_chain, needs_intersection;
    NeeState<SPEC> nee;
};

enum { EV_NULL = 0, EV_SCATTER = 1, EV_NEE = 2, EV_BSDF = 3, EV_DONE = 4 };

template <int SPEC>
MIB_D void walk_init(Pcg32 &rng, const Ray &ray, WalkState<SPEC> &st) {
    T_init<SPEC>(st.T);
#pragma unroll
    for (int i = 0; i < ST<SPEC>::NR; ++i) st.R[i] = 0.f;
    st.ray = ray;
    st.eta = 1.f;
    st.medium = sc.d.sensor.medium;
    st.specular_chain = !sc.d.integrator.hide_emitters;
    st.needs_intersection = true;
    st.depth = 0; st.channel = 0;
    if (SPEC == MIB_SPEC_RGB) st.channel = (uint32_t) fminf(rng.next() * 3.f, 2.f);
    st.si.t = inf_f(); st.si.shape = -1; st.si.face = 0; st.si.p = v3(0.f); st.si.wi = v3(0.f);
    st.last_pdf = 1.f;
    st.last_p = v3(0.f);
    st.nee.active = false;
}

// ---- shadow walk: setup (volpath.cpp:334-372 / volpathaos.cpp:299-345) ----
template <int SPEC, bool AOS>
MIB_D void nee_begin(Pcg32 &rng, V3 ref_p, V3 ref_n, bool surface_ref, const SurfIt *ref_si, int medium,
                     NeeState<SPEC> &ns, DirSample<ST<SPEC>::NC> &ds) {
    constexpr int NC = ST<SPEC>::NC;
#pragma unroll
    for (int c = 0; c < NC; ++c) ns.tr[c] = 1.f;
    float sx = rng.next(), sy = rng.next();
    sample_emitter_direction<NC>(ref_p, sx, sy, ds, ns.ev);
    ns.emitter_d = AOS ? normalize(ds.d) : ds.d;
    ns.active = true;
    ns.medium = medium;
    ns.ds_dist = ds.dist;
    ns.total_dist = 0.f;
    ns.max_dist = 0.f;
    ns.si.t = inf_f(); ns.si.shape = -1; ns.si.face = 0; ns.si.p = v3(0.f);
    ns.needs_intersection = true;
    ns.ray.o = ref_p; ns.ray.d = ds.d; ns.ray.maxt = 0.f;
    if (ds.pdf == 0.f) {
#pragma unroll
        for (int c = 0; c < NC; ++c) ns.ev[c] = 0.f;
        ns.active = false;
        return;
    }
    if (AOS) { ns.ray = spawn_ray(ref_p, ref_n, ds.d); ns.max_dist = ds.dist * (1.f - kShadowEps); }
    else {
        ns.ray = spawn_ray_to(ref_p, ref_n, ds.p);
        ns.max_dist = ns.ray.maxt;
        if (surface_ref && is_medium_transition(*ref_si)) ns.medium = target_medium(*ref_si, ns.ray.d);
    }
}

// ---- shadow walk: one iteration (volpath.cpp:374-447 / volpathaos.cpp:347-452) ----
template <int SPEC, bool AOS>
MIB_D void nee_step(Pcg32 &rng, Counters &cnt, uint32_t channel, NeeState<SPEC> &ns) {
    constexpr int NC = ST<SPEC>::NC;
    Ray &ray = ns.ray;
    SurfIt &si = ns.si;
    float *tr = ns.tr;
    float remaining = ns.max_dist - ns.total_dist;
    ray.maxt = remaining;
    if (!(remaining > 0.f)) { ns.active = false; return; }
    bool escaped_medium = false, active_medium = ns.medium >= 0, active_surface = !active_medium;
    if (active_medium) {
        const mib_medium &md = sc.d.media[ns.medium];
        cnt.n_nee_steps++;
        MedIt<NC> mei;
        sample_interaction<NC>(ns.medium, ray, rng.next(), channel, mei, cnt);
        if (md.type == MIB_MEDIUM_HOMOGENEOUS && valid(mei)) ray.maxt = fminf(mei.t, remaining);
        if (ns.needs_intersection) ray_intersect(ray, si);
        if (si.t < mei.t) mei.t = inf_f();
        ns.needs_intersection = ns.needs_intersection && !valid(si);
        const bool is_spectral = (MIB_FT(FT_SPECEXT) && md.has_spectral_extinction != 0);
        if (is_spectral) {
            float t = fminf(remaining, fminf(mei.t, si.t)) - mei.mint;
            float trv[NC], ffp[NC];
            bool surf = si.t < mei.t || mei.t > remaining;
#pragma unroll
            for (int c = 0; c < NC; ++c) { trv[c] = exp_(-t * mei.maj[c]); ffp[c] = surf ? trv[c] : trv[c] * mei.maj[c]; }
            float tr_pdf = ch<NC>(ffp, channel), ip = rcp_(tr_pdf);
#pragma unroll
            for (int c = 0; c < NC; ++c) tr[c] *= tr_pdf > 0.f ? trv[c] * ip : 0.f;
        }
        if ((mei.t > remaining) && valid(mei)) ns.total_dist = ns.ds_dist;
        if (mei.t > remaining) mei.t = inf_f();
        escaped_medium = !valid(mei);
        active_medium = valid(mei);
        if (active_medium) {
            ns.total_dist += mei.t;
            ray.o = mei.p;
            si.t = si.t - mei.t;
#pragma unroll
            for (int c = 0; c < NC; ++c) tr[c] *= is_spectral ? mei.sigma_n[c] : div_(mei.sigma_n[c], mei.maj[c]);
        }
    }
    bool intersect = active_surface && ns.needs_intersection;
    if (intersect) ray_intersect(ray, si);
    ns.needs_intersection = ns.needs_intersection && !intersect;
    active_surface = active_surface || escaped_medium;
    if (active_surface) ns.total_dist += si.t;
    active_surface = active_surface && valid(si) && !active_medium;
    if (active_surface) {
        const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
        if (bsdf.type == MIB_BSDF_NULL) {
            // eval_null_transmission = identity; its world rotation leaves column 0 untouched
        } else if (AOS) {
            V3 wo = to_local(si.sh, ns.emitter_d);                        // volpathaos.cpp:423-431
            bool ok = si.wi.z > 0.f && wo.z > 0.f;
#pragma unroll
            for (int c = 0; c < NC; ++c) tr[c] *= ok ? bsdf.reflectance[NC == 3 ? c : 0] * kInvPi * wo.z : 0.f;
            ray.d = ns.emitter_d;
        } else {
#pragma unroll
            for (int c = 0; c < NC; ++c) tr[c] *= 0.f;                    // eval_null_transmission of a non-null BSDF
        }
        ray = spawn_ray(si.p, si.n, ray.d);
    }
    ray.maxt = remaining;
    ns.needs_intersection = ns.needs_intersection || active_surface;
    if (MIB_FT(FT_FAST) && sc.super_res > 0 && active_medium) {                          // fast mode: roulette on a small transmittance
        float tm = fabsf(tr[0]);
#pragma unroll
        for (int c = 1; c < NC; ++c) tm = fmaxf(tm, fabsf(tr[c]));
        if (tm < kNeeRouletteThreshold && tm > 0.f) {
            const float q = tm * (1.f / kNeeRouletteThreshold), u = rng.next();
            const float k = u < q ? __fdividef(1.f, q) : 0.f;
#pragma unroll
            for (int c = 0; c < NC; ++c) tr[c] *= k;
        }
    }
    bool any_tr = false;
#pragma unroll
    for (int c = 0; c < NC; ++c) any_tr = any_tr || tr[c] != 0.f;
    ns.active = (active_medium || active_surface) && any_tr;
    if (active_surface && is_medium_transition(si)) ns.medium = target_medium(si, ray.d);
}

// ---- shadow walk finished: add the pending contribution ----
template <int SPEC>
MIB_D void nee_finish(const NeeState<SPEC> &ns, float *R) {
    constexpr int NC = ST<SPEC>::NC, NR = ST<SPEC>::NR;
    float em[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) em[c] = ns.tr[c] * ns.ev[c];
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        float v = ns.a[r] * emis<SPEC>(em, r);
        R[r] += ns.mul_w_after ? v * ns.w : v;
    }
}

// ---------------------------------------------------------------------------------------------
// MAIN: head of one iteration of the reference's while loop (volpath.cpp:121-232 / volpathaos.cpp:85-196)
// up to the point where the event is known, plus the surface-event setup (emission, NEE start).
//   EV_NULL    null collision handled, call again
//   EV_SCATTER real scatter at st.scatter_p; u = albedo factor the SCATTER stage multiplies into T
//   EV_NEE     diffuse surface: st.nee holds a started shadow walk, then BSDF
//   EV_BSDF    null surface: BSDF stage next
//   EV_DONE    path ended
// T_dirty / R_dirty report whether T / R changed.
// ---------------------------------------------------------------------------------------------
template <int SPEC, bool AOS>
MIB_D int stg_main(Pcg32 &rng, Counters &cnt, WalkState<SPEC> &st, float *u_out, bool &T_dirty, bool &R_dirty) {
    constexpr int NC = ST<SPEC>::NC, NR = ST<SPEC>::NR;
    const mib_integrator &I = sc.d.integrator;
    float *T = st.T, *R = st.R;
    Ray &ray = st.ray;
    SurfIt &si = st.si;
    const uint32_t channel = st.channel;
    const uint32_t maxd = (uint32_t) I.max_depth, rrd = (uint32_t) I.rr_depth;

    bool active = T_any_nonzero<SPEC>(T);
    float q = fminf(T_max<SPEC>(T) * sqr_(st.eta), .95f);
    bool perform_rr = st.depth > rrd;
    if (active) { float u = rng.next(); active = (u < q) || !perform_rr; }
    if (perform_rr) { T_mul_f<SPEC>(T, rcp_(q)); T_dirty = true; }
    active = active && st.depth < maxd;
    if (!active) return EV_DONE;

    bool active_medium = st.medium >= 0, active_surface = !active_medium;
    bool act_null = false, act_scatter = false, escaped = false, is_spectral = false;
    MedIt<NC> mei;
    if (active_medium) {
        const mib_medium &md = sc.d.media[st.medium];
        is_spectral = (MIB_FT(FT_SPECEXT) && md.has_spectral_extinction != 0);
        sample_interaction<NC>(st.medium, ray, rng.next(), channel, mei, cnt);
        if (md.type == MIB_MEDIUM_HOMOGENEOUS && valid(mei)) ray.maxt = mei.t;
        if (st.needs_intersection) ray_intersect(ray, si);
        st.needs_intersection = st.needs_intersection && !valid(si);
        if (si.t < mei.t) mei.t = inf_f();
        if (is_spectral) {                                               // medium.cpp:99-111
            float t = fminf(mei.t, si.t) - mei.mint;
            float trv[NC], ffp[NC];
            bool surf = si.t < mei.t;
#pragma unroll
            for (int c = 0; c < NC; ++c) { trv[c] = exp_(-t * mei.maj[c]); ffp[c] = surf ? trv[c] : trv[c] * mei.maj[c]; }
            float tr_pdf = ch<NC>(ffp, channel), ip = rcp_(tr_pdf);
            float u[NC];
#pragma unroll
            for (int c = 0; c < NC; ++c) u[c] = tr_pdf > 0.f ? trv[c] * ip : 0.f;
            T_mul_u<SPEC>(T, u);
            T_dirty = true;
        }
        escaped = !valid(mei);
        active_medium = valid(mei);
        if (active_medium) {
            float u = rng.next();
            bool null_scatter = u >= div_(ch<NC>(mei.sigma_t, channel), ch<NC>(mei.maj, channel));
            act_null = null_scatter;
            act_scatter = !null_scatter;
            if (is_spectral && act_null) {
                float mjc = ch<NC>(mei.maj, channel), isn = rcp_(ch<NC>(mei.sigma_n, channel)), u2[NC];
#pragma unroll
                for (int c = 0; c < NC; ++c) u2[c] = (mei.sigma_n[c] * mjc) * isn;
                T_mul_u<SPEC>(T, u2);
            }
            if (act_scatter) st.depth += 1;
        }
    }
    // `active` of the reference loop from here on; lanes that hit max_depth finish this iteration and stop
    const bool still = st.depth < maxd;
    act_scatter = act_scatter && still;
    if (act_null) { cnt.n_null++; ray.o = mei.p; si.t = si.t - mei.t; }

    if (act_scatter) {
        cnt.n_real++;
        if (is_spectral) {
            float mjc = ch<NC>(mei.maj, channel), ist = rcp_(ch<NC>(mei.sigma_t, channel));
#pragma unroll
            for (int c = 0; c < NC; ++c) u_out[c] = (mei.sigma_s[c] * mjc) * ist;
        } else {
#pragma unroll
            for (int c = 0; c < NC; ++c) u_out[c] = div_(mei.sigma_s[c], mei.sigma_t[c]);
        }
        st.scatter_p = mei.p;
        st.last_p = mei.p;
        return EV_SCATTER;
    }
    if (active_surface || escaped) {
        if (st.needs_intersection) ray_intersect(ray, si);
        if (!AOS) {                                                      // volpath.cpp:258-279
            bool count_direct = st.depth == 0u || st.specular_chain;
            int emitter = si_emitter(si);
            if (emitter >= 0 && !(st.depth == 0u && I.hide_emitters)) {
                float emitter_pdf = 1.0f;
                if (!count_direct) emitter_pdf = pdf_emitter_direction(emitter, st.last_p, si.p, si.sh.n, -si.wi);
                const mib_emitter &e = sc.d.emitters[emitter];
                float x[NR];
                T_col0<SPEC>(T, x);
                float w = mis_weight<SPEC>(st.last_pdf, emitter_pdf);
#pragma unroll
                for (int r = 0; r < NR; ++r) {
                    float er = emitter_eval<NC>(e, r, si.wi.z, -si.wi);
                    R[r] += count_direct ? x[r] * er : (x[r] * w) * er;
                }
                R_dirty = true;
            }
        }
        if (!valid(si)) return EV_DONE;                                  // active &= (active_surface | active_medium)
        cnt.n_surface++;
        const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
        const bool is_null = bsdf.type == MIB_BSDF_NULL;
        const uint32_t flags = is_null ? kBsdfNull : kBsdfDiffuseRefl;
        if ((flags & kBsdfSmooth) && (st.depth + 1 < maxd)) {
            DirSample<NC> ds;
            nee_begin<SPEC, AOS>(rng, si.p, si.n, true, &si, st.medium, st.nee, ds);
            V3 wo = to_local(si.sh, ds.d);
            bool ok = si.wi.z > 0.f && wo.z > 0.f;
            float x[NR];
            T_col0<SPEC>(T, x);
            float bsdf_pdf = ok ? kInvPi * wo.z : 0.f;
            float w = AOS ? 1.f : mis_weight<SPEC>(ds.pdf, ds.delta ? 0.f : bsdf_pdf);
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                float bv = ok ? bsdf.reflectance[NC == 3 ? r : 0] * kInvPi * wo.z : 0.f;
                st.nee.a[r] = AOS ? x[r] * bv : (x[r] * bv) * w;             // (T * depol(bv)) col 0 = T(r,0) * bv
            }
            st.nee.mul_w_after = false;
            st.nee.w = 1.f;
            return EV_NEE;
        }
        return EV_BSDF;
    }
    // null collision (or a scatter cut off by max_depth): next iteration, unless the path just ended
    return (still && active_medium) ? EV_NULL : EV_DONE;
}

// ---------------------------------------------------------------------------------------------
// SCATTER: albedo factor, NEE start + phase evaluation toward the emitter (volpath.cpp:206-234 /
// volpathaos.cpp:170-200).  Returns true when st.nee holds a shadow walk (then PHASE), false -> PHASE directly.
// ---------------------------------------------------------------------------------------------
template <int SPEC, bool AOS>
MIB_D bool stg_scatter(Pcg32 &rng, WalkState<SPEC> &st, const float *u) {
    constexpr int NC = ST<SPEC>::NC, NT = ST<SPEC>::NT;
    constexpr bool POL = ST<SPEC>::POL;
    const mib_medium &md = sc.d.media[st.medium];
    T_mul_u<SPEC>(st.T, u);
    const bool sample_emitters = md.sample_emitters != 0;
    st.specular_chain = !sample_emitters;
    if (!sample_emitters) return false;
    Frame3 sh = make_frame(st.ray.d);
    DirSample<NC> ds;
    nee_begin<SPEC, AOS>(rng, st.scatter_p, v3(0.f), false, nullptr, st.medium, st.nee, ds);
    V3 wo = to_local(sh, ds.d);
    float W[NT];
    float ppdf = phase_eval<SPEC>(md.phase, wo, W);
    if constexpr (POL) m4_to_world(sh, W, -wo, v3(0.f, 0.f, -1.f));
    TW_col0<SPEC>(st.T, W, st.nee.a);
    st.nee.mul_w_after = !AOS;
    st.nee.w = AOS ? 1.f : mis_weight<SPEC>(ds.pdf, ds.delta ? 0.f : ppdf);
    return true;
}

// ---- PHASE: sample the phase function, continue the walk (volpath.cpp:236-256 / volpathaos.cpp:202-218) ----
template <int SPEC, bool AOS>
MIB_D void stg_phase(Pcg32 &rng, WalkState<SPEC> &st) {
    constexpr int NT = ST<SPEC>::NT;
    constexpr bool POL = ST<SPEC>::POL;
    const mib_integrator &I = sc.d.integrator;
    const bool ltr = I.rng_arg_order == MIB_RNG_ARGS_LEFT_TO_RIGHT;
    const mib_medium &md = sc.d.media[st.medium];
    float s1, s2x, s2y;
    if (ltr) { s1 = rng.next(); s2x = rng.next(); s2y = rng.next(); }
    else { s2x = rng.next(); s2y = rng.next(); s1 = rng.next(); }
    (void) s1;
    V3 wo;
    float W[NT];
    Frame3 sh = make_frame(st.ray.d);
    float ppdf = phase_sample<SPEC>(md.phase, s2x, s2y, wo, W);
    if constexpr (POL) m4_to_world(sh, W, -wo, v3(0.f, 0.f, -1.f));
    if (!AOS) T_mul_W<SPEC>(st.T, W);                                    // volpath.cpp:244 (before the pdf test)
    if (ppdf > 0.f) {
        if (AOS || I.double_phase_quirk) T_mul_W<SPEC>(st.T, W);         // volpathaos.cpp:211 / volpath.cpp:251
        st.ray = spawn_ray(st.scatter_p, v3(0.f), to_world(sh, wo));
        st.needs_intersection = true;
        st.last_pdf = ppdf;
    }
}

// ---- BSDF: sample the surface, continue the walk (volpath.cpp:296-326 / volpathaos.cpp:262-292) ----
template <int SPEC, bool AOS>
MIB_D void stg_bsdf(Pcg32 &rng, WalkState<SPEC> &st) {
    constexpr int NC = ST<SPEC>::NC;
    constexpr bool POL = ST<SPEC>::POL;
    const mib_integrator &I = sc.d.integrator;
    const bool ltr = I.rng_arg_order == MIB_RNG_ARGS_LEFT_TO_RIGHT;
    SurfIt &si = st.si;
    const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
    const bool is_null = bsdf.type == MIB_BSDF_NULL;
    float s1, s2x, s2y;
    if (ltr) { s1 = rng.next(); s2x = rng.next(); s2y = rng.next(); }
    else { s2x = rng.next(); s2y = rng.next(); s1 = rng.next(); }
    (void) s1;
    V3 wo_l;
    float bs_pdf;
    uint32_t sampled;
    if (is_null) {
        wo_l = -si.wi; sampled = kBsdfNull; bs_pdf = 1.f;
        if constexpr (POL) {
            float b11, b12, b21, b22;
            null_block(si, b11, b12, b21, b22);
            m4_mul_right_block(st.T, b11, b12, b21, b22);
        }
    } else {
        wo_l = square_to_cosine_hemisphere(s2x, s2y);
        bs_pdf = kInvPi * wo_l.z;
        sampled = kBsdfDiffuseRefl;
        bool ok = si.wi.z > 0.f && bs_pdf > 0.f;
        float v[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) v[c] = ok ? bsdf.reflectance[NC == 3 ? c : 0] : 0.f;
        T_mul_depol<SPEC>(st.T, v);
    }
    st.ray = spawn_ray(si.p, si.n, to_world(si.sh, wo_l));
    st.needs_intersection = true;
    bool non_null = !(sampled & kBsdfNull);
    if (non_null) { st.depth += 1; st.last_pdf = bs_pdf; st.last_p = si.p; }
    st.specular_chain = st.specular_chain || (non_null && (sampled & kBsdfDelta));
    st.specular_chain = st.specular_chain && !(sampled & kBsdfSmooth);
    if (is_medium_transition(si)) st.medium = target_medium(si, st.ray.d);
}

// Whole path through the stage functions in one thread (per-sample dumps of the debug ABI: checks the
// stage decomposition itself, independent of the pool machinery).
template <int SPEC, bool AOS>
MIB_D void trace_path_staged(Pcg32 &rng, Counters &cnt, Ray ray, float *R) {
    constexpr int NC = ST<SPEC>::NC;
    WalkState<SPEC> st;
    walk_init<SPEC>(rng, ray, st);
    while (true) {
        float u[NC];
        bool td = false, rd = false;
        int ev = stg_main<SPEC, AOS>(rng, cnt, st, u, td, rd);
        if (ev == EV_DONE) break;
        if (ev == EV_NULL) continue;
        bool after_bsdf = ev != EV_SCATTER;
        bool nee = ev == EV_NEE;
        if (ev == EV_SCATTER) nee = stg_scatter<SPEC, AOS>(rng, st, u);
        if (nee) {
            while (st.nee.active) nee_step<SPEC, AOS>(rng, cnt, st.channel, st.nee);
            nee_finish<SPEC>(st.nee, st.R);
        }
        if (after_bsdf) stg_bsdf<SPEC, AOS>(rng, st);
        else stg_phase<SPEC, AOS>(rng, st);
    }
#pragma unroll
    for (int i = 0; i < ST<SPEC>::NR; ++i) R[i] = st.R[i];
}

} // namespace mib

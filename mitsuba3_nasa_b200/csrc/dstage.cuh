// dstage.cuh -- the forward walks (volpath.cpp:95-450, volpathaos.cpp:61-455) as a per-thread
// stage machine.
//
// Why: in the reference both the main walk and the shadow (NEE) walk are data-dependent loops
// and the shadow loop is nested in the main one.  Written that way on a GPU, the 32 lanes of a
// warp wait for the longest path AND, at every real scatter of any lane, for the longest shadow
// walk (C3: 7 shadow steps per real scatter on average while 3 of 4 lanes only had a null
// collision).  Here every thread carries an explicit stage and the kernel runs ONE flat loop:
//
//     IDLE --start lane--> MAIN --real scatter--> NEE_MEDIUM --done--> PHASE --> MAIN
//                            |  --surface hit---> NEE_SURFACE -done--> BSDF  --> MAIN
//                            '--path ended--> IDLE (film accumulate, next lane)
//
// One loop iteration executes, in this order, the blocks main-front / shadow-step / phase /
// bsdf, each predicated on the thread's stage; a thread whose shadow walk needs one step runs
// through all blocks in a single iteration (same instruction stream as the nested form), a
// thread with a long shadow walk just stays in NEE for a few iterations while its neighbours
// go on with their own paths.  Arithmetic, RNG draw order and all decisions are unchanged
// (the GPU parity tests compare every lane bit-for-bit against the oracle).
#pragma once
#include "dwalk.cuh"

namespace mib {

enum { STG_MAIN = 0, STG_NEE_MEDIUM = 1, STG_NEE_SURFACE = 2, STG_PHASE = 3, STG_BSDF = 4 };

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
    int medium, scatter_medium;
    uint32_t depth, channel;
    int stage;
    bool active, specular_chain, needs_intersection;
    NeeState<SPEC> nee;
};

template <int SPEC>
MIB_D void walk_init(const DScene &sc, Pcg32 &rng, const Ray &ray, WalkState<SPEC> &st) {
    T_init<SPEC>(st.T);
#pragma unroll
    for (int i = 0; i < ST<SPEC>::NR; ++i) st.R[i] = 0.f;
    st.ray = ray;
    st.eta = 1.f;
    st.medium = sc.d.sensor.medium;
    st.active = true;
    st.stage = STG_MAIN;
    st.specular_chain = !sc.d.integrator.hide_emitters;
    st.needs_intersection = true;
    st.depth = 0; st.channel = 0;
    if (SPEC == MIB_SPEC_RGB) st.channel = (uint32_t) fminf(rng.next() * 3.f, 2.f);
    st.si.t = inf_f(); st.si.shape = -1; st.si.wi = v3(0.f);
    st.last_pdf = 1.f;
    st.nee.active = false;
}

// ---- shadow walk: setup (volpath.cpp:334-372 / volpathaos.cpp:299-345) ----
template <int SPEC, bool AOS>
MIB_D void nee_begin(const DScene &sc, Pcg32 &rng, V3 ref_p, V3 ref_n, bool surface_ref, const SurfIt *ref_si, int medium,
                     NeeState<SPEC> &ns, DirSample<ST<SPEC>::NC> &ds) {
    constexpr int NC = ST<SPEC>::NC;
#pragma unroll
    for (int c = 0; c < NC; ++c) ns.tr[c] = 1.f;
    float sx = rng.next(), sy = rng.next();
    sample_emitter_direction<NC>(sc, ref_p, sx, sy, ds, ns.ev);
    ns.emitter_d = AOS ? normalize(ds.d) : ds.d;
    ns.active = true;
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
        if (surface_ref && is_medium_transition(sc, *ref_si)) medium = target_medium(sc, *ref_si, ns.ray.d);
    }
    ns.medium = medium;
    ns.ds_dist = ds.dist;
    ns.total_dist = 0.f;
    ns.si.t = inf_f(); ns.si.shape = -1;
    ns.needs_intersection = true;
}

// ---- shadow walk: one iteration (volpath.cpp:374-447 / volpathaos.cpp:347-452) ----
template <int SPEC, bool AOS>
MIB_D void nee_step(const DScene &sc, Pcg32 &rng, Counters &cnt, uint32_t channel, NeeState<SPEC> &ns) {
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
        sample_interaction<NC>(sc, ns.medium, ray, rng.next(), channel, mei, cnt);
        if (md.type == MIB_MEDIUM_HOMOGENEOUS && valid(mei)) ray.maxt = fminf(mei.t, remaining);
        if (ns.needs_intersection) ray_intersect(sc, ray, si);
        if (si.t < mei.t) mei.t = inf_f();
        ns.needs_intersection = ns.needs_intersection && !valid(si);
        const bool is_spectral = md.has_spectral_extinction != 0;
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
            for (int c = 0; c < NC; ++c) tr[c] *= is_spectral ? mei.sigma_n[c] : __fdiv_rn(mei.sigma_n[c], mei.maj[c]);
        }
    }
    bool intersect = active_surface && ns.needs_intersection;
    if (intersect) ray_intersect(sc, ray, si);
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
    bool any_tr = false;
#pragma unroll
    for (int c = 0; c < NC; ++c) any_tr = any_tr || tr[c] != 0.f;
    ns.active = (active_medium || active_surface) && any_tr;
    if (active_surface && is_medium_transition(sc, si)) ns.medium = target_medium(sc, si, ray.d);
}

// ---- shadow walk finished: add the pending contribution ----
template <int SPEC>
MIB_D void nee_finish(WalkState<SPEC> &st) {
    constexpr int NC = ST<SPEC>::NC, NR = ST<SPEC>::NR;
    const NeeState<SPEC> &ns = st.nee;
    float em[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) em[c] = ns.tr[c] * ns.ev[c];
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        float v = ns.a[r] * emis<SPEC>(em, r);
        st.R[r] += ns.mul_w_after ? v * ns.w : v;
    }
}

#include "dstage_step.inc"

} // namespace mib

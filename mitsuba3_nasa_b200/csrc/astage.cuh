// astage.cuh -- path-replay adjoint (prbvolpath.py:91-431, prb_volpathaos.py:95-503) cut into stages.
//
// Same decomposition as the forward walk (dstage.cuh), with two differences: a lane runs the walk twice
// (primal pass -> L, then the replay with the same PCG32 stream that scatters gradients), and in the
// replay every shadow walk runs twice as well (first for its value, then -- restarted from the saved
// sampler state -- for the gradients of its transmittance, prbvolpath.py:298-310).
//
//   START --> MAIN --real scatter--> SCATTER --> NEE [--> NEEG] --> PHASE --> MAIN
//     ^         |  --diffuse hit--------------> NEE [--> NEEG] --> BSDF  --> MAIN
//     |         |  --null surface---------------------------------> BSDF  --> MAIN
//     '---------'  path ended: primal pass -> START re-seeds the lane for the replay; replay -> START starts the next lane
//
// prb_step() below runs the stages back to back in one thread (k_adjoint, the prb primal modes of k_forward);
// k_adjoint_pool (apool.cuh) runs them from the stage queues.  Both are compared against the oracle.
#pragma once
#include "adjoint.cuh"
#include "dstage.cuh"

namespace mib {

// Shadow walk of the prb integrators in flight (prbvolpath.py:336-431, prb_volpathaos.py:368-503)
template <int SPEC> struct PrbNee {
    static constexpr int NC = ST<SPEC>::NC;
    Ray ray;
    SurfIt si;
    V3 emitter_d;
    float tr[NC], ev[NC];
    float total_dist, max_dist;
    int medium;
    bool active, needs_intersection;
};

template <int SPEC, bool AOS>
MIB_D void prb_nee_begin(Pcg32 &rng, V3 ref_p, V3 ref_n, const SurfIt *ref_si, int medium, PrbNee<SPEC> &ns, DirSample<ST<SPEC>::NC> &ds) {
    constexpr int NC = ST<SPEC>::NC;
#pragma unroll
    for (int c = 0; c < NC; ++c) ns.tr[c] = 1.f;
    float sx = rng.next(), sy = rng.next();
    sample_emitter_direction<NC>(ref_p, sx, sy, ds, ns.ev);
    ns.active = true;
    if (ds.pdf == 0.f) { for (int c = 0; c < NC; ++c) ns.ev[c] = 0.f; ns.active = false; }
    if (!AOS && ref_si && is_medium_transition(*ref_si)) medium = target_medium(*ref_si, ds.d);
    ns.medium = medium;
    ns.emitter_d = normalize(ds.d);
    ns.ray = spawn_ray(ref_p, ref_n, ds.d);
    ns.total_dist = 0.f;
    ns.si.t = inf_f(); ns.si.shape = -1; ns.si.face = 0; ns.si.p = v3(0.f);
    ns.needs_intersection = true;
    ns.max_dist = ds.dist * (1.f - kShadowEps);
}

// One iteration of the shadow loop.  adj = the NEE contribution this walk produced in the primal pass (per channel; the
// (0,0) entry in polarized mode), dl = adjoint radiance (per channel / Stokes column); both only read when backward.
template <int SPEC, bool AOS>
MIB_D void prb_nee_step(Pcg32 &rng, Counters &cnt, Sink &sink, uint32_t channel, PrbNee<SPEC> &ns, bool backward,
                        const float *adj, const float *dl) {
    constexpr int NC = ST<SPEC>::NC;
    Ray &ray = ns.ray;
    SurfIt &si = ns.si;
    float *tr = ns.tr;
    float &total_dist = ns.total_dist;
    const float max_dist = ns.max_dist;
    int &medium = ns.medium;
    bool &needs_intersection = ns.needs_intersection;
    const V3 emitter_d = ns.emitter_d;
    float remaining = max_dist - total_dist;
    ray.maxt = remaining;
    if (!(remaining > 0.f)) { ns.active = false; return; }
    if (needs_intersection) ray_intersect(ray, si);
    needs_intersection = false;
    bool active_medium = medium >= 0, active_surface = !active_medium;
    float trm[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) trm[c] = 1.f;
    MedIt<NC> mei;
    mei.t = inf_f();
    GridTaps taps;
    bool homog = false, het_step = false;
    float t_h = 0.f, tr_h = 1.f, md_scale = 0.f;
    if (active_medium) {
        const mib_medium &md = sc.d.media[medium];
        md_scale = md.scale;
        const bool is_h = md.type == MIB_MEDIUM_HOMOGENEOUS;
        cnt.n_nee_steps++;
        bool analytic = sc.prb_homog && is_h;
        if (AOS && sc.prb_homog) {
            sample_interaction<NC>(medium, ray, .5f, channel, mei, cnt, &taps);     // no sampler draw (prb_volpathaos.py:422)
        } else {
            sample_interaction<NC>(medium, ray, rng.next(), channel, mei, cnt, &taps);
            if (si.t < mei.t) mei.t = inf_f();
        }
        if (analytic) {
            homog = true;
            mei.t = fminf(remaining, si.t);
            t_h = fminf(mei.t, si.t) - mei.mint;
#pragma unroll
            for (int c = 0; c < NC; ++c) trm[c] = exp_(-t_h * mei.maj[c]);
            tr_h = trm[0];
            mei.t = inf_f();
        }
        het_step = !is_h;
    }
    bool escaped = active_medium && !valid(mei);
    active_medium = active_medium && valid(mei);
    if (active_medium) {
        ray.o = mei.p;
        si.t = si.t - mei.t;
#pragma unroll
        for (int c = 0; c < NC; ++c) trm[c] *= div_(mei.sigma_n[c], mei.maj[c]);
    }
    active_surface = (active_surface || escaped) && valid(si) && !active_medium;
    float cs1 = 1.f, cs2 = 1.f, cs0 = 1.f, cs3 = 1.f;      // column sums of the surface Mueller matrix (AOS backward)
    if (active_surface) {
        const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
        if (bsdf.type == MIB_BSDF_NULL) {
            if (AOS && backward && homog) {
                float b11, b12, b21, b22;
                null_block(si, b11, b12, b21, b22);
                cs1 = b11 + b21; cs2 = b12 + b22;
            }
        } else if (AOS) {
            V3 wo = to_local(si.sh, emitter_d);
            bool ok = si.wi.z > 0.f && wo.z > 0.f;
            float v = ok ? bsdf.reflectance[0] * kInvPi * wo.z : 0.f;
            trm[0] *= v;
            cs0 = v; cs1 = cs2 = cs3 = 0.f;
            ray.d = emitter_d;
        } else {
#pragma unroll
            for (int c = 0; c < NC; ++c) trm[c] *= 0.f;
        }
    }
    if (backward && (active_surface || active_medium)) {
        float cf[NC];
        bool any = false;
        if (AOS) {
            cf[0] = 0.f;
            if (trm[0] > 0.f) {
                if (homog) {
                    float sB = ((cs0 * dl[0] + cs1 * dl[1]) + cs2 * dl[2]) + cs3 * dl[3];
                    cf[0] = ((sB * adj[0]) * (-t_h * md_scale)) * (sc.d.integrator.unbiased_adjoint ? 1.f : tr_h);
                    any = true;
                } else if (active_medium && het_step) {
                    float s = ((dl[0] + dl[1]) + dl[2]) + dl[3];
                    cf[0] = div_((s * adj[0]) * (-md_scale), mei.maj[0]);
                    if (sc.d.integrator.unbiased_adjoint) cf[0] = div_(cf[0], div_(mei.sigma_n[0], mei.maj[0]));   // the step's multiplier
                    any = true;
                }
            }
        } else {
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                float g = trm[c] > 0.f ? div_(dl[c] * adj[c], trm[c]) : 0.f;
                if (homog) cf[c] = (g * (-t_h * md_scale)) * trm[c];
                else if (active_medium) cf[c] = div_(g * (-md_scale), mei.maj[c]);
                else cf[c] = 0.f;
                any = any || cf[c] != 0.f;
            }
        }
        if (any) sink_volume<NC>(sink, sc.d.media[medium].sigma_t, medium, false, taps, cf);
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) tr[c] *= trm[c];
    if (MIB_FT(FT_FAST) && sc.super_res > 0 && active_medium) {       // fast mode: roulette on a small transmittance (dscene.cuh);
        float tm = fabsf(tr[0]);                                      // the gradient walk replays the same decisions
#pragma unroll
        for (int c = 1; c < NC; ++c) tm = fmaxf(tm, fabsf(tr[c]));
        if (tm < kNeeRouletteThreshold && tm > 0.f) {
            const float q = tm * (1.f / kNeeRouletteThreshold), u = rng.next();
            const float k = u < q ? __fdividef(1.f, q) : 0.f;
#pragma unroll
            for (int c = 0; c < NC; ++c) tr[c] *= k;
        }
    }
    if (active_surface) ray = spawn_ray(si.p, si.n, ray.d);
    ray.maxt = remaining;
    needs_intersection = needs_intersection || active_surface;
    bool any_tr = false;
#pragma unroll
    for (int c = 0; c < NC; ++c) any_tr = any_tr || tr[c] != 0.f;
    const bool active = (active_medium || active_surface) && any_tr;
    ns.active = active;
    if (active) total_dist += active_medium ? mei.t : si.t;
    if (active_surface && is_medium_transition(si)) medium = target_medium(si, ray.d);
}

// State of one prb walk (either pass)
template <int SPEC> struct PrbState {
    static constexpr int NR = ST<SPEC>::NR;
    float T[ST<SPEC>::NT];
    float L[NR];
    float dl[4];                  // adjoint radiance of the lane (backward pass)
    Ray ray;
    SurfIt si;
    float eta, dl_sum;
    int medium;
    uint32_t depth, channel;
    bool active, needs_intersection, backward;
    // between stages
    V3 scatter_p;                 // mei.p of the real scatter being processed
    float a[NR];                  // NEE contribution per unit of emitted radiance; after the value walk: the contribution
    Pcg32 nee_rng;                // sampler state at the start of the shadow walk (replayed for its gradients)
    bool nee_surface;
    PrbNee<SPEC> nee;
};

template <int SPEC, bool AOS>
MIB_D void prb_init(Pcg32 &rng, const Ray &ray, bool backward, const float *L_in, const float *dl, PrbState<SPEC> &st) {
    constexpr int NR = ST<SPEC>::NR;
    T_init<SPEC>(st.T);
#pragma unroll
    for (int i = 0; i < NR; ++i) st.L[i] = backward ? L_in[i] : 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) st.dl[i] = backward ? dl[i] : 0.f;
    st.ray = ray;
    st.eta = 1.f;
    st.medium = -1;                                      // "TODO: support sensors inside media" (prbvolpath.py:116)
    st.active = true; st.needs_intersection = true; st.backward = backward;
    st.depth = 0; st.channel = 0;
    if (SPEC == MIB_SPEC_RGB) st.channel = (uint32_t) fminf(rng.next() * 3.f, 2.f);
    st.si.t = inf_f(); st.si.shape = -1; st.si.face = 0; st.si.p = v3(0.f); st.si.wi = v3(0.f);
    st.dl_sum = 0.f;
    if (AOS && backward) st.dl_sum = ((dl[0] + dl[1]) + dl[2]) + dl[3];
    st.nee.active = false;
    st.nee_surface = false;
}

// MAIN: head of one iteration of sample() (prbvolpath.py:121-232 / prb_volpathaos.py:139-260) up to the event, the
// free-flight gradients of the replay, and the surface-event setup.  Events as in dstage.cuh.
template <int SPEC, bool AOS>
MIB_D int prb_main(Pcg32 &rng, Counters &cnt, Sink &sink, PrbState<SPEC> &st) {
    constexpr int NC = ST<SPEC>::NC, NR = ST<SPEC>::NR;
    static_assert(AOS == ST<SPEC>::POL, "prbvolpath is unpolarized, prb_volpathaos is polarized");
    const mib_integrator &I = sc.d.integrator;
    float *T = st.T, *L = st.L;
    const float *dl = st.dl;
    Ray &ray = st.ray;
    SurfIt &si = st.si;
    float &eta = st.eta;
    const float dl_sum = st.dl_sum;
    int &medium = st.medium;
    uint32_t &depth = st.depth;
    const uint32_t channel = st.channel;
    bool &needs_intersection = st.needs_intersection;
    const bool backward = st.backward;
    const uint32_t maxd = (uint32_t) I.max_depth, rrd = (uint32_t) I.rr_depth;
    const bool null_scattering = sc.prb_null != 0;
    bool active = T_any_nonzero<SPEC>(T);
    float q = fminf(T_max<SPEC>(T) * sqr_(eta), .99f);
    bool perform_rr = depth > rrd;
    if (active) { float u = rng.next(); active = (u < q) || !perform_rr; }
    if (perform_rr) T_mul_f<SPEC>(T, rcp_(q));
    if (!active) return EV_DONE;

    bool active_medium = medium >= 0, active_surface = !active_medium;
    MedIt<NC> mei;
    mei.t = inf_f();
    GridTaps taps;
    float weight[NC], A[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) { weight[c] = 1.f; A[c] = 1.f; }
    float t_seg = 0.f;
    if (active_medium) {
        const mib_medium &md = sc.d.media[medium];
        sample_interaction<NC>(medium, ray, rng.next(), channel, mei, cnt, &taps);
        if (md.type == MIB_MEDIUM_HOMOGENEOUS && valid(mei)) ray.maxt = mei.t;
        if (needs_intersection) ray_intersect(ray, si);
        needs_intersection = false;
        if (si.t < mei.t) mei.t = inf_f();
        t_seg = fminf(mei.t, si.t) - mei.mint;
        float trv[NC], pdf[NC];
        const bool surf = si.t < mei.t;
#pragma unroll
        for (int c = 0; c < NC; ++c) { trv[c] = exp_(-t_seg * mei.maj[c]); pdf[c] = surf ? trv[c] : trv[c] * mei.maj[c]; }
        float tr_pdf = ch<NC>(pdf, channel), ip = rcp_(tr_pdf);
#pragma unroll
        for (int c = 0; c < NC; ++c) { A[c] = tr_pdf > 0.f ? trv[c] * ip : 0.f; weight[c] *= A[c]; }
    }
    const bool was_medium = active_medium;
    const bool escaped = active_medium && !valid(mei);
    active_medium = active_medium && valid(mei);
    bool act_null = false, act_scatter = false;
    float scatter_prob = 1.f;
    if (null_scattering) {
        if (active_medium) {
            scatter_prob = div_(ch<NC>(mei.sigma_t, channel), ch<NC>(mei.maj, channel));
            float u = rng.next();
            act_null = u >= scatter_prob;
        }
        act_scatter = !act_null && active_medium;
        if (act_null) {
            float ip = rcp_(1.f - scatter_prob);
#pragma unroll
            for (int c = 0; c < NC; ++c) weight[c] *= mei.sigma_n[c] * ip;
        }
    } else {
        act_scatter = active_medium;
    }
    if (act_scatter) depth += 1;
    active = depth < maxd;
    act_scatter = act_scatter && active;
    if (null_scattering && act_null) { cnt.n_null++; ray.o = mei.p; si.t = si.t - mei.t; }
    if (act_scatter) {
        float ip = rcp_(scatter_prob);
#pragma unroll
        for (int c = 0; c < NC; ++c) weight[c] *= mei.sigma_s[c] * ip;
    }
    if (!AOS || active_medium) T_mul_u<SPEC>(T, weight);

    if (backward && was_medium) {
        const mib_medium &md = sc.d.media[medium];
        float base[NC];
        if (AOS) {
            base[0] = dl_sum * L[0];                                    // sum_all(dL @ L), no division (file header)
            if (sc.d.integrator.unbiased_adjoint) base[0] = div_(base[0], fmaxf(1e-8f, weight[0]));   // extension: prbvolpath's form
        }
        else {
#pragma unroll
            for (int c = 0; c < NC; ++c) base[c] = dl[c] * div_(L[c], fmaxf(1e-8f, weight[c]));
        }
        float cf[NC];
        if (md.type == MIB_MEDIUM_HOMOGENEOUS) {
#pragma unroll
            for (int c = 0; c < NC; ++c) cf[c] = (base[c] * weight[c]) * (-t_seg * md.scale);
            sink_volume<NC>(sink, md.sigma_t, medium, false, taps, cf);
        }
        if (act_null && md.type != MIB_MEDIUM_HOMOGENEOUS) {
            float ip = 1.f - scatter_prob;
#pragma unroll
            for (int c = 0; c < NC; ++c) cf[c] = div_((base[c] * A[c]) * (-md.scale), ip);
            sink_volume<NC>(sink, md.sigma_t, medium, false, taps, cf);
        }
        if (act_scatter) {
            float alb[NC], ca[NC];
            GridTaps ataps;
            Counters scratch = { 0, 0, 0, 0, 0 };
            eval_volume<NC>(md.albedo, mei.p, alb, scratch, &ataps);
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                float k = div_(base[c] * A[c], scatter_prob);
                cf[c] = (k * md.scale) * alb[c];
                ca[c] = k * mei.sigma_t[c];
            }
            sink_volume<NC>(sink, md.sigma_t, medium, false, taps, cf);
            sink_volume<NC>(sink, md.albedo, medium, true, ataps, ca);
        }
    }


    if (act_scatter) { st.scatter_p = mei.p; return EV_SCATTER; }
    bool active_surface2 = active_surface || escaped;
    if (active_surface2 && needs_intersection) ray_intersect(ray, si);
    active_surface2 = active_surface2 && valid(si);
    if (active_surface2) {
        cnt.n_surface++;
        const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
        const bool is_null = bsdf.type == MIB_BSDF_NULL;
        const uint32_t flags = is_null ? kBsdfNull : kBsdfDiffuseRefl;
        if ((flags & kBsdfSmooth) && (depth + 1 < maxd)) {
            st.nee_rng = rng;
            DirSample<NC> ds;
            prb_nee_begin<SPEC, AOS>(rng, si.p, si.n, &si, medium, st.nee, ds);
            V3 wo = to_local(si.sh, ds.d);
            bool ok = si.wi.z > 0.f && wo.z > 0.f;
            if constexpr (ST<SPEC>::POL) {
                float bv = ok ? bsdf.reflectance[0] * kInvPi * wo.z : 0.f;
#pragma unroll
                for (int r = 0; r < NR; ++r) st.a[r] = T[4 * r] * bv;
            } else {
                float bsdf_pdf = ok ? kInvPi * wo.z : 0.f;
                float w = mis_weight<SPEC>(ds.pdf, ds.delta ? 0.f : bsdf_pdf);
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    float bv = ok ? bsdf.reflectance[NC == 3 ? c : 0] * kInvPi * wo.z : 0.f;
                    st.a[c] = (T[c] * bv) * w;
                }
            }
            st.nee_surface = true;
            return EV_NEE;
        }
        return EV_BSDF;
    }
    // null collision, or the walk is over (scatter cut off by max_depth, ray escaped)
    return (active && active_medium) ? EV_NULL : EV_DONE;
}

// SCATTER: NEE start + phase evaluation toward the emitter.  Returns true when st.nee holds a shadow walk.
template <int SPEC, bool AOS>
MIB_D bool prb_scatter(Pcg32 &rng, PrbState<SPEC> &st) {
    constexpr int NC = ST<SPEC>::NC, NT = ST<SPEC>::NT, NR = ST<SPEC>::NR;
    const mib_medium &md = sc.d.media[st.medium];
    if (md.sample_emitters == 0) return false;
    Frame3 sh = make_frame(st.ray.d);
    st.nee_rng = rng;
    DirSample<NC> ds;
    prb_nee_begin<SPEC, AOS>(rng, st.scatter_p, v3(0.f), nullptr, st.medium, st.nee, ds);
    V3 wo = to_local(sh, ds.d);
    float W[NT];
    float ppdf = phase_eval<SPEC>(md.phase, wo, W);
    if constexpr (ST<SPEC>::POL) {
        m4_to_world(sh, W, -wo, v3(0.f, 0.f, -1.f));
        float x[NR];
        TW_col0<SPEC>(st.T, W, x);
#pragma unroll
        for (int r = 0; r < NR; ++r) st.a[r] = x[r];
    } else {
        float w = mis_weight<SPEC>(ds.pdf, ds.delta ? 0.f : ppdf);
#pragma unroll
        for (int c = 0; c < NC; ++c) st.a[c] = (st.T[c] * W[c]) * w;
    }
    st.nee_surface = false;
    return true;
}

// The value walk is over: contribution into L; st.a becomes the contribution (adj of the gradient walk).
// Returns true when the walk has to be replayed for its gradients (st.nee restarted from st.nee_rng; the sampler
// ends the replay in the state it has now, because both walks take the same decisions).
template <int SPEC, bool AOS>
MIB_D bool prb_nee_end(PrbState<SPEC> &st, Pcg32 &grad_rng) {
    constexpr int NC = ST<SPEC>::NC, NR = ST<SPEC>::NR;
    float em[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) em[c] = st.nee.ev[c] * st.nee.tr[c];
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        float contrib = st.a[r] * emis<SPEC>(em, r);
        st.a[r] = contrib;
        st.L[r] = st.backward ? st.L[r] - contrib : st.L[r] + contrib;
    }
    if (!st.backward) return false;
    grad_rng = st.nee_rng;
    DirSample<NC> ds2;
    if (st.nee_surface) prb_nee_begin<SPEC, AOS>(grad_rng, st.si.p, st.si.n, &st.si, st.medium, st.nee, ds2);
    else prb_nee_begin<SPEC, AOS>(grad_rng, st.scatter_p, v3(0.f), nullptr, st.medium, st.nee, ds2);
    return true;
}

// PHASE: sample the phase function, continue the walk
template <int SPEC, bool AOS>
MIB_D void prb_phase(Pcg32 &rng, Counters &cnt, PrbState<SPEC> &st) {
    constexpr int NT = ST<SPEC>::NT;
    const mib_medium &md = sc.d.media[st.medium];
    cnt.n_real++;
    float s1 = rng.next(), s2x = rng.next(), s2y = rng.next();
    (void) s1;
    V3 wo;
    float W[NT];
    Frame3 sh = make_frame(st.ray.d);
    float ppdf = phase_sample<SPEC>(md.phase, s2x, s2y, wo, W);
    if constexpr (ST<SPEC>::POL) m4_to_world(sh, W, -wo, v3(0.f, 0.f, -1.f));
    if (ppdf > 0.f) {
        T_mul_W<SPEC>(st.T, W);
        st.ray = spawn_ray(st.scatter_p, v3(0.f), to_world(sh, wo));
        st.needs_intersection = true;
    }
}

// BSDF: sample the surface.  Returns false when the walk ends here (pdf 0).
template <int SPEC, bool AOS>
MIB_D bool prb_bsdf(Pcg32 &rng, PrbState<SPEC> &st) {
    constexpr int NC = ST<SPEC>::NC;
    SurfIt &si = st.si;
    const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
    const bool is_null = bsdf.type == MIB_BSDF_NULL;
    float s1 = rng.next(), s2x = rng.next(), s2y = rng.next();
    (void) s1;
    V3 wo_l;
    float bs_pdf;
    bool non_null;
    if (is_null) { wo_l = -si.wi; bs_pdf = 1.f; non_null = false; }
    else { wo_l = square_to_cosine_hemisphere(s2x, s2y); bs_pdf = kInvPi * wo_l.z; non_null = true; }
    if (!(bs_pdf > 0.f)) return false;
    if (is_null) {
        if constexpr (ST<SPEC>::POL) {
            float b11, b12, b21, b22;
            null_block(si, b11, b12, b21, b22);
            m4_mul_right_block(st.T, b11, b12, b21, b22);
        }
    } else {
        bool ok = si.wi.z > 0.f && bs_pdf > 0.f;
        float v[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) v[c] = ok ? bsdf.reflectance[NC == 3 ? c : 0] : 0.f;
        T_mul_depol<SPEC>(st.T, v);
    }
    st.ray = spawn_ray(si.p, si.n, to_world(si.sh, wo_l));
    st.needs_intersection = true;
    if (non_null) st.depth += 1;
    if (is_medium_transition(si)) st.medium = target_medium(si, st.ray.d);
    return true;
}

// One iteration of the while loop of sample(): the stages back to back in one thread.
template <int SPEC, bool AOS>
MIB_D void prb_step(Pcg32 &rng, Counters &cnt, Sink &sink, PrbState<SPEC> &st) {
    int ev = prb_main<SPEC, AOS>(rng, cnt, sink, st);
    if (ev == EV_DONE) { st.active = false; return; }
    if (ev == EV_NULL) return;
    bool nee = ev == EV_NEE;
    if (ev == EV_SCATTER) nee = prb_scatter<SPEC, AOS>(rng, st);
    if (nee) {
        while (st.nee.active) prb_nee_step<SPEC, AOS>(rng, cnt, sink, st.channel, st.nee, false, nullptr, nullptr);
        Pcg32 grad_rng;
        if (prb_nee_end<SPEC, AOS>(st, grad_rng)) {
            while (st.nee.active) prb_nee_step<SPEC, AOS>(grad_rng, cnt, sink, st.channel, st.nee, true, st.a, st.dl);
        }
    }
    if (ev == EV_SCATTER) prb_phase<SPEC, AOS>(rng, cnt, st);
    else if (!prb_bsdf<SPEC, AOS>(rng, st)) st.active = false;
}

template <int SPEC, bool AOS>
MIB_D void prb_walk(Pcg32 &rng, Counters &cnt, Ray ray, bool backward, const float *L_in,
                    const float *dl, Sink &sink, float *R) {
    PrbState<SPEC> st;
    prb_init<SPEC, AOS>(rng, ray, backward, L_in, dl, st);
    while (st.active) prb_step<SPEC, AOS>(rng, cnt, sink, st);
#pragma unroll
    for (int i = 0; i < ST<SPEC>::NR; ++i) R[i] = st.L[i];
}

} // namespace mib

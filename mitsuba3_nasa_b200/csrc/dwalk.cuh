// dwalk.cuh -- the forward random walks as straight-line device code.
//   volpath     src/integrators/volpath.cpp:95-450
//   volpathaos  src/integrators/volpathaos.cpp:61-455
// One thread walks one path; all state lives in registers.  Decision points
// (Russian roulette, free flight, real-vs-null, hit-vs-miss) consume the lane's
// PCG32 stream in the reference's order so a lane reproduces the
// reference lane with the same seed.
#pragma once
#include "dbsdf.cuh"

namespace mib {

template <int SPEC> MIB_D float mis_weight(float a, float b) {               // volpath.cpp:463-468
    a *= a; b *= b;
    float w = div_(a, a + b);
    return isfinite(w) ? w : 0.f;
}

// Null-BSDF Mueller matrix in world space: to_world_mueller(I, wi, wi) -> 2x2 block (see dspec.cuh)
MIB_D void null_block(const SurfIt &si, float &b11, float &b12, float &b21, float &b22) {
    // in and out direction coincide (wo = -wi of a null BSDF is evaluated as to_world_mueller(I, wi, wi)): the two basis
    // rotations are the same function of the same arguments, so one evaluation serves both (identical floats)
    const float2 r = world_rotation1(si.sh, si.wi);
    const float ci = r.x, s_i = r.y, co = r.x, so = r.y;
    // X = R_out * I, Y = X * R_in^T
    float x11 = co, x12 = so, x21 = -so, x22 = co;
    b11 = fma_(x12, s_i, x11 * ci); b12 = fma_(x12, ci, x11 * (-s_i));
    b21 = fma_(x22, s_i, x21 * ci); b22 = fma_(x22, ci, x21 * (-s_i));
}

// ---------------------------------------------------------------------------
// Shadow walk (ratio tracking toward the emitter).  AOS = volpathaos.cpp:299-455,
// else volpath.cpp:334-450.  Returns the per-channel emitted radiance * transmittance.
// The transmittance only needs its (0,0) entry in polarized mode (dspec.cuh notes).
// ---------------------------------------------------------------------------
template <int SPEC, bool AOS>
MIB_D void sample_emitter(Pcg32 &rng, Counters &cnt, V3 ref_p, V3 ref_n, bool surface_ref,
                          const SurfIt *ref_si, int medium, uint32_t channel, DirSample<ST<SPEC>::NC> &ds, float *emitted) {
    constexpr int NC = ST<SPEC>::NC;
    float tr[NC], ev[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) tr[c] = 1.f;
    float sx = rng.next(), sy = rng.next();
    sample_emitter_direction<NC>(ref_p, sx, sy, ds, ev);
    V3 emitter_d = AOS ? normalize(ds.d) : ds.d;
    bool active = true;
    if (ds.pdf == 0.f) { for (int c = 0; c < NC; ++c) ev[c] = 0.f; active = false; }
    if (!active) { for (int c = 0; c < NC; ++c) emitted[c] = ev[c]; return; }

    Ray ray;
    float max_dist;
    if (AOS) { ray = spawn_ray(ref_p, ref_n, ds.d); max_dist = ds.dist * (1.f - kShadowEps); }
    else {
        ray = spawn_ray_to(ref_p, ref_n, ds.p);
        max_dist = ray.maxt;
        if (surface_ref && is_medium_transition(*ref_si)) medium = target_medium(*ref_si, ray.d);
    }
    float total_dist = 0.f;
    SurfIt si;
    si.t = inf_f(); si.shape = -1;
    bool needs_intersection = true;
    while (active) {
        float remaining = max_dist - total_dist;
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
            bool is_spectral = (MIB_FT(FT_SPECEXT) && md.has_spectral_extinction != 0), not_spectral = !is_spectral;
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
            if ((mei.t > remaining) && valid(mei)) total_dist = ds.dist;
            if (mei.t > remaining) mei.t = inf_f();
            escaped_medium = !valid(mei);
            active_medium = valid(mei);
            if (active_medium) {
                total_dist += mei.t;
                ray.o = mei.p;
                si.t = si.t - mei.t;
#pragma unroll
                for (int c = 0; c < NC; ++c) tr[c] *= is_spectral ? mei.sigma_n[c] : div_(mei.sigma_n[c], mei.maj[c]);
            }
            (void) not_spectral;
        }
        bool intersect = active_surface && needs_intersection;
        if (intersect) ray_intersect(ray, si);
        needs_intersection = needs_intersection && !intersect;
        active_surface = active_surface || escaped_medium;
        if (active_surface) total_dist += si.t;
        active_surface = active_surface && valid(si) && !active_medium;
        if (active_surface) {
            const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
            if (bsdf.type == MIB_BSDF_NULL) {
                // eval_null_transmission = identity; its world rotation leaves column 0 untouched
            } else if (AOS) {
                // volpathaos.cpp:423-431: bsdf->eval(ctx, si, to_local(emitter_d))
                V3 wo = to_local(si.sh, emitter_d);
                bool ok = si.wi.z > 0.f && wo.z > 0.f;
#pragma unroll
                for (int c = 0; c < NC; ++c) tr[c] *= ok ? bsdf_reflectance(bsdf, si, NC == 3 ? c : 0) * kInvPi * wo.z : 0.f;
                ray.d = emitter_d;
            } else {
                const float nt = MIB_FT(FT_MESH) ? bsdf_null_transmission(bsdf, si) : 0.f;      // thindielectric: 1 - r'
#pragma unroll
                for (int c = 0; c < NC; ++c) tr[c] *= nt;           // eval_null_transmission of a non-null BSDF
            }
            ray = spawn_ray(si.p, si.n, ray.d);
        }
        ray.maxt = remaining;
        needs_intersection = needs_intersection || active_surface;
        if (MIB_FT(FT_FAST) && sc.super_res > 0 && active_medium) {                      // fast mode: roulette on a small transmittance (dscene.cuh)
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
        active = active && (active_medium || active_surface) && any_tr;
        if (active_surface && is_medium_transition(si)) medium = target_medium(si, ray.d);
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) emitted[c] = tr[c] * ev[c];
}

// ---------------------------------------------------------------------------
// One camera path as explicit state + one-iteration step, so that the kernel can run a
// flat loop and start the thread's next lane the moment a path ends (no lane of a warp
// waits for the longest path of the warp).  R receives the radiance (Stokes 4-vector in
// polarized mode).
// ---------------------------------------------------------------------------
template <int SPEC> struct PathState {
    float T[ST<SPEC>::NT];
    float R[ST<SPEC>::NR];
    Ray ray;
    SurfIt si;
    float eta, last_pdf;
    V3 last_p;                    // position of the last scattering event (MIS pdf of an area light that is hit)
    int medium;
    int refractive_bsdf;          // docean.cuh: BSDF of the interface while the path is below it (volpathaos.cpp:80), else -1
    uint32_t depth, channel;
    bool active, specular_chain, needs_intersection;
};

template <int SPEC>
MIB_D void path_init(Pcg32 &rng, const Ray &ray, PathState<SPEC> &st) {
    constexpr int NR = ST<SPEC>::NR;
    T_init<SPEC>(st.T);
#pragma unroll
    for (int i = 0; i < NR; ++i) st.R[i] = 0.f;
    st.ray = ray;
    st.eta = 1.f;
    st.medium = sc.d.sensor.medium;
    st.refractive_bsdf = -1;
    st.active = true;
    st.specular_chain = !sc.d.integrator.hide_emitters;
    st.needs_intersection = true;
    st.depth = 0; st.channel = 0;
    if (SPEC == MIB_SPEC_RGB) st.channel = (uint32_t) fminf(rng.next() * 3.f, 2.f);
    st.si.t = inf_f(); st.si.shape = -1; st.si.wi = v3(0.f);
    st.last_pdf = 1.f;
    st.last_p = v3(0.f);
}

// One iteration of the reference's while loop (volpath.cpp:121-328 / volpathaos.cpp:85-294).
template <int SPEC, bool AOS>
MIB_D void path_step(Pcg32 &rng, Counters &cnt, PathState<SPEC> &st) {
    constexpr int NC = ST<SPEC>::NC, NT = ST<SPEC>::NT, NR = ST<SPEC>::NR;
    constexpr bool POL = ST<SPEC>::POL;
    const mib_integrator &I = sc.d.integrator;
    float *T = st.T, *R = st.R;
    Ray &ray = st.ray;
    SurfIt &si = st.si;
    float &eta = st.eta, &last_pdf = st.last_pdf;
    int &medium = st.medium;
    uint32_t &depth = st.depth;
    const uint32_t channel = st.channel;
    bool &active = st.active, &specular_chain = st.specular_chain, &needs_intersection = st.needs_intersection;
    const uint32_t maxd = (uint32_t) I.max_depth, rrd = (uint32_t) I.rr_depth;
    const bool ltr = I.rng_arg_order == MIB_RNG_ARGS_LEFT_TO_RIGHT;
    MedIt<NC> mei;
    {
        active = active && T_any_nonzero<SPEC>(T);
        float q = fminf(T_max<SPEC>(T) * sqr_(eta), .95f);
        bool perform_rr = depth > rrd;
        if (active) { float u = rng.next(); active = (u < q) || !perform_rr; }
        if (perform_rr) T_mul_f<SPEC>(T, rcp_(q));
        active = active && depth < maxd;
        if (!active) return;

        bool active_medium = medium >= 0, active_surface = !active_medium;
        bool act_null = false, act_scatter = false, escaped = false, is_spectral = false, not_spectral = false;
        if (active_medium) {
            const mib_medium &md = sc.d.media[medium];
            is_spectral = (MIB_FT(FT_SPECEXT) && md.has_spectral_extinction != 0);
            not_spectral = !is_spectral;
            sample_interaction<NC>(medium, ray, rng.next(), channel, mei, cnt);
            if (md.type == MIB_MEDIUM_HOMOGENEOUS && valid(mei)) ray.maxt = mei.t;
            if (needs_intersection) ray_intersect(ray, si);
            needs_intersection = needs_intersection && !valid(si);
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
                if (act_scatter) { depth += 1; st.last_p = mei.p; }
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
                float mjc = ch<NC>(mei.maj, channel), ist = rcp_(ch<NC>(mei.sigma_t, channel));
#pragma unroll
                for (int c = 0; c < NC; ++c) u[c] = (mei.sigma_s[c] * mjc) * ist;
            } else {
#pragma unroll
                for (int c = 0; c < NC; ++c) u[c] = div_(mei.sigma_s[c], mei.sigma_t[c]);
            }
            (void) not_spectral;
            T_mul_u<SPEC>(T, u);
            const int phase = md.phase;
            const Frame3 mei_sh = make_frame(ray.d);   // the reference's mei.sh_frame = Frame(ray.d), wi = (0,0,-1) (medium.cpp:60)
            const bool sample_emitters = md.sample_emitters != 0;
            specular_chain = !sample_emitters;
            float W[NT];
            float2 out_rot = make_float2(1.f, 0.f);
            bool have_out_rot = false;
            if (sample_emitters) {
                DirSample<NC> ds;
                float em[NC];
                sample_emitter<SPEC, AOS>(rng, cnt, mei.p, v3(0.f), false, nullptr, medium, channel, ds, em);
                V3 wo = to_local(mei_sh, ds.d);
                float ppdf = phase_eval<SPEC>(phase, wo, W);
                if constexpr (POL) {
                    out_rot = world_rotation1(mei_sh, v3(0.f, 0.f, -1.f));
                    have_out_rot = true;
                    m4_to_world_with(mei_sh, W, -wo, out_rot);
                }
                float x[NR];
                TW_col0<SPEC>(T, W, x);
                if (AOS) {
#pragma unroll
                    for (int r = 0; r < NR; ++r) R[r] += x[r] * emis<SPEC>(em, r);
                } else {
                    float w = mis_weight<SPEC>(ds.pdf, ds.delta ? 0.f : ppdf);
#pragma unroll
                    for (int r = 0; r < NR; ++r) R[r] += (x[r] * emis<SPEC>(em, r)) * w;
                }
            }
            float s1, s2x, s2y;
            if (ltr) { s1 = rng.next(); s2x = rng.next(); s2y = rng.next(); }
            else { s2x = rng.next(); s2y = rng.next(); s1 = rng.next(); }
            (void) s1;
            V3 wo;
            float ppdf = phase_sample<SPEC>(phase, s2x, s2y, wo, W);
            if constexpr (POL) {
                // the rotation of the outgoing side (local (0,0,-1) in the frame of ray.d) does not depend on wo: the one
                // computed for the emitter direction a few lines up is reused (identical floats)
                if (!have_out_rot) out_rot = world_rotation1(mei_sh, v3(0.f, 0.f, -1.f));
                m4_to_world_with(mei_sh, W, -wo, out_rot);
            }
            if (!AOS) T_mul_W<SPEC>(T, W);                                   // volpath.cpp:244 (before the pdf test)
            act_scatter = act_scatter && ppdf > 0.f;
            if (act_scatter) {
                if (AOS || I.double_phase_quirk) T_mul_W<SPEC>(T, W);        // volpathaos.cpp:211 / volpath.cpp:251
                ray = spawn_ray(mei.p, v3(0.f), to_world(mei_sh, wo));
                needs_intersection = true;
                last_pdf = ppdf;
            }
        }

        active_surface = active_surface || escaped;
        if (active_surface && needs_intersection) ray_intersect(ray, si);
        if (!AOS && active_surface) {                                        // volpath.cpp:258-279
            bool count_direct = depth == 0u || specular_chain;
            int emitter = si_emitter(si);
            if (emitter >= 0 && !(depth == 0u && I.hide_emitters)) {
                float emitter_pdf = 1.0f;
                if (!count_direct) emitter_pdf = pdf_emitter_direction(emitter, st.last_p, si.p, si.sh.n, -si.wi);
                const mib_emitter &e = sc.d.emitters[emitter];
                float x[NR];
                T_col0<SPEC>(T, x);
                float w = mis_weight<SPEC>(last_pdf, emitter_pdf);
#pragma unroll
                for (int r = 0; r < NR; ++r) {
                    float er = emitter_eval<NC>(e, r, si.wi.z, -si.wi);
                    R[r] += count_direct ? x[r] * er : (x[r] * w) * er;
                }
            }
        }
        active_surface = active_surface && valid(si);
        if (active_surface) {
            cnt.n_surface++;
            const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
            if (MIB_FT(FT_MESH) && bsdf.type >= MIB_BSDF_DIELECTRIC) {        // dielectric, roughdielectric, thindielectric, plastic, roughplastic
                // every other BSDF kind through the general interface (dbsdf.cuh): volpath.cpp:281-326 as written
                const uint32_t gflags = bsdf_flags(bsdf);
                float V[NT];
                if ((gflags & kBsdfSmooth) && (depth + 1 < maxd)) {
                    DirSample<NC> ds;
                    float em[NC];
                    sample_emitter<SPEC, AOS>(rng, cnt, si.p, si.n, true, &si, medium, channel, ds, em);
                    const V3 wo = to_local(si.sh, ds.d);
                    bsdf_eval<SPEC>(bsdf, si, wo, V);
                    if constexpr (POL) m4_to_world(si.sh, V, -wo, si.wi);
                    const float w = AOS ? 1.f : mis_weight<SPEC>(ds.pdf, ds.delta ? 0.f : bsdf_pdf(bsdf, si, wo));
                    float x[NR];
                    TW_col0<SPEC>(T, V, x);
#pragma unroll
                    for (int r = 0; r < NR; ++r) R[r] += AOS ? x[r] * emis<SPEC>(em, r) : (x[r] * w) * emis<SPEC>(em, r);
                }
                float g1, g2x, g2y;
                if (ltr) { g1 = rng.next(); g2x = rng.next(); g2y = rng.next(); }
                else { g2x = rng.next(); g2y = rng.next(); g1 = rng.next(); }
                BsdfS bs;
                bsdf_sample<SPEC>(bsdf, si, g1, g2x, g2y, true, true, bs, V);
                if constexpr (POL) m4_to_world(si.sh, V, -bs.wo, si.wi);
                T_mul_W<SPEC>(T, V);
                eta *= bs.eta;
                ray = spawn_ray(si.p, si.n, to_world(si.sh, bs.wo));
                needs_intersection = true;
                const bool gnon_null = !(bs.sampled & kBsdfNull);
                if (gnon_null) { depth += 1; last_pdf = bs.pdf; st.last_p = si.p; }
                specular_chain = specular_chain || (gnon_null && (bs.sampled & kBsdfDelta));
                specular_chain = specular_chain && !(bs.sampled & kBsdfSmooth);
                if (is_medium_transition(si)) medium = target_medium(si, ray.d);
                active = active && (active_surface || active_medium);
                return;
            }
            const bool is_null = bsdf.type == MIB_BSDF_NULL;
            const uint32_t flags = is_null ? kBsdfNull : kBsdfDiffuseRefl;
            bool active_e = (flags & kBsdfSmooth) && (depth + 1 < maxd);
            if (active_e) {
                DirSample<NC> ds;
                float em[NC];
                sample_emitter<SPEC, AOS>(rng, cnt, si.p, si.n, true, &si, medium, channel, ds, em);
                V3 wo = to_local(si.sh, ds.d);
                bool ok = si.wi.z > 0.f && wo.z > 0.f;
                float x[NR];
                T_col0<SPEC>(T, x);
                float bsdf_pdf = ok ? kInvPi * wo.z : 0.f;
                float w = AOS ? 1.f : mis_weight<SPEC>(ds.pdf, ds.delta ? 0.f : bsdf_pdf);
#pragma unroll
                for (int r = 0; r < NR; ++r) {
                    float bv = ok ? bsdf_reflectance(bsdf, si, NC == 3 ? r : 0) * kInvPi * wo.z : 0.f;
                    // (T * depol(bv)) col 0 = T(r,0) * bv
                    if (AOS) R[r] += (x[r] * bv) * emis<SPEC>(em, r);
                    else R[r] += ((x[r] * bv) * w) * emis<SPEC>(em, r);
                }
            }
            float s1, s2x, s2y;
            if (ltr) { s1 = rng.next(); s2x = rng.next(); s2y = rng.next(); }
            else { s2x = rng.next(); s2y = rng.next(); s1 = rng.next(); }
            (void) s1;
            V3 wo_l;
            float bs_eta = 1.f, bs_pdf;
            uint32_t sampled;
            if (is_null) {
                wo_l = -si.wi; sampled = kBsdfNull; bs_pdf = 1.f;
                if constexpr (POL) {
                    float b11, b12, b21, b22;
                    null_block(si, b11, b12, b21, b22);
                    m4_mul_right_block(T, b11, b12, b21, b22);
                }
            } else {
                wo_l = square_to_cosine_hemisphere(s2x, s2y);
                bs_pdf = kInvPi * wo_l.z;
                sampled = kBsdfDiffuseRefl;
                bool ok = si.wi.z > 0.f && bs_pdf > 0.f;
                float v[NC];
#pragma unroll
                for (int c = 0; c < NC; ++c) v[c] = ok ? bsdf_reflectance(bsdf, si, NC == 3 ? c : 0) : 0.f;
                T_mul_depol<SPEC>(T, v);
            }
            eta *= bs_eta;
            ray = spawn_ray(si.p, si.n, to_world(si.sh, wo_l));
            needs_intersection = true;
            bool non_null = !(sampled & kBsdfNull);
            if (non_null) { depth += 1; last_pdf = bs_pdf; st.last_p = si.p; }
            specular_chain = specular_chain || (non_null && (sampled & kBsdfDelta));
            specular_chain = specular_chain && !(sampled & kBsdfSmooth);
            if (is_medium_transition(si)) medium = target_medium(si, ray.d);
        }
        active = active && (active_surface || active_medium);
    }
}

template <int SPEC, bool AOS>
MIB_D void trace_path(Pcg32 &rng, Counters &cnt, Ray ray, float *R) {
    PathState<SPEC> st;
    path_init<SPEC>(rng, ray, st);
    while (st.active) path_step<SPEC, AOS>(rng, cnt, st);
#pragma unroll
    for (int i = 0; i < ST<SPEC>::NR; ++i) R[i] = st.R[i];
}

} // namespace mib

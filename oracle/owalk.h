// owalk.h -- TEST INFRASTRUCTURE (parity oracle). Not part of the product.
// Scalar restatement of the forward random walks:
//   VolumetricPathIntegrator::sample / sample_emitter    src/integrators/volpath.cpp:95-450
//   VolumetricPathAOSIntegrator::sample / sample_emitter src/integrators/volpathaos.cpp:61-455
// with the masked-lane semantics of the llvm_ad_* variants (a PCG32 stream only
// advances when its lane's mask is set, ext/drjit/include/drjit/random.h:78-88).
#pragma once
#include "obsdf.h"

namespace orc {

struct Counters {
    uint64_t n_samples = 0, n_lookups = 0, n_real = 0, n_null = 0, n_nee_steps = 0, n_surface = 0;
    void add(const Counters &o) { n_samples += o.n_samples; n_lookups += o.n_lookups; n_real += o.n_real;
        n_null += o.n_null; n_nee_steps += o.n_nee_steps; n_surface += o.n_surface; }
};

// ORC_TRACE=1 in the environment: sample_volpath prints its loop iterations and BSDF samples to stderr (debugging aid; a traced
// copy of the reference's volpath.cpp prints the same lines, which is how the integrator-level differences were found)
static const bool g_trace = getenv("ORC_TRACE") != nullptr;

template <int SPEC> struct Walk {
    using Tr = SpecTraits<SPEC>;
    using Spec = typename Tr::Spec;
    using USpec = typename Tr::USpec;
    static constexpr int NC = Tr::NC;
    using MI = MediumInteraction<NC>;
    using DS = DirectionSample<NC>;
    using Ph = Phase<SPEC>;
    using Bs = BSDF<SPEC>;

    const Scene &sc;
    Sampler &sampler;
    Counters &cnt;
    Walk(const Scene &s, Sampler &sm, Counters &c) : sc(s), sampler(sm), cnt(c) {}

    static Spec zero_spec() { if constexpr (Tr::POL) return Mueller(); else return Spec(0.f); }
    static Spec depol(USpec u) { return Depol<Spec, NC>::make(u); }
    static Spec mul_f(const Spec &a, float f) { return a * f; }

    Spec mei_to_world_mueller(const MI &mei, const Spec &M, Vec3 in_l, Vec3 out_l) const {
        if constexpr (Tr::POL) return to_world_mueller(mei.sh_frame, M, in_l, out_l); else return M;
    }
    Spec si_to_world_mueller(const SurfaceInteraction &si, const Spec &M, Vec3 in_l, Vec3 out_l) const {
        if constexpr (Tr::POL) return to_world_mueller(si.sh_frame, M, in_l, out_l); else return M;
    }

    // volpath.cpp:463-468
    static float mis_weight(float pdf_a, float pdf_b) {
        pdf_a *= pdf_a; pdf_b *= pdf_b;
        float w = pdf_a / (pdf_a + pdf_b);
        return std::isfinite(w) ? w : 0.f;
    }

    uint32_t max_depth() const { return (uint32_t) sc.d.integrator.max_depth; } // -1 -> 2^32-1 (integrator.cpp:1290)

    // phase->sample(ctx, mei, next_1d(), next_2d()) with the configured C++ argument evaluation order
    void draw_1d_2d(float *s1, float *s2x, float *s2y) {
        if (sc.d.integrator.rng_arg_order == MIB_RNG_ARGS_LEFT_TO_RIGHT) {
            *s1 = sampler.next_1d(); sampler.next_2d(s2x, s2y);
        } else {
            sampler.next_2d(s2x, s2y); *s1 = sampler.next_1d();
        }
    }

    // ------------------------------------------------------------------
    // volpath.cpp:334-450.  ref is (p, n) of the reference interaction; surface_ref
    // enables the medium switch at line 354-356.
    // ------------------------------------------------------------------
    Spec sample_emitter_volpath(Vec3 ref_p, Vec3 ref_n, const SurfaceInteraction *ref_si, int medium,
                                uint32_t channel, DS *ds_out) {
        Spec transmittance = Spec(1.0f);
        float sx, sy; sampler.next_2d(&sx, &sy);
        DS ds;
        USpec ev_u = sample_emitter_direction<NC>(sc, ref_p, sx, sy, &ds);
        Spec emitter_val = depol(ev_u);
        bool active = true;
        if (ds.pdf == 0.f) emitter_val = zero_spec();
        active = active && ds.pdf != 0.f;
        *ds_out = ds;
        if (!active) return emitter_val;

        Ray ray = spawn_ray_to(ref_p, ref_n, ds.p);
        float max_dist = ray.maxt;
        if (ref_si && is_medium_transition(sc, *ref_si))
            medium = target_medium(sc, *ref_si, ray.d);

        float total_dist = 0.f;
        SurfaceInteraction si;
        bool needs_intersection = true;
        while (active) {
            float remaining_dist = max_dist - total_dist;
            ray.maxt = remaining_dist;
            active = active && remaining_dist > 0.f;
            if (!active) break;

            bool escaped_medium = false;
            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            if (active_medium) {
                const mib_medium &md = sc.d.media[medium];
                cnt.n_nee_steps++;
                MI mei = sample_interaction<NC>(sc, medium, ray, sampler.next_1d(), channel, &cnt.n_lookups);
                if (md.type == MIB_MEDIUM_HOMOGENEOUS && mei.is_valid())
                    ray.maxt = minimum(mei.t, remaining_dist);
                bool intersect = needs_intersection;
                if (intersect) si = ray_intersect(sc, ray);
                if (si.t < mei.t) mei.t = Inf;
                needs_intersection = needs_intersection && !si.is_valid();

                bool is_spectral = md.has_spectral_extinction != 0, not_spectral = !is_spectral;
                if (is_spectral) {
                    float t = minimum(remaining_dist, minimum(mei.t, si.t)) - mei.mint;
                    USpec tr = cexp(-t * mei.combined_extinction);
                    USpec free_flight_pdf = (si.t < mei.t || mei.t > remaining_dist) ? tr : tr * mei.combined_extinction;
                    float tr_pdf = index_spectrum(free_flight_pdf, channel);
                    transmittance = smul_u(transmittance, tr_pdf > 0.f ? div_scalar(tr, tr_pdf) : USpec(0.f));
                }
                if ((mei.t > remaining_dist) && mei.is_valid()) total_dist = ds.dist;
                if (mei.t > remaining_dist) mei.t = Inf;

                escaped_medium = !mei.is_valid();
                active_medium = mei.is_valid();
                is_spectral = is_spectral && active_medium;
                not_spectral = not_spectral && active_medium;
                if (active_medium) {
                    total_dist += mei.t;
                    ray.o = mei.p;
                    si.t = si.t - mei.t;
                    if (is_spectral) transmittance = smul_u(transmittance, mei.sigma_n);
                    if (not_spectral) transmittance = smul_u(transmittance, mei.sigma_n / mei.combined_extinction);
                }
            }
            bool intersect = active_surface && needs_intersection;
            if (intersect) si = ray_intersect(sc, ray);
            needs_intersection = needs_intersection && !intersect;
            active_surface = active_surface || escaped_medium;
            if (active_surface) total_dist += si.t;

            active_surface = active_surface && si.is_valid() && active && !active_medium;
            if (active_surface) {
                const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
                Spec bsdf_val = Bs::eval_null_transmission(bsdf, si);
                bsdf_val = si_to_world_mueller(si, bsdf_val, si.wi, si.wi);
                transmittance = smul(transmittance, bsdf_val);
            }
            if (active_surface) ray = spawn_ray(si.p, si.n, ray.d);
            ray.maxt = remaining_dist;
            needs_intersection = needs_intersection || active_surface;

            active = active && (active_medium || active_surface) && any_neq0(unpol(transmittance));
            bool has_medium_trans = active_surface && is_medium_transition(sc, si);
            if (has_medium_trans) medium = target_medium(sc, si, ray.d);
        }
        return smul(transmittance, emitter_val);
    }

    // ------------------------------------------------------------------
    // volpath.cpp:95-330
    // ------------------------------------------------------------------
    Spec sample_volpath(Ray ray, int initial_medium, bool *valid_out) {
        const mib_integrator &I = sc.d.integrator;
        bool valid_ray = !I.hide_emitters && sc.d.environment >= 0;
        float eta = 1.f;
        Spec throughput = Spec(1.f), result = zero_spec();
        int medium = initial_medium;
        MI mei;
        bool active = true;
        bool specular_chain = active && !I.hide_emitters;
        uint32_t depth = 0, channel = 0;
        if (SPEC == MIB_SPEC_RGB) {
            uint32_t n_channels = 3;
            channel = (uint32_t) minimum(sampler.next_1d() * n_channels, (float) (n_channels - 1));
        }
        SurfaceInteraction si;
        bool needs_intersection = true;
        Vec3 last_scatter_p(0.f), last_scatter_n(0.f);
        float last_scatter_direction_pdf = 1.f;
        const uint32_t maxd = max_depth();

        while (active) {
            active = active && any_neq0(unpol(throughput));
            float q = minimum(cmax(unpol(throughput)) * sqr(eta), .95f);
            bool perform_rr = depth > (uint32_t) I.rr_depth;
            if (active) { float u = sampler.next_1d(); active = (u < q) || !perform_rr; }
            if (perform_rr) throughput = mul_f(throughput, rcp(q));
            active = active && depth < maxd;
            if (!active) break;
            if (g_trace) fprintf(stderr, " it depth %u medium %d o %g %g %g d %g %g %g T %g res %g\n", depth, medium, ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z, unpol(throughput).v[0], unpol(result).v[0]);

            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            bool act_null_scatter = false, act_medium_scatter = false, escaped_medium = false;
            bool is_spectral = active_medium, not_spectral = false;
            if (active_medium) {
                is_spectral = is_spectral && sc.d.media[medium].has_spectral_extinction;
                not_spectral = !is_spectral && active_medium;
            }
            if (active_medium) {
                const mib_medium &md = sc.d.media[medium];
                mei = sample_interaction<NC>(sc, medium, ray, sampler.next_1d(), channel, &cnt.n_lookups);
                if (md.type == MIB_MEDIUM_HOMOGENEOUS && mei.is_valid()) ray.maxt = mei.t;
                bool intersect = needs_intersection && active_medium;
                if (intersect) si = ray_intersect(sc, ray);
                needs_intersection = needs_intersection && !(active_medium && si.is_valid());
                if (si.t < mei.t) mei.t = Inf;
                if (is_spectral) {
                    // medium.cpp:99-111 transmittance_eval_pdf
                    float t = minimum(mei.t, si.t) - mei.mint;
                    USpec tr = cexp(-t * mei.combined_extinction);
                    USpec free_flight_pdf = (si.t < mei.t) ? tr : tr * mei.combined_extinction;
                    float tr_pdf = index_spectrum(free_flight_pdf, channel);
                    throughput = smul_u(throughput, tr_pdf > 0.f ? div_scalar(tr, tr_pdf) : USpec(0.f));
                }
                escaped_medium = active_medium && !mei.is_valid();
                active_medium = active_medium && mei.is_valid();
                bool null_scatter = false;
                if (active_medium) {
                    float u = sampler.next_1d();
                    null_scatter = u >= index_spectrum(mei.sigma_t, channel) / index_spectrum(mei.combined_extinction, channel);
                }
                act_null_scatter = null_scatter && active_medium;
                act_medium_scatter = !act_null_scatter && active_medium;
                if (is_spectral && act_null_scatter)
                    throughput = smul_u(throughput, div_scalar(mei.sigma_n * index_spectrum(mei.combined_extinction, channel),
                                                               index_spectrum(mei.sigma_n, channel)));
                if (act_medium_scatter) { depth += 1; last_scatter_p = mei.p; last_scatter_n = Vec3(0.f); }
            }
            active = active && depth < maxd;
            act_medium_scatter = act_medium_scatter && active;
            if (act_null_scatter) { cnt.n_null++; ray.o = mei.p; si.t = si.t - mei.t; }

            if (act_medium_scatter) {
                cnt.n_real++;
                const mib_medium &md = sc.d.media[mei.medium];
                if (is_spectral)
                    throughput = smul_u(throughput, div_scalar(mei.sigma_s * index_spectrum(mei.combined_extinction, channel),
                                                               index_spectrum(mei.sigma_t, channel)));
                if (not_spectral)
                    throughput = smul_u(throughput, mei.sigma_s / mei.sigma_t);
                int phase = md.phase;
                bool sample_emitters = md.sample_emitters != 0;
                valid_ray = true;
                specular_chain = false;
                specular_chain = specular_chain || !sample_emitters;
                if (sample_emitters) {
                    DS ds;
                    Spec emitted = sample_emitter_volpath(mei.p, Vec3(0.f), nullptr, medium, channel, &ds);
                    Vec3 wo = mei.sh_frame.to_local(ds.d);
                    PhaseEval<Spec> pe = Ph::eval_pdf(sc, phase, wo, mei.wi);
                    Spec phase_val = mei_to_world_mueller(mei, pe.val, -wo, mei.wi);
                    result = result + mul_f(smul(smul(throughput, phase_val), emitted),
                                            mis_weight(ds.pdf, ds.delta ? 0.f : pe.pdf));
                }
                float s1, s2x, s2y; draw_1d_2d(&s1, &s2x, &s2y);
                PhaseSample<Spec> ps = Ph::sample(sc, phase, mei.wi, s1, s2x, s2y);
                Spec phase_weight = mei_to_world_mueller(mei, ps.weight, -ps.wo, mei.wi);
                throughput = smul(throughput, phase_weight);
                act_medium_scatter = act_medium_scatter && ps.pdf > 0.f;
                if (act_medium_scatter) {
                    ray = spawn_ray(mei.p, Vec3(0.f), mei.sh_frame.to_world(ps.wo));
                    needs_intersection = true;
                    last_scatter_direction_pdf = ps.pdf;
                    if (I.double_phase_quirk) throughput = smul(throughput, phase_weight); // volpath.cpp:251
                }
            }

            active_surface = active_surface || escaped_medium;
            bool intersect = active_surface && needs_intersection;
            if (intersect) si = ray_intersect(sc, ray);
            if (active_surface) {
                bool ray_from_camera = depth == 0u;
                bool count_direct = ray_from_camera || specular_chain;
                int emitter = si_emitter(sc, si);
                bool active_e = emitter >= 0 && !(depth == 0u && I.hide_emitters);
                if (active_e) {
                    float emitter_pdf = 1.0f;
                    if (!count_direct) emitter_pdf = pdf_emitter_direction<NC>(sc, emitter, last_scatter_p, si.p, si.sh_frame.n, -si.wi);
                    Spec emitted = depol(emitter_eval<NC>(sc, emitter, si.wi.z, -si.wi));
                    Spec contrib = count_direct ? smul(throughput, emitted)
                                                : smul(mul_f(throughput, mis_weight(last_scatter_direction_pdf, emitter_pdf)), emitted);
                    result = result + contrib;
                }
            }
            active_surface = active_surface && si.is_valid();
            if (active_surface) {
                cnt.n_surface++;
                const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
                bool active_e = (bsdf_flags(bsdf) & BSDF_Smooth) && (depth + 1 < maxd);
                if (active_e) {
                    DS ds;
                    Spec emitted = sample_emitter_volpath(si.p, si.n, &si, medium, channel, &ds);
                    Vec3 wo = si.sh_frame.to_local(ds.d);
                    Spec bsdf_val = Bs::eval(bsdf, si, wo);
                    bsdf_val = si_to_world_mueller(si, bsdf_val, -wo, si.wi);
                    float bsdf_pdf = Bs::pdf(bsdf, si, wo);
                    result = result + smul(mul_f(smul(throughput, bsdf_val), mis_weight(ds.pdf, ds.delta ? 0.f : bsdf_pdf)), emitted);
                }
                float s1, s2x, s2y; draw_1d_2d(&s1, &s2x, &s2y);
                BSDFSample<Spec> bs = Bs::sample(bsdf, si, s1, s2x, s2y);
                if (g_trace) fprintf(stderr, "  surf depth %u wi.z %g wo %g %g %g pdf %g type %x val %g (shape %d bsdf %d)\n", depth, si.wi.z, bs.wo.x, bs.wo.y, bs.wo.z, bs.pdf, bs.sampled_type, unpol(bs.value).v[0], si.shape, bsdf.type);
                Spec bsdf_val = si_to_world_mueller(si, bs.value, -bs.wo, si.wi);
                throughput = smul(throughput, bsdf_val);
                eta *= bs.eta;
                ray = spawn_ray(si.p, si.n, si.sh_frame.to_world(bs.wo));
                needs_intersection = true;
                bool non_null_bsdf = !(bs.sampled_type & BSDF_Null);
                if (non_null_bsdf) {
                    depth += 1;
                    last_scatter_p = si.p; last_scatter_n = si.n;
                    last_scatter_direction_pdf = bs.pdf;
                    valid_ray = true;
                }
                specular_chain = specular_chain || (non_null_bsdf && (bs.sampled_type & BSDF_Delta));
                specular_chain = specular_chain && !(bs.sampled_type & BSDF_Smooth);
                act_null_scatter = act_null_scatter || (bs.sampled_type & BSDF_Null);
                if (is_medium_transition(sc, si)) medium = target_medium(sc, si, ray.d);
            }
            active = active && (active_surface || active_medium);
        }
        (void) last_scatter_p; (void) last_scatter_n;
        *valid_out = valid_ray;
        return result;
    }

    // ------------------------------------------------------------------
    // volpathaos.cpp:299-455.  refractive_bsdf: index of the refractive BSDF between this interaction and the sensor
    // (-1: none) -- then the shadow ray leaves in a direction that the interface refracts toward the emitter (:312-337).
    // ------------------------------------------------------------------
    Spec sample_emitter_aos(Vec3 ref_p, Vec3 ref_n, int medium, uint32_t channel, DS *ds_out, int refractive_bsdf = -1) {
        Spec transmittance = Spec(1.0f);
        float sx, sy; sampler.next_2d(&sx, &sy);
        DS ds;
        USpec ev_u = sample_emitter_direction<NC>(sc, ref_p, sx, sy, &ds);
        Spec emitter_val = depol(ev_u);
        Vec3 emitter_d = normalize(ds.d);
        bool active = true;
        if (ds.pdf == 0.f) emitter_val = zero_spec();
        active = active && ds.pdf != 0.f;
        if (active && refractive_bsdf >= 0) {
            // a direction that the interface (normal (0, 0, 1) in world space) refracts into emitter_d: the BSDF is sampled
            // with wi = emitter_d and the transmission lobe only
            SurfaceInteraction rsi;
            rsi.wi = emitter_d;
            float s1, s2x, s2y; draw_1d_2d(&s1, &s2x, &s2y);
            BSDFSample<Spec> bs = Bs::sample(sc.d.bsdfs[refractive_bsdf], rsi, s1, s2x, s2y, 2);
            bool valid_sample = bs.pdf > EpsilonF;                       // dr::Epsilon<Float> (:329)
            if (!valid_sample) emitter_val = zero_spec();
            else { ds.d = -bs.wo; emitter_val = mul_f(emitter_val, rcp(bs.pdf)); }     // Spectrum / Float = * rcp
        }
        *ds_out = ds;
        if (!active) return emitter_val;

        Ray ray = spawn_ray(ref_p, ref_n, ds.d);
        float total_dist = 0.f;
        SurfaceInteraction si;
        bool needs_intersection = true;
        while (active) {
            float remaining_dist = ds.dist * (1.f - ShadowEpsilon) - total_dist;
            ray.maxt = remaining_dist;
            active = active && remaining_dist > 0.f;
            if (!active) break;
            bool escaped_medium = false;
            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            if (active_medium) {
                const mib_medium &md = sc.d.media[medium];
                cnt.n_nee_steps++;
                MI mi = sample_interaction<NC>(sc, medium, ray, sampler.next_1d(), channel, &cnt.n_lookups);
                if (md.type == MIB_MEDIUM_HOMOGENEOUS && mi.is_valid()) ray.maxt = minimum(mi.t, remaining_dist);
                bool intersect = needs_intersection;
                if (intersect) si = ray_intersect(sc, ray);
                if (si.t < mi.t) mi.t = Inf;
                needs_intersection = needs_intersection && !si.is_valid();
                bool is_spectral = md.has_spectral_extinction != 0, not_spectral = !is_spectral;
                if (is_spectral) {
                    float t = minimum(remaining_dist, minimum(mi.t, si.t)) - mi.mint;
                    USpec tr = cexp(-t * mi.combined_extinction);
                    USpec free_flight_pdf = (si.t < mi.t || mi.t > remaining_dist) ? tr : tr * mi.combined_extinction;
                    float tr_pdf = index_spectrum(free_flight_pdf, channel);
                    transmittance = smul_u(transmittance, tr_pdf > 0.f ? div_scalar(tr, tr_pdf) : USpec(0.f));
                }
                if ((mi.t > remaining_dist) && mi.is_valid()) total_dist = ds.dist;
                if (mi.t > remaining_dist) mi.t = Inf;
                escaped_medium = !mi.is_valid();
                active_medium = mi.is_valid();
                is_spectral = is_spectral && active_medium;
                not_spectral = not_spectral && active_medium;
                if (active_medium) {
                    total_dist += mi.t;
                    ray.o = mi.p;
                    si.t = si.t - mi.t;
                    if (is_spectral) transmittance = smul_u(transmittance, mi.sigma_n);
                    if (not_spectral) transmittance = smul_u(transmittance, mi.sigma_n / mi.combined_extinction);
                }
            }
            bool intersect = active_surface && needs_intersection;
            if (intersect) si = ray_intersect(sc, ray);
            needs_intersection = needs_intersection && !intersect;
            active_surface = active_surface || escaped_medium;
            if (active_surface) total_dist += si.t;
            active_surface = active_surface && si.is_valid() && active && !active_medium;
            if (active_surface) {
                const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
                Vec3 wo = si.sh_frame.to_local(emitter_d);
                bool is_null = (bsdf_flags(bsdf) & BSDF_Null) != 0;
                if (is_null) {
                    Spec bsdf_val = Bs::eval_null_transmission(bsdf, si);
                    bsdf_val = si_to_world_mueller(si, bsdf_val, si.wi, si.wi);
                    transmittance = smul(transmittance, bsdf_val);
                } else {
                    Spec bsdf_val = Bs::eval(bsdf, si, wo);
                    bsdf_val = si_to_world_mueller(si, bsdf_val, -wo, si.wi);
                    transmittance = smul(transmittance, bsdf_val);
                    ray.d = emitter_d;
                }
            }
            if (active_surface) ray = spawn_ray(si.p, si.n, ray.d);
            ray.maxt = remaining_dist;
            needs_intersection = needs_intersection || active_surface;
            active = active && (active_medium || active_surface) && any_neq0(unpol(transmittance));
            bool has_medium_trans = active_surface && is_medium_transition(sc, si);
            if (has_medium_trans) medium = target_medium(sc, si, ray.d);
        }
        return smul(transmittance, emitter_val);
    }

    // ------------------------------------------------------------------
    // volpathaos.cpp:61-296
    // ------------------------------------------------------------------
    Spec sample_volpathaos(Ray ray, int initial_medium, bool *valid_out) {
        const mib_integrator &I = sc.d.integrator;
        bool valid_ray = !I.hide_emitters && sc.d.environment >= 0;
        float eta = 1.f;
        Spec throughput = Spec(1.f), result = zero_spec();
        int medium = initial_medium;
        MI mei;
        bool active = true;
        bool specular_chain = active && !I.hide_emitters;
        uint32_t depth = 0, channel = 0;
        if (SPEC == MIB_SPEC_RGB)
            channel = (uint32_t) minimum(sampler.next_1d() * 3, 2.f);
        SurfaceInteraction si;
        bool needs_intersection = true;
        const uint32_t maxd = max_depth();
        int refractive_bsdf = -1;            // volpathaos.cpp:80: the BSDF of the interface while the path is below it

        while (active) {
            active = active && any_neq0(unpol(throughput));
            float q = minimum(cmax(unpol(throughput)) * sqr(eta), .95f);
            bool perform_rr = depth > (uint32_t) I.rr_depth;
            if (active) { float u = sampler.next_1d(); active = (u < q) || !perform_rr; }
            if (perform_rr) throughput = mul_f(throughput, rcp(q));
            active = active && depth < maxd;
            if (!active) break;

            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            bool act_null_scatter = false, act_medium_scatter = false, escaped_medium = false;
            bool is_spectral = active_medium, not_spectral = false;
            if (active_medium) {
                is_spectral = is_spectral && sc.d.media[medium].has_spectral_extinction;
                not_spectral = !is_spectral && active_medium;
            }
            if (active_medium) {
                const mib_medium &md = sc.d.media[medium];
                mei = sample_interaction<NC>(sc, medium, ray, sampler.next_1d(), channel, &cnt.n_lookups);
                if (md.type == MIB_MEDIUM_HOMOGENEOUS && mei.is_valid()) ray.maxt = mei.t;
                bool intersect = needs_intersection && active_medium;
                if (intersect) si = ray_intersect(sc, ray);
                needs_intersection = needs_intersection && !(active_medium && si.is_valid());
                if (si.t < mei.t) mei.t = Inf;
                if (is_spectral) {
                    float t = minimum(mei.t, si.t) - mei.mint;
                    USpec tr = cexp(-t * mei.combined_extinction);
                    USpec free_flight_pdf = (si.t < mei.t) ? tr : tr * mei.combined_extinction;
                    float tr_pdf = index_spectrum(free_flight_pdf, channel);
                    throughput = smul_u(throughput, tr_pdf > 0.f ? div_scalar(tr, tr_pdf) : USpec(0.f));
                }
                escaped_medium = active_medium && !mei.is_valid();
                active_medium = active_medium && mei.is_valid();
                bool null_scatter = false;
                if (active_medium) {
                    float u = sampler.next_1d();
                    null_scatter = u >= index_spectrum(mei.sigma_t, channel) / index_spectrum(mei.combined_extinction, channel);
                }
                act_null_scatter = null_scatter && active_medium;
                act_medium_scatter = !act_null_scatter && active_medium;
                if (is_spectral && act_null_scatter)
                    throughput = smul_u(throughput, div_scalar(mei.sigma_n * index_spectrum(mei.combined_extinction, channel),
                                                               index_spectrum(mei.sigma_n, channel)));
                if (act_medium_scatter) depth += 1;
            }
            active = active && depth < maxd;
            act_medium_scatter = act_medium_scatter && active;
            if (act_null_scatter) { cnt.n_null++; ray.o = mei.p; si.t = si.t - mei.t; }

            if (act_medium_scatter) {
                cnt.n_real++;
                const mib_medium &md = sc.d.media[mei.medium];
                if (is_spectral)
                    throughput = smul_u(throughput, div_scalar(mei.sigma_s * index_spectrum(mei.combined_extinction, channel),
                                                               index_spectrum(mei.sigma_t, channel)));
                if (not_spectral)
                    throughput = smul_u(throughput, mei.sigma_s / mei.sigma_t);
                int phase = md.phase;
                bool sample_emitters = md.sample_emitters != 0;
                valid_ray = true;
                specular_chain = !sample_emitters;
                if (sample_emitters) {
                    DS ds;
                    Spec emitted = sample_emitter_aos(mei.p, Vec3(0.f), medium, channel, &ds, refractive_bsdf);
                    Vec3 wo = mei.sh_frame.to_local(ds.d);
                    PhaseEval<Spec> pe = Ph::eval_pdf(sc, phase, wo, mei.wi);
                    Spec phase_val = mei_to_world_mueller(mei, pe.val, -wo, mei.wi);
                    result = result + smul(smul(throughput, phase_val), emitted);
                }
                float s1, s2x, s2y; draw_1d_2d(&s1, &s2x, &s2y);
                PhaseSample<Spec> ps = Ph::sample(sc, phase, mei.wi, s1, s2x, s2y);
                Spec phase_weight = mei_to_world_mueller(mei, ps.weight, -ps.wo, mei.wi);
                act_medium_scatter = act_medium_scatter && ps.pdf > 0.f;
                if (act_medium_scatter) {
                    throughput = smul(throughput, phase_weight);
                    ray = spawn_ray(mei.p, Vec3(0.f), mei.sh_frame.to_world(ps.wo));
                    needs_intersection = true;
                }
            }

            active_surface = active_surface || escaped_medium;
            bool intersect = active_surface && needs_intersection;
            if (intersect) si = ray_intersect(sc, ray);
            active_surface = active_surface && si.is_valid();
            if (active_surface) {
                cnt.n_surface++;
                const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
                uint32_t flags = bsdf_flags(bsdf);
                bool active_e = (flags & BSDF_Smooth) && (depth + 1 < maxd);
                if (active_e) {
                    bool diffuse = (flags & BSDF_DiffuseReflection) != 0;
                    bool refractive = (flags & BSDF_Transmission) != 0;          // volpathaos.cpp:234-236
                    bool reflect_e = diffuse || (refractive && si.wi.z > 0.f);
                    if (reflect_e) {
                        DS ds;
                        Spec emitted = sample_emitter_aos(si.p, si.n, medium, channel, &ds, refractive_bsdf);
                        Vec3 wo = si.sh_frame.to_local(ds.d);
                        Spec bsdf_val = Bs::eval(bsdf, si, wo);
                        bsdf_val = si_to_world_mueller(si, bsdf_val, -wo, si.wi);
                        result = result + smul(smul(throughput, bsdf_val), emitted);
                    }
                }
                float s1, s2x, s2y; draw_1d_2d(&s1, &s2x, &s2y);
                BSDFSample<Spec> bs = Bs::sample(bsdf, si, s1, s2x, s2y);
                Spec bsdf_val = si_to_world_mueller(si, bs.value, -bs.wo, si.wi);
                throughput = smul(throughput, bsdf_val);
                eta *= bs.eta;
                bool non_null_bsdf = !(bs.sampled_type & BSDF_Null);
                if (non_null_bsdf) depth += 1;
                Ray bsdf_ray = spawn_ray(si.p, si.n, si.sh_frame.to_world(bs.wo));
                // "Refraction occurred, intentionally or not.  Assumes surface is aligned with x-y plane in world space" (:262-267)
                bool refracted = non_null_bsdf && (ray.d.z * bsdf_ray.d.z > 0.f);
                ray = bsdf_ray;
                needs_intersection = true;
                if (refracted) refractive_bsdf = refractive_bsdf < 0 ? sc.d.shapes[si.shape].bsdf : -1;     // :282-285
                valid_ray = valid_ray || non_null_bsdf;
                specular_chain = specular_chain || (non_null_bsdf && (bs.sampled_type & BSDF_Delta));
                specular_chain = specular_chain && !(bs.sampled_type & BSDF_Smooth);
                if (is_medium_transition(sc, si)) medium = target_medium(sc, si, ray.d);
            }
            active = active && (active_surface || active_medium);
        }
        (void) specular_chain;
        *valid_out = valid_ray;
        return result;
    }

#include "owalkmis.inc"
};

} // namespace orc

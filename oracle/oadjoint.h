// oadjoint.h -- TEST INFRASTRUCTURE (parity oracle). Not part of the product.
// Path-replay backpropagation restated from
//   src/python/python/ad/integrators/prbvolpath.py:91-431      (mono / rgb)
//   src/python/python/ad/integrators/prb_volpathaos.py:95-503  (mono_polarized)
// Both the primal (mode = Primal) and the adjoint (mode = Backward) walks.  The
// adjoint replays the path with the same PCG32 stream and accumulates, in double,
// d(loss)/d(parameter) for: sigma_t / albedo grids (through the 8 trilinear taps),
// constant sigma_t and constant albedo of each medium.
#pragma once
#include "owalk.h"
#include <cstdlib>

namespace orc {

struct GradSink {
    double *grid[MIB_MAX_BUFFERS] = {};       // same shape as the data buffers
    double scalar[MIB_MAX_MEDIA][4] = {};     // [0..2] d/d albedo (per channel), [3] d/d sigma_t (constant volume)
    // d(loss) += coeff_c * d(volume value at p, channel c)
    template <int NC> void add_volume(const Scene &, const mib_volume &v, int medium, bool is_albedo, Vec3 p, const double *coeff) {
        if (v.type == MIB_VOLUME_GRID) {
            if (!grid[v.buffer]) return;
            uint32_t idx[8]; float w[8];
            int n = grid_taps(v, p, idx, w);
            for (int k = 0; k < n; ++k) {
                if (v.channels == 1) { double s = 0; for (int c = 0; c < NC; ++c) s += coeff[c]; grid[v.buffer][idx[k]] += s * (double) w[k]; }
                else if (NC == 1) {
                    // a 3-channel grid seen through a 1-channel variant is its luminance (oscene.h eval_volume, grid.cpp:545-575):
                    // d value / d texel channel = the luminance weight
                    const double lum[3] = { 0.212671, 0.715160, 0.072169 };
                    for (int c = 0; c < 3; ++c) grid[v.buffer][idx[k] + c] += coeff[0] * lum[c] * (double) w[k];
                } else for (int c = 0; c < NC; ++c) grid[v.buffer][idx[k] + c] += coeff[c] * (double) w[k];
            }
        } else {
            if (is_albedo) { for (int c = 0; c < NC; ++c) scalar[medium][v.channels == 1 ? 0 : c] += coeff[c]; }
            else { for (int c = 0; c < NC; ++c) scalar[medium][3] += coeff[c]; }
        }
    }
};

template <int SPEC> struct Adjoint {
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
    GradSink *sink = nullptr;
    bool handle_null_scattering = false, nee_handle_homogeneous = false;
    // prb_volpathaos with the division by the step's weight (mib_integrator.unbiased_adjoint, an extension; ORC_PRB_AOS_UNBIASED in the
    // environment forces it): see tests/test_oracle_adjoint.py
    bool unbiased_check = std::getenv("ORC_PRB_AOS_UNBIASED") != nullptr || sc.d.integrator.unbiased_adjoint != 0;

    Adjoint(const Scene &s, Sampler &sm, Counters &c) : sc(s), sampler(sm), cnt(c) {
        // prepare_scene (prbvolpath.py:76-89)
        for (int i = 0; i < sc.d.n_shapes; ++i)
            for (int m : { sc.d.shapes[i].interior_medium, sc.d.shapes[i].exterior_medium })
                if (m >= 0) {
                    bool hom = sc.d.media[m].type == MIB_MEDIUM_HOMOGENEOUS;
                    nee_handle_homogeneous = nee_handle_homogeneous || hom;
                    handle_null_scattering = handle_null_scattering || !hom;
                }
    }
    uint32_t max_depth() const { return (uint32_t) sc.d.integrator.max_depth; }
    static float mis_weight(float a, float b) { a *= a; b *= b; float w = a / (a + b); return std::isfinite(w) ? w : 0.f; }

    // d(sigma_t(p)) and d(albedo(p)) seeds for medium m at p
    void seed_sigma_t(int m, Vec3 p, const double *coeff) { if (sink) sink->add_volume<NC>(sc, sc.d.media[m].sigma_t, m, false, p, coeff); }
    void seed_albedo(int m, Vec3 p, const double *coeff) { if (sink) sink->add_volume<NC>(sc, sc.d.media[m].albedo, m, true, p, coeff); }

    // ------------------------------------------------------------------
    // prbvolpath.py:336-431
    // ------------------------------------------------------------------
    USpec sample_emitter_prb(Vec3 ref_p, Vec3 ref_n, const SurfaceInteraction *ref_si, int medium, uint32_t channel,
                             DS *ds_out, bool backward, const USpec *adj_emitted, const USpec *dL) {
        float sx, sy; sampler.next_2d(&sx, &sy);
        DS ds;
        USpec emitter_val = sample_emitter_direction<NC>(sc, ref_p, sx, sy, &ds);
        bool active = true;
        if (ds.pdf == 0.f) { emitter_val = USpec(0.f); active = false; }
        *ds_out = ds;
        if (!active) medium = -1;
        if (ref_si && is_medium_transition(sc, *ref_si)) medium = target_medium(sc, *ref_si, ds.d);
        Ray ray = spawn_ray(ref_p, ref_n, ds.d);
        float total_dist = 0.f;
        SurfaceInteraction si;
        bool needs_intersection = true;
        USpec transmittance(1.f);
        while (active) {
            float remaining_dist = ds.dist * (1.0f - ShadowEpsilon) - total_dist;
            ray.maxt = remaining_dist;
            active = active && remaining_dist > 0.f;
            needs_intersection = needs_intersection && active;
            if (needs_intersection) si = ray_intersect(sc, ray);
            needs_intersection = false;
            bool active_medium = active && medium >= 0, active_surface = active && !active_medium;
            MI mei;
            USpec tr_multiplier(1.f);
            bool homog = false;
            float t_h = 0.f;
            if (active_medium) {
                cnt.n_nee_steps++;
                mei = sample_interaction<NC>(sc, medium, ray, sampler.next_1d(), channel, &cnt.n_lookups);
                if (si.t < mei.t) mei.t = Inf;
                if (nee_handle_homogeneous && sc.d.media[medium].type == MIB_MEDIUM_HOMOGENEOUS) {
                    homog = true;
                    mei.t = minimum(remaining_dist, si.t);
                    t_h = minimum(mei.t, si.t) - mei.mint;            // medium.cpp:99-111
                    tr_multiplier = cexp(-t_h * mei.combined_extinction);
                    mei.t = Inf;
                }
            }
            bool escaped_medium = active_medium && !mei.is_valid();
            active_medium = active_medium && mei.is_valid();
            if (active_medium) {
                ray.o = mei.p;
                si.t = si.t - mei.t;
                tr_multiplier = tr_multiplier * (mei.sigma_n / mei.combined_extinction);
            }
            active_surface = active_surface || escaped_medium;
            active_surface = active_surface && si.is_valid() && !active_medium;
            if (active_surface) {
                const mib_bsdf &bsdf = sc.d.bsdfs[sc.d.shapes[si.shape].bsdf];
                tr_multiplier = tr_multiplier * (bsdf.type == MIB_BSDF_NULL ? USpec(1.f) : USpec(0.f));
            }
            if (backward && sink && (active_surface || active_medium || homog)) {
                // dr.backward(tr_multiplier * detach(select(active_adj, dL * adj_emitted / tr_multiplier, 0)))
                double coeff[NC];
                bool any = false;
                for (int c = 0; c < NC; ++c) {
                    bool adj = (active_surface || active_medium) && tr_multiplier.v[c] > 0.f;
                    float g = adj ? (dL->v[c] * adj_emitted->v[c]) / tr_multiplier.v[c] : 0.f;
                    coeff[c] = g; any = any || g != 0.f;
                }
                if (any && homog) {
                    // tr = exp(-t * sigma_t * scale): d tr / d sigma_t = -t * scale * tr
                    const mib_medium &md = sc.d.media[medium];
                    double cf[NC];
                    for (int c = 0; c < NC; ++c) cf[c] = coeff[c] * (double) (-t_h * md.scale) * (double) tr_multiplier.v[c];
                    seed_sigma_t(medium, mei.p, cf);
                } else if (any && active_medium) {
                    // sigma_n / maj with maj detached: d = -scale * d sigma_t(p) / maj
                    const mib_medium &md = sc.d.media[medium];
                    double cf[NC];
                    for (int c = 0; c < NC; ++c) cf[c] = coeff[c] * (double) (-md.scale) / (double) mei.combined_extinction.v[c];
                    seed_sigma_t(medium, mei.p, cf);
                }
            }
            transmittance = transmittance * tr_multiplier;
            if (active_surface) ray = spawn_ray(si.p, si.n, ray.d);
            ray.maxt = remaining_dist;
            needs_intersection = needs_intersection || active_surface;
            active = active && (active_medium || active_surface) && any_neq0(transmittance);
            if (active) total_dist += active_medium ? mei.t : si.t;
            if (active_surface && is_medium_transition(sc, si)) medium = target_medium(sc, si, ray.d);
        }
        return emitter_val * transmittance;
    }

    // ------------------------------------------------------------------
    // prbvolpath.py:91-334.  backward == false: returns L.  backward == true: L_in is the
    // primal radiance, dL the adjoint radiance; gradients go to the sink.
    // ------------------------------------------------------------------
    USpec walk_prbvolpath(Ray ray, bool backward, USpec L_in, USpec dL, bool *valid_out) {
        const mib_integrator &I = sc.d.integrator;
        uint32_t depth = 0, channel = 0;
        USpec L = backward ? L_in : USpec(0.f);
        USpec throughput(1.f);
        float eta = 1.f;
        bool active = true, needs_intersection = true, valid_ray = false;
        SurfaceInteraction si;
        int medium = -1;                                           // "TODO: support sensors inside media"
        const uint32_t maxd = max_depth();
        if (SPEC == MIB_SPEC_RGB) channel = (uint32_t) minimum(3 * sampler.next_1d(), 2.f);

        while (active) {
            active = active && any_neq0(throughput);
            float q = minimum(cmax(throughput) * sqr(eta), 0.99f);
            bool perform_rr = depth > (uint32_t) I.rr_depth;
            if (active) { float u = sampler.next_1d(); active = (u < q) || !perform_rr; }
            if (perform_rr) throughput = throughput * rcp(q);
            bool active_medium = active && medium >= 0, active_surface = active && !active_medium;

            MI mei;
            USpec weight(1.f), A(1.f);
            float t_seg = 0.f;
            if (active_medium) {
                const mib_medium &md = sc.d.media[medium];
                mei = sample_interaction<NC>(sc, medium, ray, sampler.next_1d(), channel, &cnt.n_lookups);
                if (md.type == MIB_MEDIUM_HOMOGENEOUS && mei.is_valid()) ray.maxt = mei.t;
                if (needs_intersection) si = ray_intersect(sc, ray);
            }
            needs_intersection = needs_intersection && !active_medium;
            if (active_medium) {
                if (si.t < mei.t) mei.t = Inf;
                t_seg = minimum(mei.t, si.t) - mei.mint;
                USpec tr = cexp(-t_seg * mei.combined_extinction);
                USpec pdf = (si.t < mei.t) ? tr : tr * mei.combined_extinction;
                float tr_pdf = index_spectrum(pdf, channel);
                A = tr_pdf > 0.f ? div_scalar(tr, tr_pdf) : USpec(0.f);
                weight = weight * A;
            }
            bool was_medium = active_medium;
            bool escaped_medium = active_medium && !mei.is_valid();
            active_medium = active_medium && mei.is_valid();
            bool act_null_scatter = false, act_medium_scatter = false;
            float scatter_prob = 1.f;
            if (handle_null_scattering) {
                if (active_medium) {
                    scatter_prob = index_spectrum(mei.sigma_t, channel) / index_spectrum(mei.combined_extinction, channel);
                    float u = sampler.next_1d();
                    act_null_scatter = u >= scatter_prob;
                }
                act_medium_scatter = !act_null_scatter && active_medium;
                if (act_null_scatter) weight = weight * div_scalar(mei.sigma_n, 1.f - scatter_prob);
            } else {
                act_medium_scatter = active_medium;
            }
            if (act_medium_scatter) depth += 1;
            active = active && depth < maxd;
            act_medium_scatter = act_medium_scatter && active;
            if (handle_null_scattering && act_null_scatter) { cnt.n_null++; ray.o = mei.p; si.t = si.t - mei.t; }
            if (act_medium_scatter) weight = weight * div_scalar(mei.sigma_s, scatter_prob);
            throughput = throughput * weight;

            if (backward && sink && was_medium && (active_medium || escaped_medium)) {
                // dr.backward(dL * weight * Lo), Lo = L / max(1e-8, weight)
                const mib_medium &md = sc.d.media[medium];
                double base[NC];
                for (int c = 0; c < NC; ++c) base[c] = (double) dL.v[c] * (double) (L.v[c] / maximum(1e-8f, weight.v[c]));
                // (1) transmittance part (homogeneous media only: the heterogeneous majorant is detached)
                if (md.type == MIB_MEDIUM_HOMOGENEOUS) {
                    double cf[NC];
                    for (int c = 0; c < NC; ++c) cf[c] = base[c] * (double) weight.v[c] * (double) (-t_seg * md.scale);
                    seed_sigma_t(medium, mei.p, cf);
                }
                // (2) event part
                if (act_null_scatter && md.type != MIB_MEDIUM_HOMOGENEOUS) {
                    double cf[NC];
                    for (int c = 0; c < NC; ++c) cf[c] = base[c] * (double) A.v[c] * (double) (-md.scale) / (double) (1.f - scatter_prob);
                    seed_sigma_t(medium, mei.p, cf);
                }
                if (act_medium_scatter) {
                    // sigma_s = scale * value(p) * albedo(p)
                    USpec alb = eval_volume<NC>(sc, md.albedo, mei.p, nullptr);
                    double cs[NC], ca[NC];
                    for (int c = 0; c < NC; ++c) {
                        double k = base[c] * (double) A.v[c] / (double) scatter_prob;
                        cs[c] = k * (double) md.scale * (double) alb.v[c];
                        ca[c] = k * (double) mei.sigma_t.v[c];
                    }
                    seed_sigma_t(medium, mei.p, cs);
                    seed_albedo(medium, mei.p, ca);
                }
            }

            // ---- surface interactions ----
            active_surface = active_surface || escaped_medium;
            if (active_surface && needs_intersection) si = ray_intersect(sc, ray);
            active_surface = active_surface && si.is_valid();
            const mib_bsdf *bsdf = active_surface ? &sc.d.bsdfs[sc.d.shapes[si.shape].bsdf] : nullptr;

            // ---- emitter sampling (use_nee is always true, prbvolpath.py:89) ----
            bool active_e_surface = active_surface && (bsdf_flags(*bsdf) & BSDF_Smooth) && (depth + 1 < maxd);
            bool sample_emitters = act_medium_scatter ? sc.d.media[mei.medium].sample_emitters != 0 : false;
            bool active_e_medium = act_medium_scatter && sample_emitters;
            bool active_e = active_e_surface || active_e_medium;
            int phase = act_medium_scatter ? sc.d.media[mei.medium].phase : -1;
            if (active_e) {
                Sampler nee_sampler = sampler;                      // sampler.clone()
                DS ds;
                Vec3 rp = active_e_medium ? mei.p : si.p, rn = active_e_medium ? Vec3(0.f) : si.n;
                USpec emitted = sample_emitter_prb(rp, rn, active_e_surface ? &si : nullptr, medium, channel, &ds, false, nullptr, nullptr);
                USpec nee_weight;
                float nee_pdf;
                if (active_e_surface) {
                    Vec3 wo = si.sh_frame.to_local(ds.d);
                    nee_weight = Bs::eval(*bsdf, si, wo);
                    nee_pdf = Bs::pdf(*bsdf, si, wo);
                } else {
                    Vec3 wo = mei.sh_frame.to_local(ds.d);
                    auto pe = Ph::eval_pdf(sc, phase, wo, mei.wi);
                    nee_weight = pe.val; nee_pdf = pe.pdf;
                }
                float w = mis_weight(ds.pdf, ds.delta ? 0.f : nee_pdf);
                USpec contrib = throughput * nee_weight * w * emitted;
                L = backward ? L - contrib : L + contrib;
                if (backward) {
                    Sampler saved = sampler;
                    sampler = nee_sampler;
                    DS ds2;
                    sample_emitter_prb(rp, rn, active_e_surface ? &si : nullptr, medium, channel, &ds2, true, &contrib, &dL);
                    sampler = saved;
                }
            }

            // ---- phase function sampling ----
            valid_ray = valid_ray || act_medium_scatter;
            if (act_medium_scatter) {
                cnt.n_real++;
                float s1 = sampler.next_1d(), s2x, s2y; sampler.next_2d(&s2x, &s2y); (void) s1;
                auto ps = Ph::sample(sc, phase, mei.wi, s1, s2x, s2y);
                act_medium_scatter = ps.pdf > 0.f;
                if (act_medium_scatter) {
                    throughput = throughput * ps.weight;
                    ray = spawn_ray(mei.p, Vec3(0.f), mei.sh_frame.to_world(ps.wo));
                    needs_intersection = true;
                }
            }
            // ---- BSDF sampling ----
            if (active_surface) {
                cnt.n_surface++;
                float s1 = sampler.next_1d(), s2x, s2y; sampler.next_2d(&s2x, &s2y);
                auto bs = Bs::sample(*bsdf, si, s1, s2x, s2y);
                active_surface = bs.pdf > 0.f;
                if (active_surface) {
                    throughput = throughput * bs.value;
                    eta *= bs.eta;
                    ray = spawn_ray(si.p, si.n, si.sh_frame.to_world(bs.wo));
                    needs_intersection = true;
                    bool non_null = !(bs.sampled_type & BSDF_Null);
                    if (non_null) { depth += 1; valid_ray = true; }
                    if (is_medium_transition(sc, si)) medium = target_medium(sc, si, ray.d);
                }
            }
            active = active && (active_surface || active_medium);
        }
        *valid_out = valid_ray;
        return L;
    }

    Spec primal_prbvolpath(Ray ray, int, bool *valid) {
        if constexpr (Tr::POL) throw std::runtime_error("prbvolpath: use prb_volpathaos in polarized variants");
        else return walk_prbvolpath(ray, false, USpec(0.f), USpec(0.f), valid);
    }

#include "oadjoint_aos.inc"
};

} // namespace orc

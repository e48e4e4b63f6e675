// adjoint.cuh -- path-replay backpropagation on the device.
//   prbvolpath      src/python/python/ad/integrators/prbvolpath.py:91-431      (mono, rgb)
//   prb_volpathaos  src/python/python/ad/integrators/prb_volpathaos.py:95-503  (mono_polarized)
// The reference records nothing during the primal pass: it runs the walk once to get
// L, then replays it with the same sampler state and differentiates each step locally
// (dr.backward on `weight` / `tr_multiplier`).  Here both passes run back to back in the
// same thread: pass 0 produces L in registers, pass 1 re-seeds the lane's PCG32 and
// scatters d(loss)/d(sigma_t grid), d/d(albedo grid) through the 8 trilinear taps with
// red.global.add.f32, and d/d(constant sigma_t | albedo) into per-thread doubles.
//
// Reference quirks kept on purpose (tests pin them against the oracle):
//  * prb_volpathaos back-propagates through I = replace_grad(1, weight) without
//    dividing by the weight (prb_volpathaos.py:214-217, :483-486).
//  * render_1_backward hands the sensor-frame adjoint Stokes vector to sample()
//    without rotating it back (common.py:1024-1072).
#pragma once
#include "dwalk.cuh"

namespace mib {

struct AdjParams {
    const float *grad_image;   // [H*W*Cd] image gradient (render_backward) or nullptr
    float grad1[4];            // d loss / d (I,Q,U,V)  (render_1_backward)
    int render_1;
};

// Per-thread gradient sink.  Grid gradients go straight to global memory; scalar
// (constant-volume) gradients are kept in registers for one medium at a time.
struct Sink {
    double acc[4];             // [0..2] albedo (per channel), [3] sigma_t
    int medium;
    MIB_D void init() { acc[0] = acc[1] = acc[2] = acc[3] = 0.0; medium = -1; }
    MIB_D void flush() {
        if (medium >= 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) if (acc[k] != 0.0) atomicAdd(sc.grad_scalar + medium * 4 + k, acc[k]);
        }
        acc[0] = acc[1] = acc[2] = acc[3] = 0.0;
    }
    MIB_D void scalar(int m, int k, float v) {
        if (m != medium) { flush(); medium = m; }
        acc[k] += (double) v;
    }
};

// d(loss) += coeff[c] * d(volume value at the taps, channel c)
template <int NC>
MIB_D void sink_volume(Sink &sink, const mib_volume &v, int medium, bool is_albedo, const GridTaps &taps, const float *coeff) {
    if (v.type == MIB_VOLUME_GRID) {
        float *g = sc.grad[v.buffer];
        if (!g) return;
        const int n = v.filter == MIB_FILTER_NEAREST ? 1 : 8;
        if (v.channels == 1) {
            float s = coeff[0];
#pragma unroll
            for (int c = 1; c < NC; ++c) s += coeff[c];
            if (s == 0.f) return;
#ifndef MIB_NO_GRAD_ATOMICS
            for (int k = 0; k < n; ++k) atomicAdd(g + taps.idx[k], s * taps.w[k]);
#endif
        } else {
            for (int k = 0; k < n; ++k)
#pragma unroll
                for (int c = 0; c < NC; ++c) atomicAdd(g + taps.idx[k] + c, coeff[c] * taps.w[k]);
        }
    } else {
#pragma unroll
        for (int c = 0; c < NC; ++c) sink.scalar(medium, is_albedo ? (v.channels == 1 ? 0 : c) : 3, coeff[c]);
    }
}

// ---------------------------------------------------------------------------
// Shadow walk of the prb integrators (prbvolpath.py:336-431, prb_volpathaos.py:368-503).
// adj = the NEE contribution this walk produced in the primal pass (per channel; the (0,0)
// entry in polarized mode), dl = adjoint radiance (per channel / Stokes column).
// ---------------------------------------------------------------------------
template <int SPEC, bool AOS>
MIB_D void prb_sample_emitter(Pcg32 &rng, Counters &cnt, V3 ref_p, V3 ref_n, const SurfIt *ref_si,
                              int medium, uint32_t channel, DirSample<ST<SPEC>::NC> &ds, float *emitted,
                              bool backward, Sink &sink, const float *adj, const float *dl) {
    constexpr int NC = ST<SPEC>::NC;
    float tr[NC], ev[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) tr[c] = 1.f;
    float sx = rng.next(), sy = rng.next();
    sample_emitter_direction<NC>(ref_p, sx, sy, ds, ev);
    bool active = true;
    if (ds.pdf == 0.f) { for (int c = 0; c < NC; ++c) ev[c] = 0.f; active = false; }
    if (!AOS && ref_si && is_medium_transition(*ref_si)) medium = target_medium(*ref_si, ds.d);
    V3 emitter_d = normalize(ds.d);
    Ray ray = spawn_ray(ref_p, ref_n, ds.d);
    float total_dist = 0.f;
    SurfIt si;
    si.t = inf_f(); si.shape = -1;
    bool needs_intersection = true;
    const float max_dist = ds.dist * (1.f - kShadowEps);
    while (active) {
        float remaining = max_dist - total_dist;
        ray.maxt = remaining;
        active = remaining > 0.f;
        if (!active) break;
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
                        cf[0] = ((sB * adj[0]) * (-t_h * md_scale)) * tr_h;
                        any = true;
                    } else if (active_medium && het_step) {
                        float s = ((dl[0] + dl[1]) + dl[2]) + dl[3];
                        cf[0] = div_((s * adj[0]) * (-md_scale), mei.maj[0]);
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
        if (active_surface) ray = spawn_ray(si.p, si.n, ray.d);
        ray.maxt = remaining;
        needs_intersection = needs_intersection || active_surface;
        bool any_tr = false;
#pragma unroll
        for (int c = 0; c < NC; ++c) any_tr = any_tr || tr[c] != 0.f;
        active = (active_medium || active_surface) && any_tr;
        if (active) total_dist += active_medium ? mei.t : si.t;
        if (active_surface && is_medium_transition(si)) medium = target_medium(si, ray.d);
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) emitted[c] = ev[c] * tr[c];
}

#include "adjoint_walk.inc"

} // namespace mib

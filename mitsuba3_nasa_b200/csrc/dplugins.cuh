// dplugins.cuh -- phase functions, emitters, BSDFs, sensors as inlined device code.
// Reference: src/phase/{hg,rayleigh,isotropic,tabphase,tabphase_polarized}.cpp,
// src/emitters/{constant,directional}.cpp, src/bsdfs/{null,diffuse}.cpp,
// src/sensors/{perspective,distant,radiancemeter}.cpp, include/mitsuba/core/warp.h.
#pragma once
#include "dspec.cuh"
#include "denvmap.cuh"

namespace mib {

// ---- warp.h ----
MIB_D V3 square_to_uniform_sphere(float sx, float sy) {                    // warp.h:250-255
    float z = fnma_(2.f, sy, 1.f), r = safe_sqrt_(fnma_(z, z, 1.f)), s, c;
    sincos_(2.f * kPi * sx, s, c);
    return v3(r * c, r * s, z);
}
MIB_D void square_to_uniform_disk_concentric(float sx, float sy, float &ox, float &oy) { // warp.h:54-90
    float x = fms_(2.f, sx, 1.f), y = fms_(2.f, sy, 1.f);
    bool is_zero = x == 0.f && y == 0.f, q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = div_(0.25f * kPi * rp, r);
    if (q13) phi = 0.5f * kPi - phi;
    if (is_zero) phi = 0.f;
    float s, c;
    sincos_(phi, s, c);
    ox = r * c; oy = r * s;
}
MIB_D V3 square_to_cosine_hemisphere(float sx, float sy) {                 // warp.h:412-420
    float px, py;
    square_to_uniform_disk_concentric(sx, sy, px, py);
    return v3(px, py, safe_sqrt_(1.f - fma_(py, py, px * px)));
}

// ---- phase functions: wo local (+z = ray.d), wi = (0,0,-1) ----
MIB_D float eval_hg(float g, float cos_theta) {                             // hg.cpp:66-71
    float temp = 1.f + sqr_(g) + 2.f * g * cos_theta;
    return div_(kInvFourPi * (1.f - sqr_(g)), temp * sqrt_(temp));
}
MIB_D float rayleigh_pdf(float c) { return (3.f / 16.f) * kInvPi * (1.f + sqr_(c)); }
// mueller.h:311-328, scaled matrix
// The depolarization factors of mueller.h:311-328 depend on the phase function's parameter only: commit() evaluates them once
// on the device (k_shape_frames -> sc.phase_const) and the walk reads them back (identical floats, 4 IEEE divisions saved
// per evaluation).
MIB_D void rayleigh_factors_compute(float rho, float &r1, float &r2, float &r3) {
    r1 = div_(1.f - rho, 1.f + div_(rho, 2.f)); r2 = div_(1.f + rho, 1.f - rho); r3 = div_(1.f - 2.f * rho, 1.f - rho);
}
MIB_D void rayleigh_factors(int idx, float &r1, float &r2, float &r3) {
    if (sc.frames_ready) { r1 = sc.phase_const[idx][0]; r2 = sc.phase_const[idx][1]; r3 = sc.phase_const[idx][2]; }
    else rayleigh_factors_compute(sc.d.phases[idx].depolarization, r1, r2, r3);
}
MIB_D void rayleigh_matrix(float cos_theta, float sin_theta, int idx, float *M) {
    float r1, r2, r3;
    rayleigh_factors(idx, r1, r2, r3);
    float a = r2 + sqr_(cos_theta), b = 1.f + sqr_(cos_theta), c = -sqr_(sin_theta), d = 2.f * cos_theta;
    float k = (3.f / 16.f) * kInvPi * r1;
#pragma unroll
    for (int i = 0; i < 16; ++i) M[i] = k * 0.f;
    M[0] = k * a; M[1] = k * c; M[4] = k * c; M[5] = k * b; M[10] = k * d; M[15] = k * (d * r3);
}

// eval_pdf: W = phase value (Mueller in local Stokes frames, or one float replicated), returns pdf
template <int SPEC>
MIB_DH float phase_eval(int idx, V3 wo, float *W) {
    constexpr bool POL = ST<SPEC>::POL;
    constexpr int NC = ST<SPEC>::NC;
    const mib_phase &p = sc.d.phases[idx];
    const V3 wi = v3(0.f, 0.f, -1.f);
    float val, pdf;
    int ptype = p.type;
    if (!MIB_FT(FT_TAB) && (ptype == MIB_PHASE_TABPHASE || ptype == MIB_PHASE_TABPHASE_POLARIZED)) ptype = MIB_PHASE_ISOTROPIC;   // not in this instantiation
    if (!MIB_FT(FT_ANALYTIC) && ptype != MIB_PHASE_TABPHASE) ptype = MIB_PHASE_TABPHASE_POLARIZED;
    switch (ptype) {
        case MIB_PHASE_ISOTROPIC: val = pdf = kInvFourPi; break;                        // isotropic.cpp:51-58
        case MIB_PHASE_HG: val = pdf = eval_hg(p.g, dot(wo, wi)); break;                 // hg.cpp:96-104
        case MIB_PHASE_TABPHASE: {                                                       // tabphase.cpp:121-133
            float ct = -dot(wo, wi);
            val = pdf = tab_eval(sc.tables[idx][0], ct) * sc.tables[idx][0].normalization * kInvTwoPi;
            break;
        }
        case MIB_PHASE_RAYLEIGH: {                                                       // rayleigh.cpp:118-128
            float ct = dot(wo, -wi);
            pdf = rayleigh_pdf(ct);
            if constexpr (POL) {
                rayleigh_matrix(ct, safe_sqrt_(1.0f - sqr_(ct)), idx, W);
                m4_phase_rotate(W, wo, wi);
                return pdf;
            } else {
                float r1, r2, r3;
                rayleigh_factors(idx, r1, r2, r3);
                (void) r3;
                val = (3.f / 16.f) * kInvPi * r1 * (r2 + sqr_(ct));
            }
            break;
        }
        default: {                                                                       // tabphase_polarized.cpp:146-205
            float ct = -dot(wo, wi);
            float m11 = tab_eval(sc.tables[idx][0], ct);
            pdf = m11 * sc.tables[idx][0].normalization * kInvTwoPi;
            if constexpr (POL) {
                float m12 = tab_eval(sc.tables[idx][1], ct), m33 = tab_eval(sc.tables[idx][2], ct), m34 = tab_eval(sc.tables[idx][3], ct);
#pragma unroll
                for (int i = 0; i < 16; ++i) W[i] = 0.f;
                W[0] = m11; W[1] = m12; W[4] = m12; W[5] = m11; W[10] = m33; W[11] = m34; W[14] = -m34; W[15] = m33;
                m4_phase_rotate(W, wo, wi);
                return pdf;
            } else val = m11;
            break;
        }
    }
    if constexpr (POL) m4_identity(W, val);
    else { for (int c = 0; c < NC; ++c) W[c] = val; }
    return pdf;
}

// sample: returns pdf, writes local wo and the weight W
template <int SPEC>
MIB_DH float phase_sample(int idx, float s2x, float s2y, V3 &wo, float *W) {
    constexpr bool POL = ST<SPEC>::POL;
    constexpr int NC = ST<SPEC>::NC;
    const mib_phase &p = sc.d.phases[idx];
    float pdf, ct, sp, cp;
    int ptype = p.type;
    if (!MIB_FT(FT_TAB) && (ptype == MIB_PHASE_TABPHASE || ptype == MIB_PHASE_TABPHASE_POLARIZED)) ptype = MIB_PHASE_ISOTROPIC;   // not in this instantiation
    if (!MIB_FT(FT_ANALYTIC) && ptype != MIB_PHASE_TABPHASE) ptype = MIB_PHASE_TABPHASE_POLARIZED;
    switch (ptype) {
        case MIB_PHASE_ISOTROPIC:                                                        // isotropic.cpp:40-49
            wo = square_to_uniform_sphere(s2x, s2y);
            pdf = kInvFourPi;
            break;
        case MIB_PHASE_HG: {                                                             // hg.cpp:73-94
            float g = p.g;
            float sqr_term = div_(1.f - sqr_(g), 1.f - g + 2.f * g * s2x);
            ct = div_(1.f + sqr_(g) - sqr_(sqr_term), 2.f * g);
            if (fabsf(g) < kEps) ct = 1.f - 2.f * s2x;
            float st = safe_sqrt_(1.f - sqr_(ct));
            sincos_(2.f * kPi * s2y, sp, cp);
            wo = v3(st * cp, st * sp, ct);
            pdf = eval_hg(g, -ct);
            break;
        }
        case MIB_PHASE_RAYLEIGH: {                                                       // rayleigh.cpp:100-116
            float z = 2.f * (2.f * s2x - 1.f), tmp = sqrt_(sqr_(z) + 1.f);
            float A = cbrt_(z + tmp), B = cbrt_(z - tmp);
            ct = A + B;
            float st = safe_sqrt_(1.0f - sqr_(ct));
            sincos_(kTwoPi * s2y, sp, cp);
            wo = v3(st * cp, st * sp, ct);
            pdf = rayleigh_pdf(ct);
            float ipdf = rcp_(pdf);
            if constexpr (POL) {
                rayleigh_matrix(ct, st, idx, W);
                m4_phase_rotate(W, wo, v3(0.f, 0.f, -1.f));
#pragma unroll
                for (int i = 0; i < 16; ++i) W[i] *= ipdf;
            } else {
                float r1, r2, r3;
                rayleigh_factors(idx, r1, r2, r3);
                (void) r3;
                float val = (3.f / 16.f) * kInvPi * r1 * (r2 + sqr_(ct));
                for (int c = 0; c < NC; ++c) W[c] = val * ipdf;
            }
            return pdf;
        }
        case MIB_PHASE_TABPHASE: {                                                       // tabphase.cpp:91-119
            ct = tab_sample(sc.tables[idx][0], s2x);
            float st = safe_sqrt_(1.f - ct * ct);
            sincos_(2.f * kPi * s2y, sp, cp);
            wo = v3(st * cp, st * sp, ct);
            pdf = tab_eval(sc.tables[idx][0], ct) * sc.tables[idx][0].normalization * kInvTwoPi;
            break;
        }
        default: {                                                                       // tabphase_polarized.cpp:119-144
            ct = tab_sample(sc.tables[idx][0], s2x);
            float st = safe_sqrt_(1.f - ct * ct);
            sincos_(2.f * kPi * s2y, sp, cp);
            wo = v3(st * cp, st * sp, ct);
            pdf = phase_eval<SPEC>(idx, wo, W);
            float ipdf = rcp_(pdf);
#pragma unroll
            for (int i = 0; i < ST<SPEC>::NT; ++i) W[i] *= ipdf;
            return pdf;
        }
    }
    if constexpr (POL) m4_identity(W, 1.f);
    else { for (int c = 0; c < NC; ++c) W[c] = 1.f; }
    return pdf;
}

// ---- emitters (constant.cpp:116-146, directional.cpp:150-177, scene.cpp:227-353) ----
template <int NC> struct DirSample { V3 p, d; float dist, pdf; bool delta; int emitter; };

template <int NC>
MIB_DH void sample_emitter_direction(V3 ref_p, float sx, float sy, DirSample<NC> &ds, float *spec) {
    int n = sc.d.n_emitters;
#pragma unroll
    for (int c = 0; c < NC; ++c) spec[c] = 0.f;
    ds.pdf = 0.f; ds.dist = 0.f; ds.delta = false; ds.emitter = -1; ds.p = v3(0.f); ds.d = v3(0.f);
    if (n == 0) return;
    int index = 0;
    float weight = 1.f, pmf = 1.f;
    if (n > 1) {
        float ec = (float) n, scaled = sx * ec;
        index = (int) min((uint32_t) scaled, (uint32_t) n - 1u);
        sx = scaled - (float) index;
        weight = ec;
        pmf = rcp_(ec);
    }
    const mib_emitter &e = sc.d.emitters[index];
    if (e.type == MIB_EMITTER_CONSTANT) {
        V3 d = square_to_uniform_sphere(sx, sy);
        V3 c = v3(e.bsphere_center[0], e.bsphere_center[1], e.bsphere_center[2]);
        float radius = fmaxf(e.bsphere_radius, norm(ref_p - c)), dist = 2.f * radius;
        ds.p = vfma(d, dist, ref_p); ds.pdf = kInvFourPi; ds.delta = false; ds.d = d; ds.dist = dist;
        float ip = rcp_(ds.pdf);
#pragma unroll
        for (int c2 = 0; c2 < NC; ++c2) spec[c2] = e.radiance[NC == 3 ? c2 : 0] * ip;
    } else if (MIB_FT(FT_MESH) && e.type == MIB_EMITTER_ENVMAP) {
        // envmap.cpp:404-446
        const float *buf = sc.buffers[e.env_buffer];
        float u, v, pdf;
        env_warp_sample(e, buf, sx, sy, u, v, pdf);
        u += div_(.5f, (float) (e.env_res[0] - 1));
        const bool act = pdf > 0.f;
        float st, ct, sp, cp;
        sincos_(v * kPi, st, ct); sincos_(u * kTwoPi, sp, cp);
        const V3 dl = v3(sp * st, ct, -(cp * st));
        V3 c = v3(e.bsphere_center[0], e.bsphere_center[1], e.bsphere_center[2]);
        float radius = fmaxf(e.bsphere_radius, norm(ref_p - c)), dist = 2.f * radius;
        const float inv_sin_theta = env_inv_sin_theta(dl);
        const V3 d = xform_vector(e.env_to_world, dl);
        ds.p = vfma(d, dist, ref_p); ds.delta = false; ds.d = d; ds.dist = dist;
        ds.pdf = act ? pdf * inv_sin_theta * (1.f / (2.f * kPi * kPi)) : 0.f;
        float val[NC];
        env_eval_spectrum<NC>(e, buf, u, v, val);
        const float ip = rcp_(ds.pdf);
#pragma unroll
        for (int c2 = 0; c2 < NC; ++c2) spec[c2] = act ? val[c2] * ip : 0.f;
    } else if (MIB_FT(FT_MESH) && e.type == MIB_EMITTER_SPOT) {
        // spot.cpp:177-214, falloff_curve :143-151
        const V3 pos = v3(e.env_to_world[3], e.env_to_world[7], e.env_to_world[11]);
        V3 d = pos - ref_p;
        const float dist = norm(d), inv_dist = rcp_(dist);
        d = d * inv_dist;
        const V3 local = normalize(xform_vector(e.env_to_local, -d));
        const float cos_theta = local.z;
        const float beam_res = cos_theta >= e.spot_cos_beam ? 1.f : (e.spot_cutoff_angle - acos_(cos_theta)) * e.spot_inv_transition;
        const float falloff = cos_theta > e.spot_cos_cutoff ? beam_res : 0.f;
        ds.p = pos; ds.pdf = 1.f; ds.delta = true; ds.d = d; ds.dist = dist;
        const float k = falloff * (inv_dist * inv_dist);
#pragma unroll
        for (int c2 = 0; c2 < NC; ++c2) spec[c2] = falloff > 0.f ? e.radiance[NC == 3 ? c2 : 0] * k : 0.f;
    } else if (e.type == MIB_EMITTER_DIRECTIONAL) {
        V3 d = v3(e.direction[0], e.direction[1], e.direction[2]);
        float dist = inf_f();
        ds.p = ref_p - d * dist; ds.pdf = 1.f; ds.delta = true; ds.d = -d; ds.dist = dist;
#pragma unroll
        for (int c2 = 0; c2 < NC; ++c2) spec[c2] = e.radiance[NC == 3 ? c2 : 0];
    }
    else if (MIB_FT(FT_AREA) && e.type == MIB_EMITTER_AREA) {
        // area.cpp:119-143 (uniform radiance) -> Shape::sample_direction (shape.cpp:358-375) -> Rectangle::sample_position
        // (rectangle.cpp:166-180); scene.py admits area emitters on rectangles only
        const mib_shape &sh = sc.d.shapes[e.shape];
        V3 p, nn;
        if (MIB_FT(FT_MESH) && sh.type == MIB_SHAPE_MESH) {
            // Mesh::sample_position (mesh.cpp:742-795): a triangle by area (DiscreteDistribution::sample_reuse, distr_1d.h:116-182),
            // a uniform point on it (warp.h:153-156), the interpolated vertex normal when the mesh has normals
            const MeshView m = mesh_view(sh, sc.buffers[sh.mesh_buffer]);
            const uint32_t F = (uint32_t) sh.mesh_faces;
            const float *area = reinterpret_cast<const float *>(m.order + F), *cdf = area + F;
            const float sum = __ldg(cdf + F - 1), normalization = sh.inv_surface_area;
            uint32_t first = 0, last = F - 1;
            while (first < F && !(__ldg(area + first) > 0.f)) ++first;
            while (last > 0 && !(__ldg(area + last) > 0.f)) --last;
            const float value = sy * sum;
            uint32_t start = first, end = last;
            const uint32_t iterations = start < end ? (uint32_t) (31 - __clz((int) (end - start))) + 1u : 0u;
            for (uint32_t it = 0; it < iterations; ++it) {
                const uint32_t middle = (start + end) >> 1;
                if (__ldg(cdf + middle) < value) start = min(middle + 1u, end); else end = middle;
            }
            const uint32_t f = start;
            const float pmf = __ldg(area + f) * normalization, c0 = f > 0 ? __ldg(cdf + f - 1) * normalization : 0.f;
            const float ry = div_(sy - c0, pmf);
            const uint32_t *fi = m.faces + 3 * (size_t) f;
            const uint32_t i0 = __ldg(fi), i1 = __ldg(fi + 1), i2 = __ldg(fi + 2);
            const V3 p0 = mesh_vec3(m.pos, i0), p1 = mesh_vec3(m.pos, i1), p2 = mesh_vec3(m.pos, i2);
            const V3 e0 = p1 - p0, e1 = p2 - p0;
            const float t = safe_sqrt_(1.f - sx), bx = 1.f - t, by = t * ry;
            p = vfma(e0, bx, vfma(e1, by, p0));
            if (m.nrm) {
                const V3 n0 = mesh_vec3(m.nrm, i0), n1 = mesh_vec3(m.nrm, i1), n2 = mesh_vec3(m.nrm, i2);
                nn = vfma(n0, 1.f - bx - by, vfma(n1, bx, n2 * by));
            } else nn = cross(e0, e1);
            nn = normalize(nn);
        } else {
            p = xform_point_affine(sh.to_world, v3(sx * 2.f - 1.f, sy * 2.f - 1.f, 0.f));
            nn = v3(sh.frame_n[0], sh.frame_n[1], sh.frame_n[2]);
        }
        V3 d = p - ref_p;
        float d2 = sqnorm(d), dist = sqrt_(d2);
        d = d * rcp_(dist);                                  // Vector / Float = multiplication by rcp (array_router.h:226-229)
        float dp = fabsf(dot(d, nn)), x = div_(d2, dp);
        float pdf = sh.inv_surface_area * (isfinite(x) ? x : 0.f);
        bool act = dot(d, nn) < 0.f && pdf != 0.f;
        ds.p = p; ds.pdf = pdf; ds.delta = false; ds.d = d; ds.dist = dist;
        float ip = rcp_(pdf);
#pragma unroll
        for (int c2 = 0; c2 < NC; ++c2) spec[c2] = act ? e.radiance[NC == 3 ? c2 : 0] * ip : 0.f;
    }
    ds.emitter = index;
    if (n > 1) {
        ds.pdf *= pmf;
#pragma unroll
        for (int c2 = 0; c2 < NC; ++c2) spec[c2] *= weight;
    }
}
// scene.cpp:355-366 with constant.cpp:139-144 / directional (0) / area.cpp:145-173: DirectionSample3f(scene, si, ref)
// (d = (si.p - ref.p) / dist, n = si.sh_frame.n) and Shape::pdf_direction (shape.cpp:377-388)
MIB_D float pdf_emitter_direction(int emitter, V3 ref_p, V3 si_p, V3 si_n, V3 minus_wi) {
    const mib_emitter &e = sc.d.emitters[emitter];
    float pmf = sc.d.n_emitters > 1 ? rcp_((float) sc.d.n_emitters) : 1.f;
    float pdf = 0.f;
    if (e.type == MIB_EMITTER_CONSTANT) pdf = kInvFourPi;
    else if (MIB_FT(FT_MESH) && e.type == MIB_EMITTER_ENVMAP) {
        // envmap.cpp:448-463 with ds.d = -si.wi of the ray that left the scene
        float u, v;
        V3 d;
        env_dir_to_uv(e, minus_wi, u, v, d);
        u -= div_(.5f, (float) (e.env_res[0] - 1));
        u -= floorf(u); v -= floorf(v);
        pdf = env_warp_eval(e, sc.buffers[e.env_buffer], u, v) * env_inv_sin_theta(d) * (1.f / (2.f * kPi * kPi));
    }
    else if (MIB_FT(FT_AREA) && e.type == MIB_EMITTER_AREA) {
        V3 rel = si_p - ref_p;
        float dist = norm(rel);
        V3 d = rel * rcp_(dist);
        float dp = dot(d, si_n), adp = fabsf(dp);
        float v = sc.d.shapes[e.shape].inv_surface_area * (adp != 0.f ? div_(dist * dist, adp) : 0.f);
        pdf = dp < 0.f ? v : 0.f;
    }
    return pdf * pmf;
}
// Emitter::eval at a hit, channel r: the area light is one-sided (area.cpp:83-90)
template <int NC> MIB_D float emitter_eval(const mib_emitter &e, int r, float cos_theta_wi, V3 minus_wi) {
    if (MIB_FT(FT_AREA) && e.type == MIB_EMITTER_AREA && !(cos_theta_wi > 0.f)) return 0.f;
    if (MIB_FT(FT_MESH) && e.type == MIB_EMITTER_ENVMAP) {                   // envmap.cpp:343-354
        float u, v, val[NC];
        V3 l;
        env_dir_to_uv(e, minus_wi, u, v, l);
        env_eval_spectrum<NC>(e, sc.buffers[e.env_buffer], u, v, val);
        return val[NC == 3 ? r : 0];
    }
    return e.radiance[NC == 3 ? r : 0];
}
MIB_D int si_emitter(const SurfIt &si) {
    return valid(si) ? sc.d.shapes[si.shape].emitter : sc.d.environment;
}

// ---- BSDF flags (include/mitsuba/render/bsdf.h:37-64) ----
constexpr uint32_t kBsdfNull = 0x1, kBsdfDiffuseRefl = 0x2, kBsdfSmooth = 0x2 | 0x4 | 0x8 | 0x10, kBsdfDelta = 0x1 | 0x20 | 0x40;

// ---- sensors ----
MIB_DH Ray sensor_ray(float px, float py, float ax, float ay) {
    const mib_sensor &s = sc.d.sensor;
    Ray ray;
    if (s.type == MIB_SENSOR_PERSPECTIVE) {                                  // perspective.cpp:201-235
        float ox = div_((float) s.film_width * s.principal_point_offset[0], (float) s.crop_width),
              oy = div_((float) s.film_height * s.principal_point_offset[1], (float) s.crop_height);
        V3 near_p = xform_point(s.sample_to_camera, v3(px + ox, py + oy, 0.f));
        V3 d = normalize(near_p);
        ray.o = v3(s.to_world[3], s.to_world[7], s.to_world[11]);
        ray.d = xform_vector(s.to_world, d);
        float inv_z = rcp_(d.z), near_t = s.near_clip * inv_z, far_t = s.far_clip * inv_z;
        ray.o = ray.o + ray.d * near_t;
        ray.maxt = far_t - near_t;
    } else if (s.type == MIB_SENSOR_DISTANT) {                               // distant.cpp:193-235
        ray.d = xform_vector(s.to_world, v3(0.f, 0.f, 1.f));
        if (s.target_type == 1) {
            ray.o = v3(s.target[0], s.target[1], s.target[2]) - 2.f * ray.d * s.bsphere_radius;
        } else {
            float ox, oy;
            square_to_uniform_disk_concentric(ax, ay, ox, oy);
            V3 perp = xform_vector(s.to_world, v3(ox, oy, 0.f));
            V3 c = v3(s.bsphere_center[0], s.bsphere_center[1], s.bsphere_center[2]);
            ray.o = c + perp * s.bsphere_radius - ray.d * s.bsphere_radius;
        }
        ray.maxt = largest_f();
    } else {                                                                 // radiancemeter.cpp:97-126
        ray.o = xform_point_affine(s.to_world, v3(0.f));
        ray.d = xform_vector(s.to_world, v3(0.f, 0.f, 1.f));
        ray.o = ray.o + ray.d * kRayEps;
        ray.maxt = largest_f();
    }
    return ray;
}

} // namespace mib

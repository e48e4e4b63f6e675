// oplugins.h -- TEST INFRASTRUCTURE (parity oracle). Not part of the product.
// Phase functions, emitters, BSDFs and sensors restated from the reference.
#pragma once
#include "oscene.h"

namespace orc {

// ------------------------------------------------------------------
// warp (include/mitsuba/core/warp.h)
// ------------------------------------------------------------------
static inline float circ(float x) { return safe_sqrt(fnmadd(x, x, 1.f)); }               // warp.h:23
static inline Vec3 square_to_uniform_sphere(float sx, float sy) {                           // warp.h:250-255
    float z = fnmadd(2.f, sy, 1.f), r = circ(z), s, c;
    sincos(2.f * Pi * sx, &s, &c);
    return { r * c, r * s, z };
}
static inline void square_to_uniform_disk_concentric(float sx, float sy, float *ox, float *oy) { // warp.h:54-90
    float x = fmsub(2.f, sx, 1.f), y = fmsub(2.f, sy, 1.f);
    bool is_zero = x == 0.f && y == 0.f, quadrant_1_or_3 = std::fabs(x) < std::fabs(y);
    float r = quadrant_1_or_3 ? y : x, rp = quadrant_1_or_3 ? x : y;
    float phi = 0.25f * Pi * rp / r;
    if (quadrant_1_or_3) phi = 0.5f * Pi - phi;
    if (is_zero) phi = 0.f;
    float s, c; sincos(phi, &s, &c);
    *ox = r * c; *oy = r * s;
}
static inline Vec3 square_to_cosine_hemisphere(float sx, float sy) {                        // warp.h:412-420
    float px, py; square_to_uniform_disk_concentric(sx, sy, &px, &py);
    // squared_norm of a Point2f
    float z = safe_sqrt(1.f - fmadd(py, py, sqr(px)));
    return { px, py, z };
}

// ------------------------------------------------------------------
// Phase functions.  wo is LOCAL (+z = ray.d), mei.wi = (0,0,-1) (medium.cpp:48-54)
// ------------------------------------------------------------------
template <typename Spec> struct PhaseSample { Vec3 wo; Spec weight; float pdf; };
template <typename Spec> struct PhaseEval { Spec val; float pdf; };

template <typename Spec> static inline Spec phase_scalar(float v) { return Spec(v); }

// rayleigh.cpp:50-95 / tabphase_polarized.cpp:169-197: rotate from the scattering plane to the implicit bases
static inline Mueller phase_rotate(const Mueller &M, Vec3 wo, Vec3 wi) {
    Vec3 wo_hat = wo, wi_hat = wi;
    Vec3 x_hat = cross(-wo_hat, wi_hat),
         p_axis_in = normalize(cross(x_hat, -wo_hat)),
         p_axis_out = normalize(cross(x_hat, wi_hat));
    Mueller R = rotate_mueller_basis(M, -wo_hat, p_axis_in, stokes_basis(-wo_hat),
                                     wi_hat, p_axis_out, stokes_basis(wi_hat));
    zero_nans(R);
    return R;
}

template <int SPEC> struct Phase {
    using Tr = SpecTraits<SPEC>;
    using Spec = typename Tr::Spec;

    // hg.cpp:66-71
    static float eval_hg(float g, float cos_theta) {
        float temp = 1.f + sqr(g) + 2.f * g * cos_theta;
        return InvFourPi * (1.f - sqr(g)) / (temp * std::sqrt(temp));
    }
    static float rayleigh_pdf(float c) { return (3.f / 16.f) * InvPi * (1.f + sqr(c)); }   // rayleigh.cpp:50-52
    static Spec eval_rayleigh(float rho, Vec3 wo, Vec3 wi, float cos_theta) {                // rayleigh.cpp:54-98
        if constexpr (Tr::POL) {
            float sin_theta = safe_sqrt(1.0f - sqr(cos_theta));
            return phase_rotate(rayleigh_scatter(cos_theta, sin_theta, rho), wo, wi);
        } else {
            float r1 = (1.f - rho) / (1.f + rho / 2.f), r2 = (1.f + rho) / (1.f - rho);
            return Spec((3.f / 16.f) * InvPi * r1 * (r2 + sqr(cos_theta)));
        }
    }

    static PhaseEval<Spec> eval_pdf(const Scene &sc, int idx, Vec3 wo, Vec3 wi) {
        const mib_phase &p = sc.d.phases[idx];
        switch (p.type) {
            case MIB_PHASE_ISOTROPIC: return { Spec(InvFourPi), InvFourPi };                 // isotropic.cpp:51-58
            case MIB_PHASE_HG: { float pdf = eval_hg(p.g, dot(wo, wi)); return { Spec(pdf), pdf }; } // hg.cpp:96-104
            case MIB_PHASE_RAYLEIGH: {                                                        // rayleigh.cpp:118-128
                float cos_theta = dot(wo, -wi);
                return { eval_rayleigh(p.depolarization, wo, wi, cos_theta), rayleigh_pdf(cos_theta) };
            }
            case MIB_PHASE_TABPHASE: {                                                        // tabphase.cpp:121-133
                float cos_theta = -dot(wo, wi);
                float pdf = sc.tables[idx].m11.eval_pdf_normalized(cos_theta) * InvTwoPi;
                return { Spec(pdf), pdf };
            }
            default: {                                                                        // tabphase_polarized.cpp:146-205
                const PhaseTables &t = sc.tables[idx];
                float cos_theta = -dot(wo, wi);
                float m11 = t.m11.eval_pdf(cos_theta);
                Spec val;
                if constexpr (Tr::POL) {
                    float m12 = t.m12.eval_pdf(cos_theta), m33 = t.m33.eval_pdf(cos_theta), m34 = t.m34.eval_pdf(cos_theta);
                    Mueller M;
                    M.m[0][0] = m11; M.m[0][1] = m12; M.m[1][0] = m12; M.m[1][1] = m11;
                    M.m[2][2] = m33; M.m[2][3] = m34; M.m[3][2] = -m34; M.m[3][3] = m33;
                    val = phase_rotate(M, wo, wi);
                } else {
                    val = Spec(m11);
                }
                float pdf = t.m11.eval_pdf_normalized(cos_theta) * InvTwoPi;
                return { val, pdf };
            }
        }
    }

    static PhaseSample<Spec> sample(const Scene &sc, int idx, Vec3 wi, float /*sample1*/, float s2x, float s2y) {
        const mib_phase &p = sc.d.phases[idx];
        switch (p.type) {
            case MIB_PHASE_ISOTROPIC: {                                                       // isotropic.cpp:40-49
                Vec3 wo = square_to_uniform_sphere(s2x, s2y);
                return { wo, Spec(1.f), InvFourPi };
            }
            case MIB_PHASE_HG: {                                                              // hg.cpp:73-94
                float g = p.g;
                float sqr_term = (1.f - sqr(g)) / (1.f - g + 2.f * g * s2x),
                      cos_theta = (1.f + sqr(g) - sqr(sqr_term)) / (2.f * g);
                if (std::fabs(g) < EpsilonF) cos_theta = 1.f - 2.f * s2x;
                float sin_theta = safe_sqrt(1.f - sqr(cos_theta)), sp, cp;
                sincos(2.f * Pi * s2y, &sp, &cp);
                Vec3 wo(sin_theta * cp, sin_theta * sp, cos_theta);
                return { wo, Spec(1.f), eval_hg(g, -cos_theta) };
            }
            case MIB_PHASE_RAYLEIGH: {                                                        // rayleigh.cpp:100-116
                float z = 2.f * (2.f * s2x - 1.f), tmp = std::sqrt(sqr(z) + 1.f);
                float A = cbrt_(z + tmp), B = cbrt_(z - tmp), cos_theta = A + B;
                float sin_theta = safe_sqrt(1.0f - sqr(cos_theta)), sp, cp;
                sincos(TwoPi * s2y, &sp, &cp);
                Vec3 wo(sin_theta * cp, sin_theta * sp, cos_theta);
                float pdf = rayleigh_pdf(cos_theta);
                Spec w = eval_rayleigh(p.depolarization, wo, wi, cos_theta) * rcp(pdf);
                return { wo, w, pdf };
            }
            case MIB_PHASE_TABPHASE: {                                                        // tabphase.cpp:91-119
                float ctp = sc.tables[idx].m11.sample(s2x);
                float stp = safe_sqrt(1.f - ctp * ctp), sp, cp;
                sincos(2.f * Pi * s2y, &sp, &cp);
                Vec3 wo(stp * cp, stp * sp, ctp);
                float pdf = sc.tables[idx].m11.eval_pdf_normalized(ctp) * InvTwoPi;
                return { wo, Spec(1.f), pdf };
            }
            default: {                                                                        // tabphase_polarized.cpp:119-144
                float ctp = sc.tables[idx].m11.sample(s2x);
                float stp = safe_sqrt(1.f - ctp * ctp), sp, cp;
                sincos(2.f * Pi * s2y, &sp, &cp);
                Vec3 wo(stp * cp, stp * sp, ctp);
                PhaseEval<Spec> e = eval_pdf(sc, idx, wo, wi);
                return { wo, e.val * rcp(e.pdf), e.pdf };
            }
        }
    }
};

// ------------------------------------------------------------------
// Emitters (constant.cpp:116-146, directional.cpp:150-177), scene.cpp:299-366
// ------------------------------------------------------------------
template <int NC> struct DirectionSample {
    Vec3 p, n, d; float dist = 0.f, pdf = 0.f; bool delta = false; int emitter = -1;
};

template <int NC> static inline Color<NC> emitter_radiance(const mib_emitter &e) {
    Color<NC> r; for (int c = 0; c < NC; ++c) r.v[c] = e.radiance[NC == 3 ? c : 0]; return r;
}

// ---- environment map (src/emitters/envmap.cpp:343-560; Hierarchical2D<Float, 0>, include/mitsuba/core/distr_2d.h) ----
static inline float lerp_(float a, float b, float t) { return fmadd(b, t, fnmadd(a, t, a)); }         // array_router.h lerp
// warp.h:447-453
static inline float interval_to_linear(float v0, float v1, float sample) {
    return std::fabs(v0 - v1) > 1e-4f * (v0 + v1) ? (v0 - safe_sqrt(lerp_(sqr(v0), sqr(v1), sample))) / (v0 - v1) : sample;
}
static inline uint32_t env_level_index(uint32_t x, uint32_t y, uint32_t width) {                      // distr_2d.h:766-770
    return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + ((y & ~1u) * width);
}
static inline float clamp01(float v) { return minimum(maximum(v, 0.f), 1.f); }
// Hierarchical2D::sample (distr_2d.h:492-575): -> uv in [0,1]^2 and the density
static inline void env_warp_sample(const mib_emitter &e, const float *buf, float sx, float sy, float *u_out, float *v_out, float *pdf_out) {
    sx = clamp01(sx); sy = clamp01(sy);
    uint32_t ox = 0, oy = 0;
    for (int l = e.env_levels - 2; l > 0; --l) {
        const float *lv = buf + e.env_level_offset[l];
        ox <<= 1; oy <<= 1;
        uint32_t i = env_level_index(ox, oy, (uint32_t) e.env_level_width[l]);
        float v00 = lv[i], v10 = lv[i + 1], v01 = lv[i + 2], v11 = lv[i + 3];
        sx = clamp01(sx); sy = clamp01(sy);
        float r0 = v00 + v10, r1 = v01 + v11;
        sy *= r0 + r1;
        bool mask = sy > r0;
        if (mask) { oy += 1u; sy -= r0; }
        sy /= mask ? r1 : r0;
        float c0 = mask ? v01 : v00, c1 = mask ? v11 : v10;
        sx *= c0 + c1;
        mask = sx > c0;
        if (mask) sx -= c0;
        sx /= mask ? c1 : c0;
        if (mask) ox += 1u;
    }
    const float *l0 = buf + e.env_level_offset[0];
    const uint32_t w = (uint32_t) e.env_level_width[0], i = ox + oy * w;
    float v00 = l0[i], v10 = l0[i + 1], v01 = l0[i + w], v11 = l0[i + w + 1];
    // warp::square_to_bilinear (warp.h:480-494)
    float r0 = v00 + v10, r1 = v01 + v11;
    sy = interval_to_linear(r0, r1, sy);
    float c0 = lerp_(v00, v01, sy), c1 = lerp_(v10, v11, sy);
    sx = interval_to_linear(c0, c1, sx);
    *pdf_out = lerp_(c0, c1, sx);
    const float psx = 1.f / (float) (e.env_res[0] - 1), psy = 1.f / (float) (e.env_res[1] - 1);      // m_patch_size = 1 / n_patches
    *u_out = ((float) (int) ox + sx) * psx;
    *v_out = ((float) (int) oy + sy) * psy;
}
// Hierarchical2D::eval (distr_2d.h:668-700)
static inline float env_warp_eval(const mib_emitter &e, const float *buf, float px, float py) {
    px = clamp01(px); py = clamp01(py);
    px *= (float) (e.env_res[0] - 1); py *= (float) (e.env_res[1] - 1);
    uint32_t ox = std::min((uint32_t) (int) px, (uint32_t) (e.env_res[0] - 2)), oy = std::min((uint32_t) (int) py, (uint32_t) (e.env_res[1] - 2));
    px -= (float) (int) ox; py -= (float) (int) oy;
    const float *l0 = buf + e.env_level_offset[0];
    const uint32_t w = (uint32_t) e.env_level_width[0], i = ox + oy * w;
    float v00 = l0[i], v10 = l0[i + 1], v01 = l0[i + w], v11 = l0[i + w + 1];
    return lerp_(lerp_(v00, v10, px), lerp_(v01, v11, px), py);                                      // square_to_bilinear_pdf
}
// envmap.cpp:517-576 (rgb / mono branch): bilinear look-up of the texel array
template <int NC> static inline Color<NC> env_eval_spectrum(const mib_emitter &e, const float *buf, float u, float v) {
    const uint32_t rx = (uint32_t) e.env_res[0], ry = (uint32_t) e.env_res[1];
    u -= .5f / (float) (rx - 1u);
    u -= std::floor(u); v -= std::floor(v);
    u *= (float) (rx - 1u); v *= (float) (ry - 1u);
    uint32_t px = std::min((uint32_t) u, rx - 2u), py = std::min((uint32_t) v, ry - 2u);
    float w1x = u - (float) px, w1y = v - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    uint32_t index = py * rx + px;
    Color<NC> out;
    for (int c = 0; c < NC; ++c) {
        float v00 = buf[3 * index + c], v10 = buf[3 * (index + 1) + c], v01 = buf[3 * (index + rx) + c], v11 = buf[3 * (index + rx + 1) + c];
        float v0 = fmadd(w0x, v00, w1x * v10), v1 = fmadd(w0x, v01, w1x * v11);
        out.v[c] = fmadd(w0y, v0, w1y * v1) * e.env_scale;
    }
    return out;
}
static inline void env_dir_to_uv(const mib_emitter &e, Vec3 d_world, float *u, float *v, Vec3 *local) {
    Vec3 d = transform_vector(mat4_from(e.env_to_local), d_world);
    *u = atan2_(d.x, -d.z) * InvTwoPi;
    *v = safe_acos_(d.y) * InvPi;
    *local = d;
}
// envmap.cpp:343-354 eval(si): si.wi is the world-space direction back along the ray for a ray that left the scene
template <int NC> static inline Color<NC> env_eval(const Scene &sc, const mib_emitter &e, Vec3 minus_wi) {
    float u, v; Vec3 l;
    env_dir_to_uv(e, minus_wi, &u, &v, &l);
    return env_eval_spectrum<NC>(e, sc.buffers[e.env_buffer], u, v);
}

// Returns the UNPOLARIZED spectrum; callers wrap with depolarizer<Spectrum>()
template <int NC>
static inline Color<NC> sample_emitter_direction(const Scene &sc, Vec3 ref_p, float sx, float sy, DirectionSample<NC> *ds_out) {
    DirectionSample<NC> ds;
    Color<NC> spec(0.f);
    int n = sc.d.n_emitters;
    if (n == 0) { *ds_out = ds; return spec; }
    int index = 0; float emitter_weight = 1.f, pmf = 1.f;
    if (n > 1) {
        // scene.cpp:227-262 sample_emitter (uniform selection, sample reuse)
        float ec = (float) n;
        float index_sample_scaled = sx * ec;
        index = (int) std::min((uint32_t) index_sample_scaled, (uint32_t) n - 1u);
        sx = index_sample_scaled - (float) index;
        emitter_weight = ec;
        pmf = rcp(ec);
    }
    const mib_emitter &e = sc.d.emitters[index];
    Vec3 c(e.bsphere_center[0], e.bsphere_center[1], e.bsphere_center[2]);
    if (e.type == MIB_EMITTER_CONSTANT) {
        Vec3 d = square_to_uniform_sphere(sx, sy);
        float radius = maximum(e.bsphere_radius, norm(ref_p - c)), dist = 2.f * radius;
        ds.p = fmadd(d, dist, ref_p); ds.n = -d; ds.pdf = InvFourPi; ds.delta = false; ds.d = d; ds.dist = dist;
        spec = div_scalar(emitter_radiance<NC>(e), ds.pdf);
    } else if (e.type == MIB_EMITTER_ENVMAP) {
        // envmap.cpp:404-446
        const float *buf = sc.buffers[e.env_buffer];
        float u, v, pdf;
        env_warp_sample(e, buf, sx, sy, &u, &v, &pdf);
        u += .5f / (float) (e.env_res[0] - 1);
        bool active = pdf > 0.f;
        float theta = v * Pi, phi = u * TwoPi, st, ct, sp, cp;
        sincos(theta, &st, &ct); sincos(phi, &sp, &cp);
        Vec3 dl(sp * st, ct, -(cp * st));                              // sphdir -> (y, z, -x)
        float radius = maximum(e.bsphere_radius, norm(ref_p - c)), dist = 2.f * radius;
        float inv_sin_theta = rsqrt(maximum(maximum(sqr(dl.x) + sqr(dl.z), sqr(EpsilonF)), 0.f));
        Vec3 d = transform_vector(mat4_from(e.env_to_world), dl);
        ds.p = fmadd(d, dist, ref_p); ds.n = -d; ds.delta = false; ds.d = d; ds.dist = dist;
        ds.pdf = active ? pdf * inv_sin_theta * (1.f / (2.f * sqr(Pi))) : 0.f;
        Color<NC> w = div_scalar(env_eval_spectrum<NC>(e, buf, u, v), ds.pdf);
        spec = active ? w : Color<NC>(0.f);
    } else if (e.type == MIB_EMITTER_SPOT) {
        // spot.cpp:177-214, falloff_curve :143-151
        Vec3 pos(e.env_to_world[3], e.env_to_world[7], e.env_to_world[11]);
        Vec3 d = pos - ref_p;
        float dist = norm(d), inv_dist = rcp(dist);
        d = d * inv_dist;
        Vec3 local = normalize(transform_vector(mat4_from(e.env_to_local), -d));
        float cos_theta = local.z;
        float beam_res = cos_theta >= e.spot_cos_beam ? 1.f : (e.spot_cutoff_angle - acos_(cos_theta)) * e.spot_inv_transition;
        float falloff = cos_theta > e.spot_cos_cutoff ? beam_res : 0.f;
        ds.p = pos; ds.n = Vec3(0.f); ds.pdf = 1.f; ds.delta = true; ds.d = d; ds.dist = dist;
        spec = falloff > 0.f ? emitter_radiance<NC>(e) * (falloff * sqr(inv_dist)) : Color<NC>(0.f);
    } else if (e.type == MIB_EMITTER_DIRECTIONAL) {
        Vec3 d(e.direction[0], e.direction[1], e.direction[2]);
        float dist = Inf;
        ds.p = ref_p - d * dist; ds.n = d; ds.pdf = 1.f; ds.delta = true; ds.d = -d; ds.dist = dist;
        spec = emitter_radiance<NC>(e);
    } else {
        // area.cpp:119-143 (uniform radiance) -> Shape::sample_direction (shape.cpp:358-375) ->
        // Rectangle::sample_position (rectangle.cpp:166-180)
        const mib_shape &sh = sc.d.shapes[e.shape];
        Vec3 p, n;
        if (sh.type == MIB_SHAPE_MESH) {
            // Mesh::sample_position (mesh.cpp:742-795): a triangle by area (DiscreteDistribution::sample_reuse, distr_1d.h:116-182),
            // a uniform point on it (warp.h:153-156), the interpolated vertex normal when the mesh has normals
            MeshView m = mesh_view(sh, sc.buffers[sh.mesh_buffer]);
            const uint32_t F = (uint32_t) sh.mesh_faces;
            const float *area = reinterpret_cast<const float *>(m.order + F), *cdf = area + F;
            const float sum = cdf[F - 1], normalization = sh.inv_surface_area;
            uint32_t first = 0, last = F - 1;
            while (first < F && !(area[first] > 0.f)) ++first;                       // m_valid: first / last bin with mass
            while (last > 0 && !(area[last] > 0.f)) --last;
            float value = sy * sum;
            uint32_t start = first, end = last;                                       // dr::binary_search (array_router.h)
            uint32_t iterations = start < end ? (uint32_t) (31 - __builtin_clz(end - start)) + 1u : 0u;
            for (uint32_t it = 0; it < iterations; ++it) {
                uint32_t middle = (start + end) >> 1;
                if (cdf[middle] < value) start = std::min(middle + 1u, end); else end = middle;
            }
            const uint32_t f = start;
            float pmf = area[f] * normalization, c0 = f > 0 ? cdf[f - 1] * normalization : 0.f;
            float ry = (sy - c0) / pmf;
            const uint32_t *fi = m.faces + 3 * (size_t) f;
            Vec3 p0 = mesh_vec3(m.pos, fi[0]), p1 = mesh_vec3(m.pos, fi[1]), p2 = mesh_vec3(m.pos, fi[2]);
            Vec3 e0 = p1 - p0, e1 = p2 - p0;
            float t = safe_sqrt(1.f - sx), bx = 1.f - t, by = t * ry;
            p = fmadd(e0, bx, fmadd(e1, by, p0));
            if (m.nrm) {
                Vec3 n0 = mesh_vec3(m.nrm, fi[0]), n1 = mesh_vec3(m.nrm, fi[1]), n2 = mesh_vec3(m.nrm, fi[2]);
                n = fmadd(n0, 1.f - bx - by, fmadd(n1, bx, n2 * by));
            } else n = cross(e0, e1);
            n = normalize(n);
        } else if (sh.type == MIB_SHAPE_RECTANGLE) {
            p = transform_point_affine(mat4_from(sh.to_world), Vec3(sx * 2.f - 1.f, sy * 2.f - 1.f, 0.f));
            n = Vec3(sh.frame_n[0], sh.frame_n[1], sh.frame_n[2]);
        } else throw std::runtime_error("oracle: area emitters are supported on rectangles and meshes only");
        Vec3 d = p - ref_p;
        float dist_squared = squared_norm(d), dist = std::sqrt(dist_squared);
        d = d * rcp(dist);                                   // Vector / Float = multiplication by rcp (array_router.h:226-229)
        float dp = std::fabs(dot(d, n)), x = dist_squared / dp;
        float pdf = sh.inv_surface_area * (std::isfinite(x) ? x : 0.f);
        bool active = dot(d, n) < 0.f && pdf != 0.f;
        ds.p = p; ds.n = n; ds.d = d; ds.dist = dist; ds.pdf = pdf; ds.delta = false;
        spec = active ? div_scalar(emitter_radiance<NC>(e), pdf) : Color<NC>(0.f);
    }
    ds.emitter = index;
    if (n > 1) { ds.pdf *= pmf; spec = spec * emitter_weight; }
    *ds_out = ds;
    return spec;
}

// scene.cpp:355-366 + constant.cpp:139-144 / directional.cpp pdf_direction (= 0) / area.cpp:145-173 with
// DirectionSample3f(scene, si, ref) (records.h: d = (si.p - ref.p) / dist, n = si.sh_frame.n) and
// Shape::pdf_direction (shape.cpp:377-388)
template <int NC> static inline float pdf_emitter_direction(const Scene &sc, int emitter, Vec3 ref_p, Vec3 si_p, Vec3 si_n, Vec3 minus_wi = Vec3(0.f)) {
    const mib_emitter &e = sc.d.emitters[emitter];
    float pmf = sc.d.n_emitters > 1 ? rcp((float) sc.d.n_emitters) : 1.f;
    float pdf = 0.f;
    if (e.type == MIB_EMITTER_CONSTANT) pdf = InvFourPi;
    else if (e.type == MIB_EMITTER_ENVMAP) {
        // envmap.cpp:448-463 with ds.d = -si.wi of the ray that left the scene (records.h DirectionSample(scene, si, ref))
        float u, v; Vec3 d;
        env_dir_to_uv(e, minus_wi, &u, &v, &d);
        u -= .5f / (float) (e.env_res[0] - 1);
        u -= std::floor(u); v -= std::floor(v);
        float inv_sin_theta = rsqrt(maximum(maximum(sqr(d.x) + sqr(d.z), sqr(EpsilonF)), 0.f));
        pdf = env_warp_eval(e, sc.buffers[e.env_buffer], u, v) * inv_sin_theta * (1.f / (2.f * sqr(Pi)));
    }
    else if (e.type == MIB_EMITTER_AREA) {
        Vec3 rel = si_p - ref_p;
        float dist = norm(rel);
        Vec3 d = rel * rcp(dist);
        float dp = dot(d, si_n);
        float adp = std::fabs(dp);
        float v = sc.d.shapes[e.shape].inv_surface_area * (adp != 0.f ? (dist * dist) / adp : 0.f);
        pdf = dp < 0.f ? v : 0.f;
    }
    return pdf * pmf;
}

// Emitter::eval at a hit: constant / area radiance; the area light is one-sided (area.cpp:83-90)
template <int NC> static inline Color<NC> emitter_eval(const Scene &sc, int emitter, float cos_theta_wi, Vec3 minus_wi = Vec3(0.f)) {
    const mib_emitter &e = sc.d.emitters[emitter];
    if (e.type == MIB_EMITTER_ENVMAP) return env_eval<NC>(sc, e, minus_wi);
    if (e.type == MIB_EMITTER_AREA && !(cos_theta_wi > 0.f)) return Color<NC>(0.f);
    return emitter_radiance<NC>(e);
}

// si.emitter(scene): shape emitter if valid, else the environment (interaction.h)
static inline int si_emitter(const Scene &sc, const SurfaceInteraction &si) {
    if (si.is_valid()) return sc.d.shapes[si.shape].emitter;
    return sc.d.environment;
}

// (BSDFs: obsdf.h)

// ------------------------------------------------------------------
// Sensors (perspective.cpp:201-235, distant.cpp:193-235)
// ------------------------------------------------------------------
static inline Ray sensor_sample_ray(const Scene &sc, float px, float py, float ax, float ay) {
    const mib_sensor &s = sc.d.sensor;
    Mat4 to_world = mat4_from(s.to_world);
    Ray ray;
    if (s.type == MIB_SENSOR_PERSPECTIVE) {
        Mat4 s2c = mat4_from(s.sample_to_camera);
        // film size * offset / crop size (principal point, 0 by default)
        float ox = (float) s.film_width * s.principal_point_offset[0] / (float) s.crop_width,
              oy = (float) s.film_height * s.principal_point_offset[1] / (float) s.crop_height;
        Vec3 near_p = transform_point(s2c, Vec3(px + ox, py + oy, 0.f));
        Vec3 d = normalize(near_p);
        ray.o = Vec3(to_world.m[0][3], to_world.m[1][3], to_world.m[2][3]);
        ray.d = transform_vector(to_world, d);
        float inv_z = rcp(d.z), near_t = s.near_clip * inv_z, far_t = s.far_clip * inv_z;
        ray.o = ray.o + ray.d * near_t;
        ray.maxt = far_t - near_t;
    } else if (s.type == MIB_SENSOR_DISTANT) {
        ray.d = transform_vector(to_world, Vec3(0.f, 0.f, 1.f));
        Vec3 c(s.bsphere_center[0], s.bsphere_center[1], s.bsphere_center[2]);
        if (s.target_type == 1) {
            ray.o = Vec3(s.target[0], s.target[1], s.target[2]) - 2.f * ray.d * s.bsphere_radius;
        } else {
            float ox, oy; square_to_uniform_disk_concentric(ax, ay, &ox, &oy);
            Vec3 perp = transform_vector(to_world, Vec3(ox, oy, 0.f));
            ray.o = c + perp * s.bsphere_radius - ray.d * s.bsphere_radius;
        }
        ray.maxt = Largest;
    } else { // radiancemeter (radiancemeter.cpp:97-126)
        ray.o = transform_point_affine(to_world, Vec3(0.f));
        ray.d = transform_vector(to_world, Vec3(0.f, 0.f, 1.f));
        ray.o = ray.o + ray.d * RayEpsilon;
        ray.maxt = Largest;
    }
    return ray;
}

} // namespace orc

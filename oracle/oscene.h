// oscene.h -- TEST INFRASTRUCTURE (parity oracle). Not part of the product.
// Scene-side plugins of the reference restated on the CPU: shapes, media,
// volumes, phase functions, emitters, BSDFs and sensors.  Input is the same POD
// scene description the C ABI consumes (include/mi_b200.h).
#pragma once
#include "ospec.h"
#include "../include/mi_b200.h"
#include <vector>
#include <stdexcept>
#include <algorithm>

namespace orc {

struct Ray { Vec3 o, d; float maxt; Vec3 at(float t) const { return fmadd(d, t, o); } };

// Prepared irregular distribution (include/mitsuba/core/distr_1d.h:618-989)
struct IrregularDistr {
    const float *nodes = nullptr, *pdf = nullptr;
    std::vector<float> cdf;
    uint32_t n = 0, valid_x = 0, valid_y = 0;
    float integral = 0.f, normalization = 0.f, range_x = 0.f, range_y = 0.f;
    bool enable_sampling = true;

    // distr_1d.h:916-975 compute_cdf_scalar (double accumulation)
    void prepare(const float *nodes_, const float *pdf_, uint32_t size, bool sampling) {
        nodes = nodes_; pdf = pdf_; n = size; enable_sampling = sampling;
        if (size < 2) throw std::runtime_error("IrregularContinuousDistribution: needs at least two entries!");
        cdf.assign(size, 0.f);
        uint32_t vx = (uint32_t) -1, vy = (uint32_t) -1;
        range_x = Inf; range_y = -Inf;
        double integral_d = 0.;
        for (uint32_t i = 0; i < size - 1; ++i) {
            double x0 = nodes[i], x1 = nodes[i + 1], y0 = pdf[i], y1 = pdf[i + 1];
            double value = 0.5 * (x1 - x0) * (y0 + y1);
            range_x = minimum(range_x, (float) x0);
            range_y = maximum(range_y, (float) x1);
            integral_d += value;
            cdf[i] = (float) integral_d;
            if (!(x1 > x0)) throw std::runtime_error("IrregularContinuousDistribution: node positions must be strictly increasing!");
            if (sampling) {
                if (y0 < 0. || y1 < 0.) throw std::runtime_error("IrregularContinuousDistribution: entries must be non-negative!");
                else if (value > 0.) { if (vx == (uint32_t) -1) vx = i; vy = i; }
            }
        }
        if (sampling) {
            if (vx == (uint32_t) -1) throw std::runtime_error("IrregularContinuousDistribution: no probability mass found!");
            valid_x = vx; valid_y = vy;
            integral = cdf[vy];
            normalization = rcp(integral);
        }
    }
    // drjit/util.h:172-220
    template <typename Pred> static uint32_t binary_search(uint32_t start, uint32_t end, Pred pred) {
        uint32_t iterations = (start < end) ? (31 - __builtin_clz(end - start)) + 1 : 0;
        for (uint32_t i = 0; i < iterations; ++i) {
            uint32_t middle = (start + end) >> 1;
            if (pred(middle)) start = std::min(middle + 1, end);
            else end = middle;
        }
        return start;
    }
    // distr_1d.h:707-729
    float eval_pdf(float x) const {
        bool active = x >= range_x && x <= range_y;
        uint32_t index = binary_search(0, n, [&](uint32_t i) { return (active && i < n ? nodes[i] : 0.f) < x; });
        index = std::max(std::min(index, n - 1u), 1u) - 1u;
        float x0 = nodes[index], x1 = nodes[index + 1], y0 = pdf[index], y1 = pdf[index + 1];
        x = (x - x0) / (x1 - x0);
        return active ? fmadd(x, y1 - y0, y0) : 0.f;
    }
    float eval_pdf_normalized(float x) const { return eval_pdf(x) * normalization; }
    // distr_1d.h:781-814 (JIT predicate)
    float sample(float sample) const {
        sample *= integral;
        uint32_t index = binary_search(valid_x, valid_y, [&](uint32_t i) {
            float value = cdf[i];
            return ((value < sample) || (value == 0.f)) && (value != integral);
        });
        float x0 = nodes[index], x1 = nodes[index + 1], y0 = pdf[index], y1 = pdf[index + 1],
              c0 = index > 0 ? cdf[index - 1] : 0.f, w = x1 - x0;
        sample = (sample - c0) / w;
        float t_linear = (y0 - safe_sqrt(sqr(y0) + 2.f * sample * (y1 - y0))) / (y0 - y1),
              t_const = sample / y0, t = (y0 == y1) ? t_const : t_linear;
        return fmadd(t, w, x0);
    }
};

struct PhaseTables { IrregularDistr m11, m12, m33, m34; };

struct Scene {
    mib_scene_desc d;
    const float *buffers[MIB_MAX_BUFFERS] = {};
    size_t buffer_count[MIB_MAX_BUFFERS] = {};
    PhaseTables tables[MIB_MAX_PHASES];
    Mat4 shape_to_object[MIB_MAX_SHAPES];
    Vec3 cube_normals[MIB_MAX_SHAPES][6];   // world-space outward face normals (-x,+x,-y,+y,-z,+z)

    static float grid_max(const float *data, size_t n) { float m = -Inf; for (size_t i = 0; i < n; ++i) m = maximum(m, data[i]); return m; }

    // heterogeneous.cpp:163,176-178; grid.cpp:306-319; tabphase*.cpp ctor
    // Reconstruction filter (src/rfilters/gaussian.cpp:50-98 non-CUDA branch, tent.cpp:33-45)
    float rf_coeff[10] = {};
    float rf_inv_radius = 1.f;
    void prepare_rfilter() {
        const mib_sensor &s = d.sensor;
        if (s.rfilter == MIB_RFILTER_GAUSSIAN) {
            const double coeff[10] = { 9.992604880e-1, -4.977025247e-1, 1.222248550e-1, -1.932406282e-2, 2.136713061e-3,
                                       -1.679873860e-4, 9.202145248e-6, -3.329417433e-7, 7.128382794e-9, -6.821193280e-11 };
            double scale = 1;
            for (int i = 0; i < 10; ++i) { rf_coeff[i] = (float) (coeff[i] * scale); scale /= (double) s.rfilter_stddev * (double) s.rfilter_stddev; }
            rf_coeff[0] -= estrin(sqr(s.rfilter_radius), rf_coeff, 10);
        } else if (s.rfilter == MIB_RFILTER_TENT) {
            rf_inv_radius = 1.f / s.rfilter_radius;
        }
    }
    float rfilter_eval(float x) const {
        if (d.sensor.rfilter == MIB_RFILTER_GAUSSIAN) return maximum(estrin(sqr(x), rf_coeff, 10), 0.f);
        return maximum(0.f, 1.f - std::fabs(x * rf_inv_radius));
    }

    void prepare() {
        prepare_rfilter();
        for (int i = 0; i < d.n_media; ++i) {
            mib_medium &m = d.media[i];
            mib_volume *vols[2] = { &m.sigma_t, &m.albedo };
            for (mib_volume *v : vols) {
                if (v->type == MIB_VOLUME_GRID) {
                    size_t n = (size_t) v->res[0] * v->res[1] * v->res[2] * v->channels;
                    if (!v->fixed_max) v->max_value = grid_max(buffers[v->buffer], n);      // grid.cpp:296-297
                } else {
                    v->max_value = v->value[0];
                    for (int c = 1; c < v->channels; ++c) v->max_value = maximum(v->max_value, v->value[c]);
                }
            }
            m.max_density = m.scale * m.sigma_t.max_value;
        }
        for (int i = 0; i < d.n_phases; ++i) {
            const mib_phase &p = d.phases[i];
            if (p.type == MIB_PHASE_TABPHASE || p.type == MIB_PHASE_TABPHASE_POLARIZED) {
                tables[i].m11.prepare(buffers[p.buf_nodes], buffers[p.buf_m11], p.n_nodes, true);
                if (p.type == MIB_PHASE_TABPHASE_POLARIZED) {
                    tables[i].m12.prepare(buffers[p.buf_nodes], buffers[p.buf_m12], p.n_nodes, false);
                    tables[i].m33.prepare(buffers[p.buf_nodes], buffers[p.buf_m33], p.n_nodes, false);
                    tables[i].m34.prepare(buffers[p.buf_nodes], buffers[p.buf_m34], p.n_nodes, false);
                }
            }
        }
        for (int i = 0; i < d.n_shapes; ++i) {
            shape_to_object[i] = mat4_from(d.shapes[i].to_object);
            if (d.shapes[i].type == MIB_SHAPE_CUBE) {
                // normal transform = inverse transpose of to_world = transpose(to_object) (transform.h:146-155)
                const Mat4 &T = shape_to_object[i];
                for (int f = 0; f < 6; ++f) {
                    Vec3 nl(0.f); nl[f / 2] = (f & 1) ? 1.f : -1.f;
                    Vec3 nw;
                    for (int r = 0; r < 3; ++r)
                        nw[r] = fmadd(T.m[2][r], nl.z, fmadd(T.m[1][r], nl.y, T.m[0][r] * nl.x));
                    cube_normals[i][f] = normalize(nw);
                }
            }
        }
    }
};

// ------------------------------------------------------------------
// Surface interaction (include/mitsuba/render/interaction.h)
// ------------------------------------------------------------------
struct SurfaceInteraction {
    float t = Inf;
    Vec3 p, n, wi;
    Frame sh_frame;
    int shape = -1;
    float uv[2] = { 0.f, 0.f };         // texture coordinates (meshes, rectangles)
    bool is_valid() const { return t != Inf; }
};

// interaction.h:133-162
static inline Vec3 offset_p(Vec3 p, Vec3 n, Vec3 d) {
    float mag = (1.f + max3(abs3(p))) * RayEpsilon;
    mag = mulsign_ad(mag, dot(n, d));
    return Vec3(fmadd(mag, n.x, p.x), fmadd(mag, n.y, p.y), fmadd(mag, n.z, p.z));
}
static inline Ray spawn_ray(Vec3 p, Vec3 n, Vec3 d) { return Ray{ offset_p(p, n, d), d, Largest }; }
static inline Ray spawn_ray_to(Vec3 p, Vec3 n, Vec3 target) {
    Vec3 o = offset_p(p, n, target - p);
    Vec3 d = target - o;
    float dist = norm(d);
    d = d / dist;
    return Ray{ o, d, dist * (1.f - ShadowEpsilon) };
}

// interaction.h:245-255 + :464-484
static inline void finalize_si(SurfaceInteraction &si, Vec3 dp_du, const Ray &ray) {
    Vec3 n = si.sh_frame.n;
    si.sh_frame.s = normalize(fmadd(n, -dot(n, dp_du), dp_du));
    if (dp_du.x == 0.f && dp_du.y == 0.f && dp_du.z == 0.f) {
        Vec3 s, t; coordinate_system(n, &s, &t); si.sh_frame.s = s;
    }
    si.sh_frame.t = cross(n, si.sh_frame.s);
    si.wi = si.sh_frame.to_local(-ray.d);
}

} // namespace orc
#include "omesh.h"
namespace orc {

// Shapes.  rectangle: src/shapes/rectangle.cpp:449-470,497-568 (AD branch: p = ray(t)).
// cube / sphere: analytic (the reference's cube is a 12-triangle mesh traced by
// Embree/kd-tree, whose arithmetic is outside the tree: unpinned).
static inline float intersect_shape(const Scene &sc, int idx, const Ray &ray, int *face) {
    const mib_shape &s = sc.d.shapes[idx];
    const Mat4 &T = sc.shape_to_object[idx];
    Vec3 o = transform_point_affine(T, ray.o), d = transform_vector(T, ray.d);
    *face = 0;
    if (s.type == MIB_SHAPE_RECTANGLE) {
        float t = -o.z / d.z;
        Vec3 local = fmadd(d, t, o);
        bool hit = t >= 0.f && t <= ray.maxt && std::fabs(local.x) <= 1.f && std::fabs(local.y) <= 1.f;
        return hit ? t : Inf;
    } else if (s.type == MIB_SHAPE_CUBE) {
        float tn = -Inf, tf = Inf; int fn = 0, ff = 0;
        for (int a = 0; a < 3; ++a) {
            float inv = rcp(d[a]);
            float t1 = (-1.f - o[a]) * inv, t2 = (1.f - o[a]) * inv;
            int f1 = 2 * a, f2 = 2 * a + 1;
            if (t1 > t2) { std::swap(t1, t2); std::swap(f1, f2); }
            if (t1 > tn) { tn = t1; fn = f1; }
            if (t2 < tf) { tf = t2; ff = f2; }
        }
        if (!(tn <= tf)) return Inf;
        float t;
        if (tn >= 0.f) { t = tn; *face = fn; } else { t = tf; *face = ff; }
        if (!(t >= 0.f && t <= ray.maxt)) return Inf;
        return t;
    } else { // sphere (object space: unit sphere at origin)
        float A = squared_norm(d), B = 2.f * dot(o, d), C = squared_norm(o) - 1.f;
        float disc = fmsub(B, B, 4.f * A * C);
        if (!(disc >= 0.f)) return Inf;
        float sq = std::sqrt(disc);
        float q = -0.5f * (B + (B >= 0.f ? sq : -sq));
        float t0 = q / A, t1 = C / q;
        if (t0 > t1) std::swap(t0, t1);
        float t = t0 >= 0.f ? t0 : t1;
        if (!(t >= 0.f && t <= ray.maxt)) return Inf;
        return t;
    }
}

// TEST-ONLY switch (orc_set_scalar_shapes): reproduce the branches of the reference that only its NON-differentiable
// variants take, so that the oracle can be compared with the scalar_* build of the reference (oracle/_ref/scalar):
//   rectangle.cpp:549-553  the hit is re-projected onto the rectangle ("to improve accuracy"); the *_ad_* variants -- the
//                          parity target of the kernels -- take p = ray(t) (:541-546).
//   diffuse.cpp:108-111    `none_or<false>(active)` is an early exit in scalar variants only: a hit on the back side
//                          returns a zeroed sample (pdf 0); JIT variants fill pdf / wo and mask the value (volpathmis
//                          stops on pdf == 0, so the two differ by one Russian-roulette draw there).
// Off by default; the kernels and every other oracle entry point follow the llvm_ad_* branches.
static bool g_scalar_shapes = false;

static inline SurfaceInteraction ray_intersect(const Scene &sc, const Ray &ray) {
    SurfaceInteraction si;
    float best = Inf, best_u = 0.f, best_v = 0.f; int best_shape = -1, best_face = 0;
    for (int i = 0; i < sc.d.n_shapes; ++i) {
        int face = 0;
        float t, bu = 0.f, bv = 0.f;
        if (sc.d.shapes[i].type == MIB_SHAPE_MESH)
            // (the search is cut at the nearest hit of the shapes before this one: same result -- a later shape only wins with a
            //  strictly smaller t -- and the kernels prune their traversal the same way)
            t = intersect_mesh(sc.d.shapes[i], sc.buffers[sc.d.shapes[i].mesh_buffer], ray, minimum(ray.maxt, best), &face, &bu, &bv);
        else
            t = intersect_shape(sc, i, ray, &face);
        if (t < best) { best = t; best_shape = i; best_face = face; best_u = bu; best_v = bv; }
    }
    if (best_shape < 0) { si.wi = -ray.d; return si; }
    const mib_shape &s = sc.d.shapes[best_shape];
    si.t = best; si.shape = best_shape;
    si.p = ray.at(best);
    Vec3 dp_du;
    if (s.type == MIB_SHAPE_MESH) {
        dp_du = mesh_surface_interaction(s, sc.buffers[s.mesh_buffer], best_face, best_u, best_v, &si.p, &si.n, &si.sh_frame.n, si.uv);
        finalize_si(si, dp_du, ray);
        return si;
    }
    if (s.type == MIB_SHAPE_RECTANGLE) {
        {   // rectangle.cpp:559-560: uv = prim_uv * 0.5 + 0.5, prim_uv = the local hit position of the intersection routine
            const Mat4 &T = sc.shape_to_object[best_shape];
            Vec3 local = fmadd(transform_vector(T, ray.d), best, transform_point_affine(T, ray.o));
            si.uv[0] = fmadd(local.x, .5f, .5f); si.uv[1] = fmadd(local.y, .5f, .5f);
        }
        si.n = Vec3(s.frame_n[0], s.frame_n[1], s.frame_n[2]);
        dp_du = Vec3(s.frame_s[0], s.frame_s[1], s.frame_s[2]);
        if (g_scalar_shapes) {
            Vec3 tr(s.to_world[3], s.to_world[7], s.to_world[11]), w = tr - si.p;
            float dist = (w.x * si.n.x + w.y * si.n.y) + w.z * si.n.z;        // scalar Array dot: left to right, no fma
            si.p = Vec3(si.p.x + dist * si.n.x, si.p.y + dist * si.n.y, si.p.z + dist * si.n.z);
        }
    } else if (s.type == MIB_SHAPE_CUBE) {
        si.n = sc.cube_normals[best_shape][best_face];
        dp_du = Vec3(0.f);
    } else {
        si.n = normalize(si.p - Vec3(s.center[0], s.center[1], s.center[2]));
        dp_du = Vec3(0.f);
    }
    si.sh_frame.n = si.n;
    finalize_si(si, dp_du, ray);
    return si;
}

// interaction.h:285-299
static inline int target_medium(const Scene &sc, const SurfaceInteraction &si, Vec3 d) {
    const mib_shape &s = sc.d.shapes[si.shape];
    return dot(d, si.n) > 0.f ? s.exterior_medium : s.interior_medium;
}
static inline bool is_medium_transition(const Scene &sc, const SurfaceInteraction &si) {
    const mib_shape &s = sc.d.shapes[si.shape];
    return s.interior_medium >= 0 || s.exterior_medium >= 0;
}

// ------------------------------------------------------------------
// Volumes (src/volumes/grid.cpp:321-356,545-575; drjit/texture.h:346-412,1143-1172,1270-1294)
// ------------------------------------------------------------------
static inline int wrap_coord(int pos, int shape, int mode) {
    if (mode == MIB_WRAP_CLAMP) return std::max(std::min(pos, shape - 1), 0);
    int shift_neg = pos < 0 ? pos + 1 : pos;
    int div = shift_neg / shape;          // divisor<int32_t> truncating division
    int mod = pos - div * shape;
    if (mod < 0) mod += shape;
    if (mode == MIB_WRAP_MIRROR)
        mod = (((div & 1) == 0) ^ (pos < 0)) ? mod : shape - 1 - mod;
    return mod;
}

template <int NC>
static inline Color<NC> eval_volume(const Scene &sc, const mib_volume &v, Vec3 p, uint64_t *n_lookups) {
    Color<NC> out;
    if (v.type == MIB_VOLUME_CONST) {
        for (int c = 0; c < NC; ++c) out.v[c] = v.value[v.channels == 1 ? 0 : c];
        return out;
    }
    if (n_lookups) (*n_lookups)++;
    Mat4 T = mat4_from(v.to_local);
    Vec3 pl = transform_point(T, p);
    const float *data = sc.buffers[v.buffer];
    const int ch = v.channels;
    float tmp[3] = { 0.f, 0.f, 0.f };
    const int rx = v.res[0], ry = v.res[1], rz = v.res[2];
    if (v.filter == MIB_FILTER_NEAREST) {
        float fx = pl.x * (float) rx, fy = pl.y * (float) ry, fz = pl.z * (float) rz;
        int ix = wrap_coord((int) std::floor(fx), rx, v.wrap), iy = wrap_coord((int) std::floor(fy), ry, v.wrap),
            iz = wrap_coord((int) std::floor(fz), rz, v.wrap);
        uint32_t idx = (((uint32_t) iz * ry + iy) * rx + ix) * ch;
        for (int c = 0; c < ch; ++c) tmp[c] = data[idx + c];
    } else {
        float fx = fmadd(pl.x, (float) rx, -.5f), fy = fmadd(pl.y, (float) ry, -.5f), fz = fmadd(pl.z, (float) rz, -.5f);
        int ix = (int) std::floor(fx), iy = (int) std::floor(fy), iz = (int) std::floor(fz);
        float w1x = fx - (float) ix, w1y = fy - (float) iy, w1z = fz - (float) iz;
        float w0x = 1.f - w1x, w0y = 1.f - w1y, w0z = 1.f - w1z;
        int x0 = wrap_coord(ix, rx, v.wrap), x1 = wrap_coord(ix + 1, rx, v.wrap);
        int y0 = wrap_coord(iy, ry, v.wrap), y1 = wrap_coord(iy + 1, ry, v.wrap);
        int z0 = wrap_coord(iz, rz, v.wrap), z1 = wrap_coord(iz + 1, rz, v.wrap);
        const int xs[2] = { x0, x1 }, ys[2] = { y0, y1 }, zs[2] = { z0, z1 };
        const float wx[2] = { w0x, w1x }, wy[2] = { w0y, w1y }, wz[2] = { w0z, w1z };
        for (int k = 0; k < 8; ++k) {
            int a = k & 1, b = (k >> 1) & 1, c3 = (k >> 2) & 1;
            uint32_t idx = (((uint32_t) zs[c3] * ry + ys[b]) * rx + xs[a]) * ch;
            float w = wx[a] * wy[b] * wz[c3];
            for (int c = 0; c < ch; ++c) tmp[c] = fmadd(data[idx + c], w, tmp[c]);
        }
    }
    if (ch == 1) { for (int c = 0; c < NC; ++c) out.v[c] = tmp[0]; }
    else if (NC == 3) { for (int c = 0; c < 3; ++c) out.v[c] = tmp[c]; }
    else { // luminance (include/mitsuba/core/spectrum.h luminance(Color3f))
        out.v[0] = tmp[0] * 0.212671f + tmp[1] * 0.715160f + tmp[2] * 0.072169f;
    }
    return out;
}

// Trilinear taps (indices into the data buffer for channel 0, weights) of a grid lookup: the adjoint
// scatters d(value)/d(voxel_k) = w_k through the same 8 taps (drjit/texture.h:366-408).
static inline int grid_taps(const mib_volume &v, Vec3 p, uint32_t *idx, float *w) {
    Mat4 T = mat4_from(v.to_local);
    Vec3 pl = transform_point(T, p);
    const int ch = v.channels, rx = v.res[0], ry = v.res[1], rz = v.res[2];
    if (v.filter == MIB_FILTER_NEAREST) {
        int ix = wrap_coord((int) std::floor(pl.x * (float) rx), rx, v.wrap), iy = wrap_coord((int) std::floor(pl.y * (float) ry), ry, v.wrap),
            iz = wrap_coord((int) std::floor(pl.z * (float) rz), rz, v.wrap);
        idx[0] = (((uint32_t) iz * ry + iy) * rx + ix) * ch; w[0] = 1.f;
        return 1;
    }
    float fx = fmadd(pl.x, (float) rx, -.5f), fy = fmadd(pl.y, (float) ry, -.5f), fz = fmadd(pl.z, (float) rz, -.5f);
    int ix = (int) std::floor(fx), iy = (int) std::floor(fy), iz = (int) std::floor(fz);
    float w1x = fx - (float) ix, w1y = fy - (float) iy, w1z = fz - (float) iz;
    float w0x = 1.f - w1x, w0y = 1.f - w1y, w0z = 1.f - w1z;
    const int xs[2] = { wrap_coord(ix, rx, v.wrap), wrap_coord(ix + 1, rx, v.wrap) },
              ys[2] = { wrap_coord(iy, ry, v.wrap), wrap_coord(iy + 1, ry, v.wrap) },
              zs[2] = { wrap_coord(iz, rz, v.wrap), wrap_coord(iz + 1, rz, v.wrap) };
    const float wx[2] = { w0x, w1x }, wy[2] = { w0y, w1y }, wz[2] = { w0z, w1z };
    for (int k = 0; k < 8; ++k) {
        int a = k & 1, b = (k >> 1) & 1, c3 = (k >> 2) & 1;
        idx[k] = (((uint32_t) zs[c3] * ry + ys[b]) * rx + xs[a]) * ch;
        w[k] = wx[a] * wy[b] * wz[c3];
    }
    return 8;
}

// bbox.h:303-326
static inline bool bbox_ray_intersect(const float *bmin, const float *bmax, const Ray &ray, float *mint, float *maxt) {
    bool active = true;
    float t1p[3], t2p[3];
    for (int a = 0; a < 3; ++a) {
        active = active && ((ray.d[a] != 0.f) || (ray.o[a] > bmin[a]) || (ray.o[a] < bmax[a]));
        float d_rcp = rcp(ray.d[a]);
        float t1 = (bmin[a] - ray.o[a]) * d_rcp, t2 = (bmax[a] - ray.o[a]) * d_rcp;
        t1p[a] = minimum(t1, t2); t2p[a] = maximum(t1, t2);
    }
    *mint = maximum(maximum(t1p[0], t1p[1]), t1p[2]);
    *maxt = minimum(minimum(t2p[0], t2p[1]), t2p[2]);
    return active && *maxt >= *mint;
}

// ------------------------------------------------------------------
// Medium interaction (src/render/medium.cpp:40-111, homogeneous.cpp:156-184, heterogeneous.cpp:180-204)
// ------------------------------------------------------------------
template <int NC> struct MediumInteraction {
    float t = Inf, mint = 0.f;
    Vec3 p, wi;
    Frame sh_frame;
    Color<NC> sigma_s, sigma_n, sigma_t, combined_extinction;
    int medium = -1;
    bool is_valid() const { return t != Inf; }
};

template <int NC> static inline float index_spectrum(const Color<NC> &s, uint32_t ch) { return NC == 3 ? s.v[ch] : s.v[0]; }

template <int NC>
static inline MediumInteraction<NC> sample_interaction(const Scene &sc, int medium_idx, const Ray &ray, float sample,
                                                       uint32_t channel, uint64_t *n_lookups) {
    const mib_medium &m = sc.d.media[medium_idx];
    MediumInteraction<NC> mei;
    mei.sh_frame = Frame(ray.d);
    mei.wi = Vec3(0.f, 0.f, -1.f);
    bool active = true;
    float mint, maxt;
    if (m.type == MIB_MEDIUM_HOMOGENEOUS) { mint = 0.f; maxt = Inf; }
    else active = bbox_ray_intersect(m.sigma_t.bbox_min, m.sigma_t.bbox_max, ray, &mint, &maxt);
    active = active && (std::isfinite(mint) || std::isfinite(maxt));
    if (!active) { mint = 0.f; maxt = Inf; }
    mint = maximum(0.f, mint);
    maxt = minimum(ray.maxt, maxt);
    Color<NC> combined;
    if (m.type == MIB_MEDIUM_HOMOGENEOUS) {
        // get_majorant = eval_sigmat & active (homogeneous.cpp:156-168)
        Color<NC> st = eval_volume<NC>(sc, m.sigma_t, Vec3(0.f), nullptr) * m.scale;
        combined = active ? st : Color<NC>(0.f);
    } else {
        combined = Color<NC>(m.max_density);
    }
    float mj = index_spectrum(combined, channel);
    float sampled_t = mint + (-log_(1.f - sample) / mj);
    bool valid_mi = active && (sampled_t <= maxt);
    mei.t = valid_mi ? sampled_t : Inf;
    mei.p = ray.at(sampled_t);
    mei.medium = medium_idx;
    mei.mint = mint;
    if (m.type == MIB_MEDIUM_HOMOGENEOUS) {
        Color<NC> st = eval_volume<NC>(sc, m.sigma_t, mei.p, nullptr) * m.scale;
        Color<NC> ss = st * eval_volume<NC>(sc, m.albedo, mei.p, valid_mi ? n_lookups : nullptr);
        mei.sigma_s = valid_mi ? ss : Color<NC>(0.f);
        mei.sigma_n = Color<NC>(0.f);
        mei.sigma_t = valid_mi ? st : Color<NC>(0.f);
    } else {
        // masked gathers return 0 for inactive lanes
        Color<NC> sv = valid_mi ? eval_volume<NC>(sc, m.sigma_t, mei.p, n_lookups) : Color<NC>(0.f);
        Color<NC> st = m.scale * sv;
        Color<NC> av = (valid_mi || m.albedo.type == MIB_VOLUME_CONST) ? eval_volume<NC>(sc, m.albedo, mei.p, n_lookups) : Color<NC>(0.f);
        mei.sigma_s = st * av;
        mei.sigma_n = Color<NC>(m.max_density) - st;
        mei.sigma_t = st;
    }
    mei.combined_extinction = combined;
    return mei;
}

} // namespace orc

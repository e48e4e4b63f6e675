// dscene.cuh -- device-side scene: analytic shapes, media, grid lookups, tables.
//
// The whole scene description travels as ONE __grid_constant__ kernel parameter
// (a few KiB, served by the constant cache): plugin types are enum tags, there
// is no virtual dispatch on the device.
#pragma once
#include "dmath.cuh"
#include "../../include/mi_b200.h"
#include "dmesh.cuh"

namespace mib {

// include/mitsuba/core/distr_1d.h:618-989, prepared on the host (mib_scene_commit)
struct DTable {
    const float *nodes, *pdf, *cdf;
    uint32_t n, valid_x, valid_y;
    float integral, normalization, range_x, range_y;
};

struct DScene {
    mib_scene_desc d;
    const float *buffers[MIB_MAX_BUFFERS];
    float *grad[MIB_MAX_BUFFERS];                 // gradient accumulators (adjoint), same shape as buffers
    DTable tables[MIB_MAX_PHASES][4];             // m11, m12, m33, m34
    float cube_normals[MIB_MAX_SHAPES][6][3];     // world-space outward face normals (-x,+x,-y,+y,-z,+z)
    double *grad_scalar;                          // [MIB_MAX_MEDIA*4] constant-volume gradients (adjoint)
    int prb_null, prb_homog;                      // prepare_scene flags (prbvolpath.py:76-89)
    float rf_coeff[10];                           // gaussian reconstruction filter polynomial (gaussian.cpp:50-98)
    float rf_inv_radius;                          // tent filter
    const float *weight_film;                     // [H*W] filter-weight film for the adjoint of non-box filters
    // "fast" mode (mib_integrator.majorant_res > 0): per heterogeneous medium a supergrid of local
    // majorants, super[m][(z*M + y)*M + x] = scale * max of the sigma_t texels any trilinear look-up inside the cell can touch
    const float *super[MIB_MAX_MEDIA];
    int super_res;                                // M (0: reference behaviour, one global majorant)
    int has_area_emitter;                         // pool kernels carry the last scattering position only when it is needed
    // Shading frames of the planar faces (rectangle: face 0; cube: 6 faces): s(3), t(3).  They are functions of scene constants
    // only, so commit() evaluates them once ON THE DEVICE with the arithmetic of si_finalize2 (k_shape_frames) and every
    // intersection reads them back instead of redoing two normalisations and a cross product (identical floats).
    float face_frame[MIB_MAX_SHAPES][6][6];
    float phase_const[MIB_MAX_PHASES][4];          // rayleigh: the three depolarization factors of mueller.h:311-328
    int frames_ready;
};

// The scene of the running launch.  One copy per translation unit (each kernel instantiation is its own object
// file); the launchers in inst.cu / mib_api.cu upload it on the launch stream right before the kernel.  Device
// code anywhere -- inlined or a real subroutine -- reads it through the constant bank (operands like c[0x3][..]).
static __constant__ DScene sc;

struct Ray { V3 o, d; float maxt; };
MIB_D V3 ray_at(const Ray &r, float t) { return vfma(r.d, t, r.o); }

struct Counters { unsigned long long n_lookups, n_real, n_null, n_nee_steps, n_surface; };

// ------------------------------------------------------------------
// Surface interaction (include/mitsuba/render/interaction.h)
// ------------------------------------------------------------------
struct SurfIt {
    float t;
    V3 p, n, wi;
    Frame3 sh;
    int shape, face;
    float tu, tv;          // texture coordinates (meshes, rectangles; only with FT_MESH)
};
MIB_D bool valid(const SurfIt &si) { return si.t != inf_f(); }

// interaction.h:133-162
MIB_D V3 offset_p(V3 p, V3 n, V3 d) {
    float mag = (1.f + fmaxf(fmaxf(fabsf(p.x), fabsf(p.y)), fabsf(p.z))) * kRayEps;
    mag = mulsign_ad(mag, dot(n, d));
    return v3(fma_(mag, n.x, p.x), fma_(mag, n.y, p.y), fma_(mag, n.z, p.z));
}
MIB_D Ray spawn_ray(V3 p, V3 n, V3 d) { Ray r; r.o = offset_p(p, n, d); r.d = d; r.maxt = largest_f(); return r; }
MIB_D Ray spawn_ray_to(V3 p, V3 n, V3 target) {
    Ray r;
    r.o = offset_p(p, n, target - p);
    V3 d = target - r.o;
    float dist = norm(d);
    float inv = rcp_(dist);
    r.d = v3(d.x * inv, d.y * inv, d.z * inv);
    r.maxt = dist * (1.f - kShadowEps);
    return r;
}

// Analytic intersection.  rectangle: src/shapes/rectangle.cpp:449-470 (+ :497-568, AD branch p = ray(t));
// cube: slab test in object space; sphere: quadratic in object space.
MIB_D float intersect_shape(int idx, const Ray &ray, int &face) {
    const mib_shape &s = sc.d.shapes[idx];
    V3 o = xform_point_affine(s.to_object, ray.o), d = xform_vector(s.to_object, ray.d);
    face = 0;
    // (shape kinds outside this instantiation's feature set never occur: the host picks a covering instantiation)
    const bool is_rect = MIB_FT(FT_RECT) && (s.type == MIB_SHAPE_RECTANGLE || !(MIB_FT(FT_CUBE) || MIB_FT(FT_SPHERE)));
    const bool is_cube = MIB_FT(FT_CUBE) && (s.type == MIB_SHAPE_CUBE || !MIB_FT(FT_SPHERE));
    if (is_rect) {
        float t = div_(-o.z, d.z);
        float lx = fma_(d.x, t, o.x), ly = fma_(d.y, t, o.y);
        bool hit = t >= 0.f && t <= ray.maxt && fabsf(lx) <= 1.f && fabsf(ly) <= 1.f;
        return hit ? t : inf_f();
    } else if (is_cube) {
        float tn = -inf_f(), tf = inf_f();
        int fn = 0, ff = 0;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            float oa = comp(o, a), inv = rcp_(comp(d, a));
            float t1 = (-1.f - oa) * inv, t2 = (1.f - oa) * inv;
            int f1 = 2 * a, f2 = 2 * a + 1;
            if (t1 > t2) { float tt = t1; t1 = t2; t2 = tt; int ft = f1; f1 = f2; f2 = ft; }
            if (t1 > tn) { tn = t1; fn = f1; }
            if (t2 < tf) { tf = t2; ff = f2; }
        }
        if (!(tn <= tf)) return inf_f();
        float t;
        if (tn >= 0.f) { t = tn; face = fn; } else { t = tf; face = ff; }
        if (!(t >= 0.f && t <= ray.maxt)) return inf_f();
        return t;
    } else {
        float A = sqnorm(d), B = 2.f * dot(o, d), C = sqnorm(o) - 1.f;
        float disc = fms_(B, B, 4.f * A * C);
        if (!(disc >= 0.f)) return inf_f();
        float sq = sqrt_(disc);
        float q = -0.5f * (B + (B >= 0.f ? sq : -sq));
        float t0 = div_(q, A), t1 = div_(C, q);
        if (t0 > t1) { float tt = t0; t0 = t1; t1 = tt; }
        float t = t0 >= 0.f ? t0 : t1;
        if (!(t >= 0.f && t <= ray.maxt)) return inf_f();
        return t;
    }
}

// finalize_surface_interaction (interaction.h:245-255, :464-484): n, shading frame and wi as functions of
// (shape, face, p, ray.d) only -- the pool kernel keeps just those in shared memory and rebuilds the rest.
// shading frame from the geometric normal and dp/du (interaction.h:464-484 with the rectangle / cube / sphere partials)
MIB_D void frame_from_normal(V3 n, V3 dp_du, V3 &fs, V3 &ft) {
    fs = normalize(vfma(n, -dot(n, dp_du), dp_du));
    if (dp_du.x == 0.f && dp_du.y == 0.f && dp_du.z == 0.f) { V3 s2, t2; coordinate_system(n, s2, t2); fs = s2; }
    ft = cross(n, fs);
}
MIB_D void face_normal_dpdu(int shape, int face, V3 &n, V3 &dp_du) {
    const mib_shape &s = sc.d.shapes[shape];
    dp_du = v3(0.f);
    if (MIB_FT(FT_RECT) && (s.type == MIB_SHAPE_RECTANGLE || !MIB_FT(FT_CUBE))) {
        n = v3(s.frame_n[0], s.frame_n[1], s.frame_n[2]);
        dp_du = v3(s.frame_s[0], s.frame_s[1], s.frame_s[2]);
    } else {
        const float *nn = sc.cube_normals[shape][face];
        n = v3(nn[0], nn[1], nn[2]);
    }
}
MIB_SUB_GEOM SurfIt si_finalize2(int shape, int face, float t, V3 p, V3 ray_d) {
    SurfIt si;
    si.t = t; si.shape = shape; si.face = face; si.p = p;
    const mib_shape &s = sc.d.shapes[shape];
    V3 dp_du = v3(0.f);
    if (MIB_FT(FT_SPHERE) && s.type == MIB_SHAPE_SPHERE) {
        si.n = normalize(si.p - v3(s.center[0], s.center[1], s.center[2]));
        si.sh.n = si.n;
        frame_from_normal(si.n, dp_du, si.sh.s, si.sh.t);
    } else {
        face_normal_dpdu(shape, face, si.n, dp_du);
        si.sh.n = si.n;
        if (sc.frames_ready) {
            const float *ff = sc.face_frame[shape][face];
            si.sh.s = v3(ff[0], ff[1], ff[2]);
            si.sh.t = v3(ff[3], ff[4], ff[5]);
        } else frame_from_normal(si.n, dp_du, si.sh.s, si.sh.t);
    }
    si.wi = to_local(si.sh, -ray_d);
    return si;
}
MIB_D void si_finalize(V3 ray_d, SurfIt &si) { si = si_finalize2(si.shape, si.face, si.t, si.p, ray_d); }

// Scene::ray_intersect: nearest hit over the (few) shapes of the scene -- analytic shapes, and triangle meshes (dmesh.cuh)
MIB_SUB_GEOM SurfIt ray_intersect2(Ray ray) {
    float best = inf_f(), bu = 0.f, bv = 0.f;
    int bs = -1, bf = 0;
    for (int i = 0; i < sc.d.n_shapes; ++i) {
        int face = 0;
        float t, u = 0.f, v = 0.f;
        if (MIB_FT(FT_MESH) && sc.d.shapes[i].type == MIB_SHAPE_MESH)
            // the search is cut at the nearest hit of the shapes before this one (a later shape only wins with a strictly smaller t)
            t = intersect_mesh(sc.d.shapes[i], sc.buffers[sc.d.shapes[i].mesh_buffer], ray.o, ray.d, fminf(ray.maxt, best), face, u, v);
        else
            t = intersect_shape(i, ray, face);
        if (t < best) { best = t; bs = i; bf = face; bu = u; bv = v; }
    }
    if (bs < 0) {
        SurfIt si;
        si.t = best; si.shape = -1; si.face = 0; si.tu = 0.f; si.tv = 0.f;
        si.p = v3(0.f); si.n = v3(0.f); si.sh.s = v3(0.f); si.sh.t = v3(0.f); si.sh.n = v3(0.f);
        si.wi = -ray.d;
        return si;
    }
    if (MIB_FT(FT_MESH) && sc.d.shapes[bs].type == MIB_SHAPE_MESH) {
        SurfIt si;
        si.t = best; si.shape = bs; si.face = bf;
        const V3 dp_du = mesh_surface_interaction(sc.d.shapes[bs], sc.buffers[sc.d.shapes[bs].mesh_buffer], bf, bu, bv, si.p, si.n, si.sh.n, si.tu, si.tv);
        frame_from_normal(si.sh.n, dp_du, si.sh.s, si.sh.t);
        si.wi = to_local(si.sh, -ray.d);
        return si;
    }
    SurfIt si = si_finalize2(bs, bf, best, ray_at(ray, best), ray.d);
    si.tu = 0.f; si.tv = 0.f;
    if (MIB_FT(FT_MESH) && sc.d.shapes[bs].type == MIB_SHAPE_RECTANGLE) {
        // rectangle.cpp:559-560: uv = prim_uv * 0.5 + 0.5, prim_uv = the local hit position of the intersection routine
        const mib_shape &s = sc.d.shapes[bs];
        const V3 o = xform_point_affine(s.to_object, ray.o), d = xform_vector(s.to_object, ray.d);
        si.tu = fma_(fma_(d.x, best, o.x), .5f, .5f);
        si.tv = fma_(fma_(d.y, best, o.y), .5f, .5f);
    }
    return si;
}
MIB_D void ray_intersect(const Ray &ray, SurfIt &si) { si = ray_intersect2(ray); }

MIB_D int target_medium(const SurfIt &si, V3 d) {              // interaction.h:285-299
    const mib_shape &s = sc.d.shapes[si.shape];
    return dot(d, si.n) > 0.f ? s.exterior_medium : s.interior_medium;
}
MIB_D bool is_medium_transition(const SurfIt &si) {
    const mib_shape &s = sc.d.shapes[si.shape];
    return s.interior_medium >= 0 || s.exterior_medium >= 0;
}

// ------------------------------------------------------------------
// Grid lookups (drjit/texture.h:346-412, :1143-1172, :1270-1294; grid.cpp:545-575)
// The 8 taps are plain read-only loads served from the 126 MB L2 (a 256^3 fp32
// grid is 64 MiB and stays resident); accumulation order 000,100,010,...,111.
// ------------------------------------------------------------------
MIB_D int wrap_coord(int pos, int shape, int mode) {
    if (mode == MIB_WRAP_CLAMP) return max(min(pos, shape - 1), 0);
    int sn = pos < 0 ? pos + 1 : pos;
    int dv = sn / shape;
    int md = pos - dv * shape;
    if (md < 0) md += shape;
    if (mode == MIB_WRAP_MIRROR) md = (((dv & 1) == 0) != (pos < 0)) ? md : shape - 1 - md;
    return md;
}

struct GridTaps { uint32_t idx[8]; float w[8]; };   // used by the adjoint to scatter through the same taps

template <int NC>
MIB_DH void eval_grid(const mib_volume &v, V3 p, float *out, GridTaps *taps) {
    V3 pl = xform_point(v.to_local, p);
    const float *__restrict__ data = sc.buffers[v.buffer];
    const int ch = v.channels, rx = v.res[0], ry = v.res[1], rz = v.res[2];
    float tmp[3] = { 0.f, 0.f, 0.f };
    if (v.filter == MIB_FILTER_NEAREST) {
        int ix = wrap_coord((int) floorf(pl.x * (float) rx), rx, v.wrap), iy = wrap_coord((int) floorf(pl.y * (float) ry), ry, v.wrap),
            iz = wrap_coord((int) floorf(pl.z * (float) rz), rz, v.wrap);
        uint32_t idx = (((uint32_t) iz * ry + iy) * rx + ix) * ch;
        for (int c = 0; c < ch; ++c) tmp[c] = __ldg(data + idx + c);
        if (taps) { for (int k = 0; k < 8; ++k) { taps->idx[k] = idx; taps->w[k] = k == 0 ? 1.f : 0.f; } }
    } else {
        float fx = fma_(pl.x, (float) rx, -.5f), fy = fma_(pl.y, (float) ry, -.5f), fz = fma_(pl.z, (float) rz, -.5f);
        float flx = floorf(fx), fly = floorf(fy), flz = floorf(fz);
        int ix = (int) flx, iy = (int) fly, iz = (int) flz;
        float w1x = fx - flx, w1y = fy - fly, w1z = fz - flz;
        float w0x = 1.f - w1x, w0y = 1.f - w1y, w0z = 1.f - w1z;
        int x0 = wrap_coord(ix, rx, v.wrap), x1 = wrap_coord(ix + 1, rx, v.wrap);
        int y0 = wrap_coord(iy, ry, v.wrap), y1 = wrap_coord(iy + 1, ry, v.wrap);
        int z0 = wrap_coord(iz, rz, v.wrap), z1 = wrap_coord(iz + 1, rz, v.wrap);
        uint32_t r00 = ((uint32_t) z0 * ry + y0) * rx, r10 = ((uint32_t) z0 * ry + y1) * rx,
                 r01 = ((uint32_t) z1 * ry + y0) * rx, r11 = ((uint32_t) z1 * ry + y1) * rx;
        uint32_t id[8] = { (r00 + x0) * ch, (r00 + x1) * ch, (r10 + x0) * ch, (r10 + x1) * ch,
                           (r01 + x0) * ch, (r01 + x1) * ch, (r11 + x0) * ch, (r11 + x1) * ch };
        float w[8] = { w0x * w0y * w0z, w1x * w0y * w0z, w0x * w1y * w0z, w1x * w1y * w0z,
                       w0x * w0y * w1z, w1x * w0y * w1z, w0x * w1y * w1z, w1x * w1y * w1z };
        if (ch == 1) {
            // issue the 8 independent loads first so they overlap in the memory pipeline
            float val[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) val[k] = __ldg(data + id[k]);
#pragma unroll
            for (int k = 0; k < 8; ++k) tmp[0] = fma_(val[k], w[k], tmp[0]);
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k)
                for (int c = 0; c < 3; ++c) tmp[c] = fma_(__ldg(data + id[k] + c), w[k], tmp[c]);
        }
        if (taps) {
#pragma unroll
            for (int k = 0; k < 8; ++k) { taps->idx[k] = id[k]; taps->w[k] = w[k]; }
        }
    }
    if (ch == 1) { for (int c = 0; c < NC; ++c) out[c] = tmp[0]; }
    else if (NC == 3) { out[0] = tmp[0]; out[1] = tmp[1]; out[2] = tmp[2]; }
    else out[0] = tmp[0] * 0.212671f + tmp[1] * 0.715160f + tmp[2] * 0.072169f;
}

template <int NC>
MIB_D void eval_volume(const mib_volume &v, V3 p, float *out, Counters &cnt, GridTaps *taps = nullptr) {
    if (!MIB_FT(FT_GRID) || v.type == MIB_VOLUME_CONST) {
#pragma unroll
        for (int c = 0; c < NC; ++c) out[c] = v.value[v.channels == 1 ? 0 : c];
        return;
    }
    cnt.n_lookups++;
    eval_grid<NC>(v, p, out, taps);
}

// include/mitsuba/core/bbox.h:303-326
MIB_D bool bbox_intersect(const float *bmin, const float *bmax, const Ray &ray, float &mint, float &maxt) {
    bool active = true;
    float lo = -inf_f(), hi = inf_f();
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float o = comp(ray.o, a), d = comp(ray.d, a);
        active = active && ((d != 0.f) || (o > bmin[a]) || (o < bmax[a]));
        float dr = rcp_(d);
        float t1 = (bmin[a] - o) * dr, t2 = (bmax[a] - o) * dr;
        float t1p = fminf(t1, t2), t2p = fmaxf(t1, t2);
        lo = a == 0 ? t1p : fmaxf(lo, t1p);
        hi = a == 0 ? t2p : fminf(hi, t2p);
    }
    mint = lo; maxt = hi;
    return active && maxt >= mint;
}

// ------------------------------------------------------------------
// Medium interaction (src/render/medium.cpp:40-86)
// ------------------------------------------------------------------
template <int NC> struct MedIt {
    float t, mint;
    V3 p;
    float sigma_s[NC], sigma_n[NC], sigma_t[NC], maj[NC];
    int medium;
};
template <int NC> MIB_D bool valid(const MedIt<NC> &m) { return m.t != inf_f(); }
template <int NC> MIB_D float ch(const float *s, uint32_t c) { return NC == 3 ? s[c] : s[0]; }

// Free-flight sampling through the majorant supergrid ("fast" mode; no counterpart in the fork, whose heterogeneous medium
// uses one global majorant, heterogeneous.cpp:163-185).  3D-DDA over the M^3 cells of the medium's local unit cube: the
// optical-depth target tau = -log(1 - u) is spent cell by cell at the cell's local majorant; the collision lands where it
// runs out.  Unbiased for any majorant that bounds sigma_t inside the cell; consumes one random number like the reference's
// sample_interaction, but the tentative collisions fall elsewhere, so results agree with parity mode only statistically.
struct DdaHit { float t, maj; bool hit; };
// Fast mode only: Russian roulette on the transmittance of a shadow walk (ratio tracking never terminates by itself in an
// optically thick region; the reference lets it run).  Unbiased; consumes one extra random number per step below the threshold.
constexpr float kNeeRouletteThreshold = 0.1f;
MIB_SUB_GEOM DdaHit supergrid_sample(int mi, Ray ray, float mint, float maxt, float tau) {
    const mib_medium &m = sc.d.media[mi];
    const mib_volume &v = m.sigma_t;
    const int M = sc.super_res;
    const float *__restrict__ sg = sc.super[mi];
    const float fM = (float) M, cell = 1.f / fM;
    V3 lo = xform_point_affine(v.to_local, ray_at(ray, mint)), ld = xform_vector(v.to_local, ray.d);
    const float tlen = maxt - mint;
    int c[3], step[3];
    float tmax[3], tdelta[3];
    // (approximate reciprocals: fast mode has no bit-parity contract, and a cell boundary that moves by an ulp is covered
    // by the one-texel slack of the supergrid)
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float o = comp(lo, a), d = comp(ld, a);
        int ci = min(max((int) floorf(o * fM), 0), M - 1);
        c[a] = ci;
        const float inv = __fdividef(1.f, d);
        if (d > 0.f) { step[a] = 1; tmax[a] = ((float) (ci + 1) * cell - o) * inv; tdelta[a] = cell * inv; }
        else if (d < 0.f) { step[a] = -1; tmax[a] = ((float) ci * cell - o) * inv; tdelta[a] = -cell * inv; }
        else { step[a] = 0; tmax[a] = inf_f(); tdelta[a] = inf_f(); }
        tmax[a] = fmaxf(tmax[a], 0.f);
    }
    DdaHit r;
    r.t = inf_f(); r.maj = m.max_density; r.hit = false;
    float t_cur = 0.f;
    for (int it = 0; it < 3 * M + 3; ++it) {
        const float mu = __ldg(sg + ((c[2] * M + c[1]) * M + c[0]));
        const float t_next = fminf(tmax[0], fminf(tmax[1], tmax[2]));
        const float t_end = fminf(t_next, tlen);
        const float seg = fmaxf(t_end - t_cur, 0.f);
        if (mu > 0.f && mu * seg >= tau) { r.t = mint + (t_cur + __fdividef(tau, mu)); r.maj = mu; r.hit = true; break; }
        tau -= mu * seg;
        t_cur = t_end;
        if (t_next >= tlen) break;
        const int a = (tmax[0] <= tmax[1] && tmax[0] <= tmax[2]) ? 0 : (tmax[1] <= tmax[2] ? 1 : 2);
        // (selects instead of dynamic indexing keep the arrays in registers)
#pragma unroll
        for (int b = 0; b < 3; ++b) if (b == a) { c[b] += step[b]; tmax[b] += tdelta[b]; }
        if (c[0] < 0 || c[0] >= M || c[1] < 0 || c[1] >= M || c[2] < 0 || c[2] >= M) break;
    }
    return r;
}

template <bool TAPS> struct TapsOpt { GridTaps t; };       // the 8 taps travel with the result only for the adjoint
template <> struct TapsOpt<false> {};
template <int NC, bool TAPS> struct MedRes : TapsOpt<TAPS> {
    MedIt<NC> mei;
    int lookups;
};
template <int NC, bool TAPS>
MIB_SUB_GEOM MedRes<NC, TAPS> sample_interaction2(int mi, Ray ray, float sample, uint32_t channel) {
    MedRes<NC, TAPS> res;
    MedIt<NC> &mei = res.mei;
    Counters cnt = { 0, 0, 0, 0, 0 };
    GridTaps *taps = nullptr;
    if constexpr (TAPS) {
        taps = &res.t;
#pragma unroll
        for (int k = 0; k < 8; ++k) { res.t.idx[k] = 0; res.t.w[k] = 0.f; }
    }
    const mib_medium &m = sc.d.media[mi];
    // media kinds outside this instantiation's feature set never occur
    const bool homog = MIB_FT(FT_HOMOG) && (m.type == MIB_MEDIUM_HOMOGENEOUS || !MIB_FT(FT_GRID));
    bool active = true;
    float mint = 0.f, maxt = inf_f();
    if (!homog)
        active = bbox_intersect(m.sigma_t.bbox_min, m.sigma_t.bbox_max, ray, mint, maxt);
    active = active && (isfinite(mint) || isfinite(maxt));
    if (!active) { mint = 0.f; maxt = inf_f(); }
    mint = fmaxf(0.f, mint);
    maxt = fminf(ray.maxt, maxt);
    if (homog) {
#pragma unroll
        for (int c = 0; c < NC; ++c) mei.maj[c] = active ? m.sigma_t.value[m.sigma_t.channels == 1 ? 0 : c] * m.scale : 0.f;
    } else {
#pragma unroll
        for (int c = 0; c < NC; ++c) mei.maj[c] = m.max_density;
    }
    float mj = ch<NC>(mei.maj, channel);
    float sampled_t;
    bool valid_mi;
    float maj_local = m.max_density;
    if (MIB_FT(FT_FAST) && !homog && sc.super_res > 0 && sc.super[mi] != nullptr) {
        sampled_t = inf_f(); valid_mi = false;
        if (active && maxt > mint) {
            DdaHit h = supergrid_sample(mi, ray, mint, maxt, -log_(1.f - sample));
            valid_mi = h.hit && h.t <= maxt;
            sampled_t = valid_mi ? h.t : inf_f();
            maj_local = h.maj;
#pragma unroll
            for (int c = 0; c < NC; ++c) mei.maj[c] = maj_local;
        }
    } else {
        sampled_t = mint + div_(-log_(1.f - sample), mj);
        valid_mi = active && (sampled_t <= maxt);
    }
    mei.t = valid_mi ? sampled_t : inf_f();
    mei.p = ray_at(ray, sampled_t);
    mei.medium = mi;
    mei.mint = mint;
    if (homog) {
        float alb[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) alb[c] = 0.f;
        if (valid_mi || m.albedo.type == MIB_VOLUME_CONST) eval_volume<NC>(m.albedo, mei.p, alb, cnt);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            float st = m.sigma_t.value[m.sigma_t.channels == 1 ? 0 : c] * m.scale;
            mei.sigma_s[c] = valid_mi ? st * alb[c] : 0.f;
            mei.sigma_n[c] = 0.f;
            mei.sigma_t[c] = valid_mi ? st : 0.f;
        }
    } else {
        float sv[NC], av[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) { sv[c] = 0.f; av[c] = 0.f; }
        if (valid_mi) eval_volume<NC>(m.sigma_t, mei.p, sv, cnt, taps);
        if (valid_mi || m.albedo.type == MIB_VOLUME_CONST) eval_volume<NC>(m.albedo, mei.p, av, cnt);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            float st = m.scale * sv[c];
            mei.sigma_s[c] = st * av[c];
            mei.sigma_n[c] = maj_local - st;
            mei.sigma_t[c] = st;
        }
    }
    res.lookups = (int) cnt.n_lookups;
    return res;
}
template <int NC>
MIB_D void sample_interaction(int mi, const Ray &ray, float sample, uint32_t channel,
                              MedIt<NC> &mei, Counters &cnt, GridTaps *taps = nullptr) {
    if (taps) {
        MedRes<NC, true> r = sample_interaction2<NC, true>(mi, ray, sample, channel);
        mei = r.mei; cnt.n_lookups += r.lookups; *taps = r.t;
    } else {
        MedRes<NC, false> r = sample_interaction2<NC, false>(mi, ray, sample, channel);
        mei = r.mei; cnt.n_lookups += r.lookups;
    }
}

// ------------------------------------------------------------------
// Irregular 1D distribution (distr_1d.h:707-729, :781-814; drjit/util.h:172-220)
// Tables are a few KiB and stay in L1/L2; binary search = log2(N) dependent loads.
// ------------------------------------------------------------------
MIB_DH float tab_eval(const DTable &t, float x) {
    bool active = x >= t.range_x && x <= t.range_y;
    uint32_t start = 0, end = t.n;
    uint32_t iters = (start < end) ? (31 - __clz(end - start)) + 1 : 0;
    for (uint32_t i = 0; i < iters; ++i) {
        uint32_t mid = (start + end) >> 1;
        float nv = (active && mid < t.n) ? __ldg(t.nodes + mid) : 0.f;
        if (nv < x) start = min(mid + 1, end); else end = mid;
    }
    uint32_t index = max(min(start, t.n - 1u), 1u) - 1u;
    float x0 = __ldg(t.nodes + index), x1 = __ldg(t.nodes + index + 1), y0 = __ldg(t.pdf + index), y1 = __ldg(t.pdf + index + 1);
    x = div_(x - x0, x1 - x0);
    return active ? fma_(x, y1 - y0, y0) : 0.f;
}
MIB_DH float tab_sample(const DTable &t, float sample) {
    sample *= t.integral;
    uint32_t start = t.valid_x, end = t.valid_y;
    uint32_t iters = (start < end) ? (31 - __clz(end - start)) + 1 : 0;
    for (uint32_t i = 0; i < iters; ++i) {
        uint32_t mid = (start + end) >> 1;
        float value = __ldg(t.cdf + mid);
        bool pred = ((value < sample) || (value == 0.f)) && (value != t.integral);
        if (pred) start = min(mid + 1, end); else end = mid;
    }
    uint32_t index = start;
    float x0 = __ldg(t.nodes + index), x1 = __ldg(t.nodes + index + 1), y0 = __ldg(t.pdf + index), y1 = __ldg(t.pdf + index + 1);
    float c0 = index > 0 ? __ldg(t.cdf + index - 1) : 0.f, w = x1 - x0;
    sample = div_(sample - c0, w);
    float t_linear = div_(y0 - safe_sqrt_(sqr_(y0) + 2.f * sample * (y1 - y0)), y0 - y1);
    float t_const = div_(sample, y0);
    float tt = (y0 == y1) ? t_const : t_linear;
    return fma_(tt, w, x0);
}

} // namespace mib

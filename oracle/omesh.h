// omesh.h -- TEST INFRASTRUCTURE (CPU oracle).  Triangle meshes of the `obj` shape:
//   ray / triangle          include/mitsuba/render/mesh.h:431-453 (moeller_trumbore), :274-300
//   surface interaction     src/render/mesh.cpp:1368-1537 (barycentric re-interpolation, face normal, texcoords, dp_du from
//                           the uv parameterisation, interpolated shading normal)
// in their JIT forms (fma-chain dot, fmsub cross).  The acceleration structure is ours (mitsuba3_nasa_b200/mesh.py: binary BVH,
// layout in include/mi_b200.h): the reference hands meshes to Embree or its own kd-tree, and which structure finds the
// nearest triangle is not part of the arithmetic.  The CUDA kernels (csrc/dmesh.cuh) walk the same nodes in the same order.
#pragma once
#include <cstring>

namespace orc {

struct MeshView {
    const float *pos, *nrm, *uv;     // nrm / uv may be null
    const uint32_t *faces, *order;
    const float *nodes;
};
static inline MeshView mesh_view(const mib_shape &s, const float *buf) {
    MeshView m;
    size_t o = 0, V = (size_t) s.mesh_vertices, F = (size_t) s.mesh_faces, N = (size_t) s.mesh_nodes;
    m.pos = buf; o += 3 * V;
    m.nrm = nullptr; m.uv = nullptr;
    if (s.mesh_flags & MIB_MESH_HAS_NORMALS) { m.nrm = buf + o; o += 3 * V; }
    if (s.mesh_flags & MIB_MESH_HAS_TEXCOORDS) { m.uv = buf + o; o += 2 * V; }
    m.faces = reinterpret_cast<const uint32_t *>(buf + o); o += 3 * F;
    m.nodes = buf + o; o += 8 * N;
    m.order = reinterpret_cast<const uint32_t *>(buf + o);
    return m;
}
static inline size_t mesh_floats(const mib_shape &s) {
    size_t V = (size_t) s.mesh_vertices, F = (size_t) s.mesh_faces, N = (size_t) s.mesh_nodes;
    return 3 * V + ((s.mesh_flags & MIB_MESH_HAS_NORMALS) ? 3 * V : 0) + ((s.mesh_flags & MIB_MESH_HAS_TEXCOORDS) ? 2 * V : 0) + 3 * F + 8 * N + F;
}
static inline Vec3 mesh_vec3(const float *a, uint32_t i) { return Vec3(a[3 * i], a[3 * i + 1], a[3 * i + 2]); }

// mesh.h:431-453
static inline bool moeller_trumbore(const Ray &ray, Vec3 p0, Vec3 p1, Vec3 p2, float maxt, float *t_out, float *u_out, float *v_out) {
    Vec3 e1 = p1 - p0, e2 = p2 - p0;
    Vec3 pvec = cross(ray.d, e2);
    float inv_det = rcp(dot(e1, pvec));
    Vec3 tvec = ray.o - p0;
    float u = dot(tvec, pvec) * inv_det;
    bool active = u >= 0.f && u <= 1.f;
    Vec3 qvec = cross(tvec, e1);
    float v = dot(ray.d, qvec) * inv_det;
    active = active && v >= 0.f && u + v <= 1.f;
    float t = dot(e2, qvec) * inv_det;
    active = active && t >= 0.f && t <= maxt;
    *t_out = t; *u_out = u; *v_out = v;
    return active;
}

// nearest triangle of one mesh within [0, maxt]: returns t (Inf: none), face index and barycentrics
static inline float intersect_mesh(const mib_shape &s, const float *buf, const Ray &ray, float maxt, int *face, float *bu, float *bv) {
    MeshView m = mesh_view(s, buf);
    const float inv[3] = { rcp(ray.d.x), rcp(ray.d.y), rcp(ray.d.z) };
    const float org[3] = { ray.o.x, ray.o.y, ray.o.z };
    float best = Inf;
    uint32_t stack[64];
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0) {
        const uint32_t ni = stack[--sp];
        const float *nd = m.nodes + 8 * (size_t) ni;
        float tmin = 0.f, tmax = minimum(best, maxt);
        for (int a = 0; a < 3; ++a) {
            float t1 = (nd[a] - org[a]) * inv[a], t2 = (nd[3 + a] - org[a]) * inv[a];
            tmin = std::fmax(tmin, std::fmin(t1, t2));
            tmax = std::fmin(tmax, std::fmax(t1, t2));
        }
        if (!(tmin <= tmax)) continue;
        uint32_t a, b;
        std::memcpy(&a, nd + 6, 4); std::memcpy(&b, nd + 7, 4);
        if (b & 0x80000000u) {
            if (sp + 2 > 64) continue;                   // (deeper than any tree mesh.py builds)
            stack[sp++] = b & 0x7fffffffu;               // right child below, left child on top: left is visited first
            stack[sp++] = a;
        } else {
            for (uint32_t k = 0; k < b; ++k) {
                const uint32_t f = m.order[a + k];
                const uint32_t *fi = m.faces + 3 * (size_t) f;
                float t, u, v;
                if (moeller_trumbore(ray, mesh_vec3(m.pos, fi[0]), mesh_vec3(m.pos, fi[1]), mesh_vec3(m.pos, fi[2]), maxt, &t, &u, &v) && t < best) {
                    best = t; *face = (int) f; *bu = u; *bv = v;
                }
            }
        }
    }
    return best;
}

// mesh.cpp:1368-1537: fills p, n, uv, sh_frame.n and returns dp_du
static inline Vec3 mesh_surface_interaction(const mib_shape &s, const float *buf, int face, float b1, float b2, Vec3 *p, Vec3 *n, Vec3 *sh_n, float *uv) {
    MeshView m = mesh_view(s, buf);
    const uint32_t *fi = m.faces + 3 * (size_t) face;
    Vec3 p0 = mesh_vec3(m.pos, fi[0]), p1 = mesh_vec3(m.pos, fi[1]), p2 = mesh_vec3(m.pos, fi[2]);
    float b0 = 1.f - b1 - b2;
    *p = fmadd(p0, b0, fmadd(p1, b1, p2 * b2));
    Vec3 dp0 = p1 - p0, dp1 = p2 - p0;
    *n = normalize(cross(dp0, dp1));
    uv[0] = b1; uv[1] = b2;
    Vec3 dp_du, dp_dv;
    coordinate_system(*n, &dp_du, &dp_dv);
    if (m.uv) {
        const float *t0 = m.uv + 2 * (size_t) fi[0], *t1 = m.uv + 2 * (size_t) fi[1], *t2 = m.uv + 2 * (size_t) fi[2];
        uv[0] = fmadd(t2[0], b2, fmadd(t1[0], b1, t0[0] * b0));
        uv[1] = fmadd(t2[1], b2, fmadd(t1[1], b1, t0[1] * b0));
        float du0x = t1[0] - t0[0], du0y = t1[1] - t0[1], du1x = t2[0] - t0[0], du1y = t2[1] - t0[1];
        float det = fmsub(du0x, du1y, du0y * du1x), inv_det = rcp(det);
        if (det != 0.f)
            dp_du = Vec3(fmsub(du1y, dp0.x, du0y * dp1.x) * inv_det, fmsub(du1y, dp0.y, du0y * dp1.y) * inv_det, fmsub(du1y, dp0.z, du0y * dp1.z) * inv_det);
    }
    *sh_n = *n;
    if (m.nrm) {
        Vec3 n0 = mesh_vec3(m.nrm, fi[0]), n1 = mesh_vec3(m.nrm, fi[1]), n2 = mesh_vec3(m.nrm, fi[2]);
        Vec3 ns = fmadd(n2, b2, fmadd(n1, b1, n0 * b0));
        float il = rsqrt(squared_norm(ns));
        *sh_n = ns * il;
    }
    return dp_du;
}

} // namespace orc

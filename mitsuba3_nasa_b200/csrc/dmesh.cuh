// dmesh.cuh -- triangle meshes of the `obj` shape on the device.
//   ray / triangle          include/mitsuba/render/mesh.h:431-453 (moeller_trumbore)
//   surface interaction     src/render/mesh.cpp:1368-1537
// Buffer layout: include/mi_b200.h (`mib_shape`), built by mitsuba3_nasa_b200/mesh.py.  The BVH walk is the oracle's
// (oracle/omesh.h), node for node, so both pick the same triangle.  Only the one-thread-one-path kernels carry this code
// (feature FT_MESH); the pool kernels keep a hit as (t, p, shape | face) and are not offered mesh scenes.
#pragma once
#include "dmath.cuh"

namespace mib {

struct MeshView {
    const float *pos, *nrm, *uv, *nodes;
    const uint32_t *faces, *order;
};
MIB_D MeshView mesh_view(const mib_shape &s, const float *buf) {
    MeshView m;
    const size_t V = (size_t) s.mesh_vertices, F = (size_t) s.mesh_faces, N = (size_t) s.mesh_nodes;
    size_t o = 3 * V;
    m.pos = buf;
    m.nrm = nullptr; m.uv = nullptr;
    if (s.mesh_flags & MIB_MESH_HAS_NORMALS) { m.nrm = buf + o; o += 3 * V; }
    if (s.mesh_flags & MIB_MESH_HAS_TEXCOORDS) { m.uv = buf + o; o += 2 * V; }
    m.faces = reinterpret_cast<const uint32_t *>(buf + o); o += 3 * F;
    m.nodes = buf + o; o += 8 * N;
    m.order = reinterpret_cast<const uint32_t *>(buf + o);
    return m;
}
MIB_D V3 mesh_vec3(const float *a, uint32_t i) { return v3(__ldg(a + 3 * i), __ldg(a + 3 * i + 1), __ldg(a + 3 * i + 2)); }

// mesh.h:431-453
MIB_D bool moeller_trumbore(V3 ro, V3 rd, V3 p0, V3 p1, V3 p2, float maxt, float &t, float &u, float &v) {
    const V3 e1 = p1 - p0, e2 = p2 - p0;
    const V3 pvec = cross(rd, e2);
    const float inv_det = rcp_(dot(e1, pvec));
    const V3 tvec = ro - p0;
    u = dot(tvec, pvec) * inv_det;
    bool active = u >= 0.f && u <= 1.f;
    const V3 qvec = cross(tvec, e1);
    v = dot(rd, qvec) * inv_det;
    active = active && v >= 0.f && u + v <= 1.f;
    t = dot(e2, qvec) * inv_det;
    return active && t >= 0.f && t <= maxt;
}

// nearest triangle of one mesh within [0, maxt]: t (inf: none), face index, barycentrics
MIB_NOI float intersect_mesh(const mib_shape &s, const float *buf, V3 ro, V3 rd, float maxt, int &face, float &bu, float &bv) {
    const MeshView m = mesh_view(s, buf);
    const float inv[3] = { rcp_(rd.x), rcp_(rd.y), rcp_(rd.z) };
    const float org[3] = { ro.x, ro.y, ro.z };
    float best = inf_f();
    uint32_t stack[64];
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0) {
        const uint32_t ni = stack[--sp];
        const float *nd = m.nodes + 8 * (size_t) ni;
        float tmin = 0.f, tmax = fminf(best, maxt);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float t1 = (__ldg(nd + a) - org[a]) * inv[a], t2 = (__ldg(nd + 3 + a) - org[a]) * inv[a];
            tmin = fmaxf(tmin, fminf(t1, t2));
            tmax = fminf(tmax, fmaxf(t1, t2));
        }
        if (!(tmin <= tmax)) continue;
        const uint32_t a = __float_as_uint(__ldg(nd + 6)), b = __float_as_uint(__ldg(nd + 7));
        if (b & 0x80000000u) {
            if (sp + 2 > 64) continue;
            stack[sp++] = b & 0x7fffffffu;
            stack[sp++] = a;
        } else {
            for (uint32_t k = 0; k < b; ++k) {
                const uint32_t f = __ldg(m.order + a + k);
                const uint32_t *fi = m.faces + 3 * (size_t) f;
                float t, u, v;
                if (moeller_trumbore(ro, rd, mesh_vec3(m.pos, __ldg(fi)), mesh_vec3(m.pos, __ldg(fi + 1)), mesh_vec3(m.pos, __ldg(fi + 2)), maxt, t, u, v) && t < best) {
                    best = t; face = (int) f; bu = u; bv = v;
                }
            }
        }
    }
    return best;
}

// mesh.cpp:1368-1537: p, geometric normal n, shading normal sh_n, uv; returns dp_du
MIB_D V3 mesh_surface_interaction(const mib_shape &s, const float *buf, int face, float b1, float b2, V3 &p, V3 &n, V3 &sh_n, float &tu, float &tv) {
    const MeshView m = mesh_view(s, buf);
    const uint32_t *fi = m.faces + 3 * (size_t) face;
    const uint32_t i0 = __ldg(fi), i1 = __ldg(fi + 1), i2 = __ldg(fi + 2);
    const V3 p0 = mesh_vec3(m.pos, i0), p1 = mesh_vec3(m.pos, i1), p2 = mesh_vec3(m.pos, i2);
    const float b0 = 1.f - b1 - b2;
    p = vfma(p0, b0, vfma(p1, b1, p2 * b2));
    const V3 dp0 = p1 - p0, dp1 = p2 - p0;
    n = normalize(cross(dp0, dp1));
    tu = b1; tv = b2;
    V3 dp_du, dp_dv;
    coordinate_system(n, dp_du, dp_dv);
    if (m.uv) {
        const float t0x = __ldg(m.uv + 2 * i0), t0y = __ldg(m.uv + 2 * i0 + 1), t1x = __ldg(m.uv + 2 * i1), t1y = __ldg(m.uv + 2 * i1 + 1),
                    t2x = __ldg(m.uv + 2 * i2), t2y = __ldg(m.uv + 2 * i2 + 1);
        tu = fma_(t2x, b2, fma_(t1x, b1, t0x * b0));
        tv = fma_(t2y, b2, fma_(t1y, b1, t0y * b0));
        const float du0x = t1x - t0x, du0y = t1y - t0y, du1x = t2x - t0x, du1y = t2y - t0y;
        const float det = fms_(du0x, du1y, du0y * du1x), inv_det = rcp_(det);
        if (det != 0.f)
            dp_du = v3(fms_(du1y, dp0.x, du0y * dp1.x) * inv_det, fms_(du1y, dp0.y, du0y * dp1.y) * inv_det, fms_(du1y, dp0.z, du0y * dp1.z) * inv_det);
    }
    sh_n = n;
    if (m.nrm) {
        const V3 n0 = mesh_vec3(m.nrm, i0), n1 = mesh_vec3(m.nrm, i1), n2 = mesh_vec3(m.nrm, i2);
        const V3 ns = vfma(n2, b2, vfma(n1, b1, n0 * b0));
        sh_n = ns * rsqrt_(sqnorm(ns));
    }
    return dp_du;
}

} // namespace mib

// denvmap.cuh -- environment-map emitter on the device.
//   eval / sample_direction / pdf_direction / eval_spectrum   src/emitters/envmap.cpp:343-354, 404-463, 517-576
//   Hierarchical2D<Float, 0>::sample / eval                   include/mitsuba/core/distr_2d.h:492-575, 668-700
//   square_to_bilinear, interval_to_linear                    include/mitsuba/core/warp.h:447-494
//   acos, atan2 (Dr.Jit polynomials)                          ext/drjit/include/drjit/math.h:459-594
// Buffer layout: include/mi_b200.h (`mib_emitter`), built by mitsuba3_nasa_b200/envmap.py.  Same arithmetic as
// oracle/oplugins.h (env_*), which reproduces the reference's scalar build sample by sample.  Feature FT_MESH (the
// one-thread-one-path forward kernels).
#pragma once
#include "dspec.cuh"

namespace mib {

MIB_D float acos_(float x) {
    const float xa = fabsf(x), x2 = x * x;
    const bool big = xa > 0.5f;
    const float x1 = 0.5f * (1.f - xa);
    const float x3 = big ? x1 : x2;
    const float x4 = big ? sqrt_(x1) : xa;
    const float t2 = x3 * x3, t4 = t2 * t2;
    const float a0 = fma_(x3, 7.4953002686e-2f, 1.666675242e-1f);
    const float a1 = fma_(x3, 2.4181311049e-2f, 4.5470025998e-2f);
    float z1 = fma_(t4, 4.2163199048e-2f, fma_(t2, a1, a0));
    z1 = fma_(z1, x3 * x4, x4);
    float z2 = z1 + z1;
    z2 = x < 0.f ? kPi - z2 : z2;
    const float z3 = (kPi * .5f) - __int_as_float((__float_as_int(z1) & 0x7fffffff) | (__float_as_int(x) & 0x80000000));
    return big ? z2 : z3;
}
MIB_D float safe_acos_(float a) { return acos_(fmaxf(fminf(a, 1.f), -1.f)); }
MIB_D float atan2_(float y, float x) {
    const float abs_x = fabsf(x), abs_y = fabsf(y), min_val = fminf(abs_y, abs_x), max_val = fmaxf(abs_x, abs_y);
    const float scaled_min = div_(min_val, max_val), z = scaled_min * scaled_min;
    const float c0 = (float) 0.99999934166683966009, c1 = (float) -0.33326497518773606976, c2 = (float) 0.19881342388439013552,
                c3 = (float) -0.13486708938456973185, c4 = (float) 0.083863120428809689910, c5 = (float) -0.037006525670417265220,
                c6 = (float) 0.0078613793713198150252;
    const float z2 = z * z, z4 = z2 * z2;
    const float r0 = fma_(z, c1, c0), r1 = fma_(z, c3, c2), r2 = fma_(z, c5, c4);
    float t = fma_(z4, fma_(z2, c6, r2), fma_(z2, r1, r0));
    t = t * scaled_min;
    t = abs_y > abs_x ? kPi * .5f - t : t;
    t = x < 0.f ? kPi - t : t;
    const float r = y < 0.f ? -t : t;
    return max_val != 0.f ? r : 0.f;
}

MIB_D float lerp_(float a, float b, float t) { return fma_(b, t, fnma_(a, t, a)); }
MIB_D float interval_to_linear(float v0, float v1, float sample) {
    return fabsf(v0 - v1) > 1e-4f * (v0 + v1) ? div_(v0 - safe_sqrt_(lerp_(v0 * v0, v1 * v1, sample)), v0 - v1) : sample;
}
MIB_D uint32_t env_level_index(uint32_t x, uint32_t y, uint32_t width) {
    return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + ((y & ~1u) * width);
}
MIB_D float clamp01(float v) { return fminf(fmaxf(v, 0.f), 1.f); }

MIB_D void env_warp_sample(const mib_emitter &e, const float *buf, float sx, float sy, float &u, float &v, float &pdf) {
    sx = clamp01(sx); sy = clamp01(sy);
    uint32_t ox = 0, oy = 0;
    for (int l = e.env_levels - 2; l > 0; --l) {
        const float *lv = buf + e.env_level_offset[l];
        ox <<= 1; oy <<= 1;
        const uint32_t i = env_level_index(ox, oy, (uint32_t) e.env_level_width[l]);
        const float v00 = __ldg(lv + i), v10 = __ldg(lv + i + 1), v01 = __ldg(lv + i + 2), v11 = __ldg(lv + i + 3);
        sx = clamp01(sx); sy = clamp01(sy);
        const float r0 = v00 + v10, r1 = v01 + v11;
        sy *= r0 + r1;
        bool mask = sy > r0;
        if (mask) { oy += 1u; sy -= r0; }
        sy = div_(sy, mask ? r1 : r0);
        const float c0 = mask ? v01 : v00, c1 = mask ? v11 : v10;
        sx *= c0 + c1;
        mask = sx > c0;
        if (mask) sx -= c0;
        sx = div_(sx, mask ? c1 : c0);
        if (mask) ox += 1u;
    }
    const float *l0 = buf + e.env_level_offset[0];
    const uint32_t w = (uint32_t) e.env_level_width[0], i = ox + oy * w;
    const float v00 = __ldg(l0 + i), v10 = __ldg(l0 + i + 1), v01 = __ldg(l0 + i + w), v11 = __ldg(l0 + i + w + 1);
    const float r0 = v00 + v10, r1 = v01 + v11;
    sy = interval_to_linear(r0, r1, sy);
    const float c0 = lerp_(v00, v01, sy), c1 = lerp_(v10, v11, sy);
    sx = interval_to_linear(c0, c1, sx);
    pdf = lerp_(c0, c1, sx);
    const float psx = div_(1.f, (float) (e.env_res[0] - 1)), psy = div_(1.f, (float) (e.env_res[1] - 1));
    u = ((float) (int) ox + sx) * psx;
    v = ((float) (int) oy + sy) * psy;
}
MIB_D float env_warp_eval(const mib_emitter &e, const float *buf, float px, float py) {
    px = clamp01(px); py = clamp01(py);
    px *= (float) (e.env_res[0] - 1); py *= (float) (e.env_res[1] - 1);
    const uint32_t ox = min((uint32_t) (int) px, (uint32_t) (e.env_res[0] - 2)), oy = min((uint32_t) (int) py, (uint32_t) (e.env_res[1] - 2));
    px -= (float) (int) ox; py -= (float) (int) oy;
    const float *l0 = buf + e.env_level_offset[0];
    const uint32_t w = (uint32_t) e.env_level_width[0], i = ox + oy * w;
    const float v00 = __ldg(l0 + i), v10 = __ldg(l0 + i + 1), v01 = __ldg(l0 + i + w), v11 = __ldg(l0 + i + w + 1);
    return lerp_(lerp_(v00, v10, px), lerp_(v01, v11, px), py);
}
template <int NC> MIB_D void env_eval_spectrum(const mib_emitter &e, const float *buf, float u, float v, float *out) {
    const uint32_t rx = (uint32_t) e.env_res[0], ry = (uint32_t) e.env_res[1];
    u -= div_(.5f, (float) (rx - 1u));
    u -= floorf(u); v -= floorf(v);
    u *= (float) (rx - 1u); v *= (float) (ry - 1u);
    const uint32_t px = min((uint32_t) u, rx - 2u), py = min((uint32_t) v, ry - 2u);
    const float w1x = u - (float) px, w1y = v - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    const uint32_t index = py * rx + px;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        const float v00 = __ldg(buf + 3 * index + c), v10 = __ldg(buf + 3 * (index + 1) + c),
                    v01 = __ldg(buf + 3 * (index + rx) + c), v11 = __ldg(buf + 3 * (index + rx + 1) + c);
        const float v0 = fma_(w0x, v00, w1x * v10), v1 = fma_(w0x, v01, w1x * v11);
        out[c] = fma_(w0y, v0, w1y * v1) * e.env_scale;
    }
}
MIB_D void env_dir_to_uv(const mib_emitter &e, V3 d_world, float &u, float &v, V3 &local) {
    local = xform_vector(e.env_to_local, d_world);
    u = atan2_(local.x, -local.z) * kInvTwoPi;
    v = safe_acos_(local.y) * kInvPi;
}
MIB_D float env_inv_sin_theta(V3 d) { return rsqrt_(fmaxf(fmaxf(d.x * d.x + d.z * d.z, kEps * kEps), 0.f)); }

} // namespace mib

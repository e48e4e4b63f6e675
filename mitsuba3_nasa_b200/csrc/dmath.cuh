// dmath.cuh -- device arithmetic of the B200 volumetric path tracer.
//
// Parity contract: the reference's llvm_ad_* variants evaluate Dr.Jit's own
// polynomial approximations (ext/drjit/include/drjit/math.h) and fuse
// multiply-adds only where the source says dr::fmadd.  This file is compiled
// with -fmad=false; every fused operation below is an explicit __fmaf_rn, so a
// decision (real vs. null collision, hit vs. miss) is taken on the same float
// as in the reference.
#pragma once
// ---------------------------------------------------------------------------------------------
// Scene-specialised instantiations.  Every walk kernel is compiled once with all features (MIB_FEAT = all bits) and again
// for the feature sets that the BASELINE scenes use; ft(F) is a compile-time constant, so `if (ft(F) && ...)` removes the
// code of a plugin the scene does not hold.  The host picks the most specific instantiation whose set covers the scene
// (mib_api.cu scene_features / pick_variant).  Fewer live branches and a smaller hot loop is the point: the walk kernels are
// bound by instruction supply, not by arithmetic or memory.
// ---------------------------------------------------------------------------------------------
#ifndef MIB_FEAT
#define MIB_FEAT 0xfffff7ffu        // everything but FT_MESH: only the forward one-thread-one-path kernels carry meshes (Makefile)
#endif
enum : unsigned {
    FT_HOMOG = 1u, FT_GRID = 2u, FT_FAST = 4u, FT_AREA = 8u, FT_TAB = 16u, FT_ANALYTIC = 32u, FT_SPHERE = 64u, FT_CUBE = 128u,
    FT_RECT = 256u, FT_FILTER = 512u, FT_SPECEXT = 1024u, FT_MESH = 2048u /* obj meshes, uv textures */, FT_ALL = 0xffffffffu
};
#define MIB_FT(f) (((MIB_FEAT) & (f)) != 0u)
#include <cuda_runtime.h>
#include <stdint.h>

namespace mib {

#define MIB_D __device__ __forceinline__
// Real subroutines.  The walk kernels are instruction-fetch bound when everything is inlined (190-440 KB of SASS
// per kernel against a 32 KB L1.5 instruction cache; ncu: stall_no_instruction 5-30x the issue rate,
// profiles/r01_*).  Helpers that are used from many places are therefore compiled ONCE per kernel as
// __noinline__ functions with by-value arguments and by-value results: within one translation unit ptxas
// allocates registers across the call (no ABI spills, no stack traffic), so a call costs CALL + RET + a few moves.
// Never pass references / pointers to locals into a MIB_SUB function: that would force the local into memory.
// Four groups, each switchable for A/B runs (-DMIB_SUBS=mask; profiles/r01_ab_variants.md):
//   1 ARITH  div_ rcp_ sqrt_ rsqrt_ normalize coordinate_system2      2 TRANS  sincos2_ exp_ log_ cbrt_ asin_
//   4 ROT    stokes_rotation2 world_rotation1 phase_rotation_angles   8 GEOM   ray_intersect2 si_finalize2 sample_interaction2
#ifndef MIB_SUBS
#define MIB_SUBS 15
#endif
#define MIB_INL static __device__ __forceinline__
#define MIB_NOI static __device__ __noinline__
#if MIB_SUBS & 1
#define MIB_SUB_ARITH MIB_NOI
#else
#define MIB_SUB_ARITH MIB_INL
#endif
#if MIB_SUBS & 2
#define MIB_SUB_TRANS MIB_NOI
#else
#define MIB_SUB_TRANS MIB_INL
#endif
#if MIB_SUBS & 4
#define MIB_SUB_ROT MIB_NOI
#else
#define MIB_SUB_ROT MIB_INL
#endif
#if MIB_SUBS & 8
#define MIB_SUB_GEOM MIB_NOI
#else
#define MIB_SUB_GEOM MIB_INL
#endif
#define MIB_DH static __device__ __forceinline__
#define MIB_DM static __device__ __forceinline__

constexpr float kPi        = 3.14159265358979323846f;
constexpr float kInvPi     = 0.31830988618379067154f;
constexpr float kInvTwoPi  = 0.15915494309189533577f;
constexpr float kInvFourPi = 0.07957747154594766788f;
constexpr float kTwoPi     = 6.28318530717958647692f;
constexpr float kRayEps    = 1500.f * 5.9604644775390625e-08f;   // include/mitsuba/core/math.h:17-23
constexpr float kShadowEps = kRayEps * 10.f;
constexpr float kEps       = 5.9604644775390625e-08f;

MIB_D float inf_f() { return __int_as_float(0x7f800000); }
MIB_D float largest_f() { return __int_as_float(0x7f7fffff); }
MIB_D float fma_(float a, float b, float c) { return __fmaf_rn(a, b, c); }
MIB_D float fms_(float a, float b, float c) { return __fmaf_rn(a, b, -c); }
MIB_D float fnma_(float a, float b, float c) { return __fmaf_rn(-a, b, c); }
MIB_D float sqr_(float a) { return a * a; }
// IEEE division / square root: ~10 SASS instructions + a slow-path call each when inlined, and the walks contain
// well over a hundred of them -> one real subroutine each.
MIB_SUB_ARITH float div_(float a, float b) { return __fdiv_rn(a, b); }
MIB_SUB_ARITH float rcp_(float a) { return __fdiv_rn(1.f, a); }            // drjit-core op.cpp:1195 (LLVM: 1/x)
MIB_SUB_ARITH float sqrt_(float a) { return __fsqrt_rn(a); }
MIB_SUB_ARITH float rsqrt_(float a) { return __fsqrt_rn(__fdiv_rn(1.f, a)); } // op.cpp:1226 (LLVM: sqrt(1/x))
MIB_D float safe_sqrt_(float a) { return sqrt_(fmaxf(a, 0.f)); }
MIB_D float mulsign_ad(float v1, float v2) { return v2 >= 0.f ? v1 : -v1; }      // array_router.h:381-387 (diff arrays)
MIB_D float mulsign_bits(float v1, float v2) { return __int_as_float(__float_as_int(v1) ^ (__float_as_int(v2) & 0x80000000)); }

// ext/drjit/include/drjit/math.h:74-180
MIB_SUB_TRANS float2 sincos2_(float x) {          // (sin, cos)
    float s_out, c_out;
    float xa = fabsf(x);
    int j = (int) (xa * 1.2732395447351626862f);
    j = (j + 1) & ~1;
    float y = (float) j;
    unsigned sign_sin = ((unsigned) j << 29) ^ (unsigned) __float_as_int(x);
    unsigned sign_cos = (unsigned) (~(j - 2)) << 29;
    y = xa - y * 0.78515625f - y * 2.4187564849853515625e-4f - y * 3.77489497744594108e-8f;
    float z = y * y;
    if (xa == inf_f()) z = __int_as_float(0xffffffff);
    float z2 = z * z;
    // estrin(z, c0, c1, c2) = fma(z^2, c2, fma(z, c1, c0))
    float s = fma_(z2, -1.9515295891e-4f, fma_(z, 8.3321608736e-3f, -1.6666654611e-1f)) * z;
    float c = fma_(z2, 2.443315711809948e-5f, fma_(z, -1.388731625493765e-3f, 4.166664568298827e-2f)) * z;
    s = fma_(s, y, y);
    c = fma_(c, z, fma_(z, -0.5f, 1.f));
    bool polymask = (j & 2) == 0;
    s_out = mulsign_bits(polymask ? s : c, __int_as_float((int) sign_sin));
    c_out = mulsign_bits(polymask ? c : s, __int_as_float((int) sign_cos));
    return make_float2(s_out, c_out);
}
MIB_D void sincos_(float x, float &s_out, float &c_out) { float2 r = sincos2_(x); s_out = r.x; c_out = r.y; }

MIB_D float ldexp_(float x, float n) { return x * __int_as_float(((int) n + 0x7f) << 23); }  // math.h:610-627
MIB_D void frexp_(float x, float &mant, float &expo) {                                        // math.h:634-678
    unsigned xi = (unsigned) __float_as_int(x);
    unsigned eb = xi & 0x7f800000u;
    bool normal = (x != 0.f) && (eb != 0x7f800000u);
    int ei = normal ? ((int) (eb >> 23) - 0x7f) : 0;
    unsigned m = (xi & ~0x7f800000u) | 0x3f000000u;
    mant = __int_as_float((int) (normal ? m : xi));
    expo = (float) ei;
}

// math.h:685-747.  estrin with 6 coefficients: level 1 pairs, level 2 on y^2, level 3 on y^4
MIB_SUB_TRANS float exp_(float x) {
    float n = floorf(fma_(1.44269504088896340736f, x, 0.5f));
    float y = fma_(n, -0.693359375f, x);
    y = fma_(n, 2.12194440e-4f, y);
    float y2 = y * y, y4 = y2 * y2;
    float p0 = fma_(y, 1.6666665459e-1f, 5.0000001201e-1f);
    float p1 = fma_(y, 8.3334519073e-3f, 4.1665795894e-2f);
    float p2 = fma_(y, 1.9875691500e-4f, 1.3981999507e-3f);
    float q0 = fma_(y2, p1, p0);
    float z = fma_(y4, p2, q0);
    z = fma_(z, y2, y + 1.f);
    if (x > 88.3762588501f) return inf_f();
    if (x < -88.3762588501f) return 0.f;
    return ldexp_(z, n);
}

// math.h:824-897.  estrin with 9 coefficients
MIB_SUB_TRANS float log_(float x) {
    float xm, e;
    frexp_(x, xm, e);
    bool ge = xm >= 0.70710678118654752440f;
    if (ge) e += 1.f;
    xm += (ge ? 0.f : xm) - 1.f;
    float x2 = xm * xm, x4 = x2 * x2, x8 = x4 * x4;
    float a0 = fma_(xm, -2.4999993993e-1f, 3.3333331174e-1f);
    float a1 = fma_(xm, -1.6668057665e-1f, 2.0000714765e-1f);
    float a2 = fma_(xm, -1.2420140846e-1f, 1.4249322787e-1f);
    float a3 = fma_(xm, -1.1514610310e-1f, 1.1676998740e-1f);
    float a4 = 7.0376836292e-2f;
    float b0 = fma_(x2, a1, a0), b1 = fma_(x2, a3, a2), b2 = a4;
    float c0 = fma_(x4, b1, b0), c1 = b2;
    float y = fma_(x8, c1, c0);
    float z = x2;
    y *= xm * z;
    y = fma_(e, -2.121944400546905827679e-4f, y);
    float r = xm + fma_(-.5f, z, y);
    r = fma_(e, 0.693359375f, r);
    if (x == inf_f()) r = inf_f();
    if (x == 0.f) r = -inf_f();
    if (!(x >= 0.f)) r = __int_as_float(0xffffffff);
    return r;
}

// math.h:1411-1468 (one Newton step in single precision).  estrin with 5 coefficients
MIB_SUB_TRANS float cbrt_(float x) {
    const float cbrt2 = 1.25992104989487316477f, cbrt4 = 1.58740105196819947475f, third = (float) (1.0 / 3.0);
    float xa = fabsf(x), xm, xe;
    frexp_(xa, xm, xe);
    xe += 1.f;
    float xea = fabsf(xe), xea1 = floorf(xea * third), rem = fnma_(xea1, 3.f, xea);
    float m2 = xm * xm, m4 = m2 * m2;
    float a0 = fma_(xm, 1.1399983354717293273738f, 0.40238979564544752126924f);
    float a1 = fma_(xm, 0.54664601366395524503440f, -0.95438224771509446525043f);
    float a2 = -0.13466110473359520655053f;
    float b0 = fma_(m2, a1, a0);
    xm = fma_(m4, a2, b0);
    float f1 = xe >= 0.f ? cbrt2 : 1.f / cbrt2, f2 = xe >= 0.f ? cbrt4 : 1.f / cbrt4;
    float f = rem == 1.f ? f1 : f2;
    if (rem != 0.f) xm *= f;
    float r = ldexp_(xm, mulsign_bits(xea1, xe));
    r = mulsign_bits(r, x);
    r -= (r - div_(x, sqr_(r))) * third;
    return (isfinite(x) && x != 0.f) ? r : x;
}

// math.h:366-456.  estrin with 5 coefficients
MIB_SUB_TRANS float asin_(float x) {
    float xa = fabsf(x), x2 = x * x;
    bool big = xa > 0.5f;
    float x1 = 0.5f * (1.f - xa);
    float x3 = big ? x1 : x2;
    float x4 = big ? sqrt_(x1) : xa;
    float t2 = x3 * x3, t4 = t2 * t2;
    float a0 = fma_(x3, 7.4953002686e-2f, 1.6666752422e-1f);
    float a1 = fma_(x3, 2.4181311049e-2f, 4.5470025998e-2f);
    float a2 = 4.2163199048e-2f;
    float z1 = fma_(t4, a2, fma_(t2, a1, a0));
    z1 = fma_(z1, x3 * x4, x4);
    float r = big ? (.5f * kPi) - (z1 + z1) : z1;
    return __int_as_float((__float_as_int(r) & 0x7fffffff) | (__float_as_int(x) & 0x80000000));
}

// ------------------------------------------------------------------
// 3-vectors (arithmetic of Vector<DiffArray<LLVMArray<float>>, 3>)
// ------------------------------------------------------------------
struct V3 { float x, y, z; };
MIB_D V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
MIB_D V3 v3(float s) { return v3(s, s, s); }
MIB_D V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
MIB_D V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
MIB_D V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
MIB_D V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
MIB_D V3 operator*(float s, V3 a) { return v3(a.x * s, a.y * s, a.z * s); }
MIB_D V3 vfma(V3 a, float b, V3 c) { return v3(fma_(a.x, b, c.x), fma_(a.y, b, c.y), fma_(a.z, b, c.z)); }
MIB_D float dot(V3 a, V3 b) { return fma_(a.z, b.z, fma_(a.y, b.y, a.x * b.x)); }              // array_base.h:603-630
MIB_D float sqnorm(V3 v) { return fma_(v.z, v.z, fma_(v.y, v.y, v.x * v.x)); }                  // array_router.h:635-644
MIB_D float norm(V3 v) { return sqrt_(sqnorm(v)); }
MIB_SUB_ARITH V3 normalize(V3 v) { return v * rsqrt_(sqnorm(v)); }
MIB_D V3 cross(V3 a, V3 b) {                                                                     // array_router.h:654-666
    return v3(fms_(a.y, b.z, a.z * b.y), fms_(a.z, b.x, a.x * b.z), fms_(a.x, b.y, a.y * b.x));
}
MIB_D float comp(V3 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

// include/mitsuba/core/vector.h:116-136
struct V3x2 { V3 a, b; };
MIB_SUB_ARITH V3x2 coordinate_system2(V3 n) {
    float sign = n.z >= 0.f ? 1.f : -1.f, a = -rcp_(sign + n.z), b = n.x * n.y * a;
    V3x2 r;
    r.a = v3(mulsign_ad(sqr_(n.x) * a, n.z) + 1.f, mulsign_ad(b, n.z), n.z >= 0.f ? -n.x : n.x);
    r.b = v3(b, fma_(n.y, n.y * a, sign), -n.y);
    return r;
}
MIB_D void coordinate_system(V3 n, V3 &s, V3 &t) { V3x2 r = coordinate_system2(n); s = r.a; t = r.b; }
MIB_D V3 stokes_basis(V3 fwd) { V3 s, t; coordinate_system(fwd, s, t); return s; }               // mueller.h:347-350

// include/mitsuba/core/frame.h:17-41
struct Frame3 { V3 s, t, n; };
MIB_D Frame3 make_frame(V3 n) { Frame3 f; f.n = n; coordinate_system(n, f.s, f.t); return f; }
MIB_D V3 to_local(const Frame3 &f, V3 v) { return v3(dot(v, f.s), dot(v, f.t), dot(v, f.n)); }
MIB_D V3 to_world(const Frame3 &f, V3 v) { return vfma(f.n, v.z, vfma(f.t, v.y, f.s * v.x)); }

// ext/drjit/include/drjit/sphere.h:53-60
MIB_D float unit_angle(V3 a, V3 b) {
    float d = dot(a, b);
    V3 am = v3(mulsign_ad(a.x, d), mulsign_ad(a.y, d), mulsign_ad(a.z, d));
    float temp = 2.f * asin_(.5f * norm(b - am));
    return d >= 0.f ? temp : kPi - temp;
}

// transform.h:98-155, 4x4 row-major
MIB_D V3 xform_point_affine(const float *M, V3 p) {
    return v3(fma_(M[2], p.z, fma_(M[1], p.y, fma_(M[0], p.x, M[3]))),
              fma_(M[6], p.z, fma_(M[5], p.y, fma_(M[4], p.x, M[7]))),
              fma_(M[10], p.z, fma_(M[9], p.y, fma_(M[8], p.x, M[11]))));
}
MIB_D V3 xform_point(const float *M, V3 p) {
    V3 r = xform_point_affine(M, p);
    float w = fma_(M[14], p.z, fma_(M[13], p.y, fma_(M[12], p.x, M[15])));
    float iw = rcp_(w);
    return v3(r.x * iw, r.y * iw, r.z * iw);
}
MIB_D V3 xform_vector(const float *M, V3 v) {
    return v3(fma_(M[2], v.z, fma_(M[1], v.y, M[0] * v.x)),
              fma_(M[6], v.z, fma_(M[5], v.y, M[4] * v.x)),
              fma_(M[10], v.z, fma_(M[9], v.y, M[8] * v.x)));
}

// ------------------------------------------------------------------
// RNG: include/mitsuba/core/random.h:76-91, ext/drjit/include/drjit/random.h:55-120
// ------------------------------------------------------------------
MIB_D void tea32(uint32_t v0, uint32_t v1, uint32_t &o0, uint32_t &o1) {
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    o0 = v0; o1 = v1;
}
struct Pcg32 {
    uint64_t state, inc;
    MIB_D uint32_t next_u32() {
        uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        uint32_t xs = (uint32_t) (((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t) (old >> 59u);
        return (xs >> rot) | (xs << ((-(int) rot) & 31));
    }
    MIB_D float next() { return __uint_as_float((next_u32() >> 9) | 0x3f800000u) - 1.f; }
    // src/render/sampler.cpp:115-134: seed(1, v0, v1) with (v0, v1) = TEA(base_seed + seed, lane)
    MIB_D void seed_lane(uint32_t seed_value, uint32_t lane) {
        uint32_t v0, v1;
        tea32(seed_value, lane, v0, v1);
        state = 0;
        inc = ((uint64_t) v1 << 1) | 1u;
        next_u32();
        state += (uint64_t) v0;
        next_u32();
    }
};

} // namespace mib

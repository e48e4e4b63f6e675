// omath.h -- TEST INFRASTRUCTURE (parity oracle). Not part of the product.
//
// CPU restatement of the Dr.Jit arithmetic the reference's llvm_ad_* variants
// execute on the volumetric path tracing hot path.  Each function cites the
// reference file:line it follows (paths relative to the reference checkout).
// Compile with -ffp-contract=off: fused multiply-adds appear ONLY where the
// reference writes dr::fmadd / fmsub / fnmadd explicitly.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <utility>

namespace orc {

static constexpr float Pi        = 3.14159265358979323846f;
static constexpr float InvPi     = 0.31830988618379067154f;
static constexpr float InvTwoPi  = 0.15915494309189533577f;
static constexpr float InvFourPi = 0.07957747154594766788f;
static constexpr float TwoPi     = 6.28318530717958647692f;
static constexpr float Inf       = std::numeric_limits<float>::infinity();
static constexpr float Largest   = std::numeric_limits<float>::max();
// include/mitsuba/core/math.h:17-23
static constexpr float RayEpsilon    = 1500.f * 5.9604644775390625e-08f; // eps/2 * 1500, eps/2 = 2^-24
static constexpr float ShadowEpsilon = RayEpsilon * 10.f;
static constexpr float EpsilonF      = 5.9604644775390625e-08f;           // dr::Epsilon<float> = 2^-24

static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

static inline float fmadd(float a, float b, float c)  { return std::fmaf(a, b, c); }
static inline float fmsub(float a, float b, float c)  { return std::fmaf(a, b, -c); }
static inline float fnmadd(float a, float b, float c) { return std::fmaf(-a, b, c); }
static inline float sqr(float a) { return a * a; }
// drjit-core/src/op.cpp:1195-1216: LLVM backend rcp(x) = 1/x
static inline float rcp(float a) { return 1.f / a; }
// drjit-core/src/op.cpp:1226-1250: LLVM backend rsqrt(x) = sqrt(1/x)
static inline float rsqrt(float a) { return std::sqrt(1.f / a); }
static inline float maximum(float a, float b) { return std::fmax(a, b); }
static inline float minimum(float a, float b) { return std::fmin(a, b); }
// drjit/array_router.h:1987-1996
static inline float safe_sqrt(float a) { return std::sqrt(maximum(a, 0.f)); }

// Sign helpers on *differentiable* arrays (llvm_ad): drjit/array_router.h:353-397
static inline float sign_ad(float v) { return v >= 0.f ? 1.f : -1.f; }
static inline float mulsign_ad(float v1, float v2) { return v2 >= 0.f ? v1 : -v1; }
static inline float mulsign_neg_ad(float v1, float v2) { return v2 >= 0.f ? -v1 : v1; }
// Bit versions used inside the math kernels (evaluated on detached arrays)
static inline float mulsign_bits(float v1, float v2) { return u2f(f2u(v1) ^ (f2u(v2) & 0x80000000u)); }
static inline float copysign_bits(float v1, float v2) { return u2f((f2u(v1) & 0x7fffffffu) | (f2u(v2) & 0x80000000u)); }

// drjit/math.h:24-41 Estrin's scheme, exactly as the recursive template evaluates it
static inline float estrin(float x, const float *coeff, int n) {
    float buf[16];
    for (int i = 0; i < n; ++i) buf[i] = coeff[i];
    while (true) {
        int n_rec = (n - 1) / 2, n_fma = n / 2;
        float rec[16];
        for (int i = 0; i < n_fma; ++i)
            rec[i] = fmadd(x, buf[2 * i + 1], buf[2 * i]);
        if (n_rec == n_fma)
            rec[n_rec] = buf[n - 1];
        if (n_rec == 0)
            return rec[0];
        x = x * x;
        n = n_rec + 1;
        for (int i = 0; i < n; ++i) buf[i] = rec[i];
    }
}

// drjit/math.h:74-180 (single precision branch)
static inline void sincos(float x, float *s_out, float *c_out) {
    float xa = std::fabs(x);
    int32_t j = (int32_t) (xa * 1.2732395447351626862f);
    j = (j + 1) & ~1;
    float y = (float) j;
    uint32_t sign_sin = ((uint32_t) j << 29) ^ f2u(x);
    uint32_t sign_cos = (uint32_t) (~(j - 2)) << 29;
    y = xa - y * 0.78515625f - y * 2.4187564849853515625e-4f - y * 3.77489497744594108e-8f;
    float z = y * y;
    if (xa == Inf) z = u2f(0xffffffffu);
    static const float cs[3] = { -1.6666654611e-1f, 8.3321608736e-3f, -1.9515295891e-4f };
    static const float cc[3] = { 4.166664568298827e-2f, -1.388731625493765e-3f, 2.443315711809948e-5f };
    float s = estrin(z, cs, 3) * z;
    float c = estrin(z, cc, 3) * z;
    s = fmadd(s, y, y);
    c = fmadd(c, z, fmadd(z, -0.5f, 1.f));
    bool polymask = (j & 2) == 0;
    *s_out = mulsign_bits(polymask ? s : c, u2f(sign_sin));
    *c_out = mulsign_bits(polymask ? c : s, u2f(sign_cos));
}

// drjit/math.h:610-627
static inline float ldexp_(float x, float n) {
    int32_t e = (int32_t) n + 0x7f;
    return x * u2f((uint32_t) e << 23);
}

// drjit/math.h:634-678
static inline void frexp_(float x, float *mant, float *expo) {
    uint32_t xi = f2u(x);
    uint32_t exponent_bits = xi & 0x7f800000u;
    bool is_normal = (x != 0.f) && (exponent_bits != 0x7f800000u);
    int32_t exponent_i = is_normal ? ((int32_t) (exponent_bits >> 23) - 0x7f) : 0;
    uint32_t mantissa = (xi & ~0x7f800000u) | f2u(.5f);
    *mant = u2f(is_normal ? mantissa : xi);
    *expo = (float) exponent_i;
}

// drjit/math.h:685-747 (single precision)
static inline float exp_(float x) {
    const float range = 88.3762588501f;
    bool over = x > range, under = x < -range;
    float n = std::floor(fmadd(1.44269504088896340736f, x, 0.5f));
    float y = x;
    y = fmadd(n, -0.693359375f, y);
    y = fmadd(n, 2.12194440e-4f, y);
    static const float c[6] = { 5.0000001201e-1f, 1.6666665459e-1f, 4.1665795894e-2f,
                                8.3334519073e-3f, 1.3981999507e-3f, 1.9875691500e-4f };
    float z = estrin(y, c, 6);
    z = fmadd(z, sqr(y), y + 1.f);
    if (over) return Inf;
    if (under) return 0.f;
    return ldexp_(z, n);
}

// drjit/math.h:824-897 (single precision)
static inline float log_(float x) {
    bool valid = x >= 0.f;
    float xm, e;
    frexp_(x, &xm, &e);
    bool ge = xm >= 0.70710678118654752440f;
    if (ge) e += 1.f;
    xm += (ge ? 0.f : xm) - 1.f;
    static const float c[9] = { 3.3333331174e-1f, -2.4999993993e-1f, 2.0000714765e-1f,
                                -1.6668057665e-1f, 1.4249322787e-1f, -1.2420140846e-1f,
                                1.1676998740e-1f, -1.1514610310e-1f, 7.0376836292e-2f };
    float y = estrin(xm, c, 9);
    float z = sqr(xm);
    y *= xm * z;
    y = fmadd(e, -2.121944400546905827679e-4f, y);
    float r = xm + fmadd(-.5f, z, y);
    r = fmadd(e, 0.693359375f, r);
    if (x == Inf) r = Inf;
    if (x == 0.f) r = -Inf;
    if (!valid) r = u2f(0xffffffffu);
    return r;
}

// drjit/math.h:1411-1468 (single precision: one Newton iteration)
static inline float cbrt_(float x) {
    const float cbrt2 = 1.25992104989487316477f, cbrt4 = 1.58740105196819947475f,
                third = (float) (1.0 / 3.0);
    float xa = std::fabs(x);
    float xm, xe;
    frexp_(xa, &xm, &xe);
    xe += 1.f;
    float xea = std::fabs(xe), xea1 = std::floor(xea * third), rem = fnmadd(xea1, 3.f, xea);
    static const float c[5] = { 0.40238979564544752126924f, 1.1399983354717293273738f,
                                -0.95438224771509446525043f, 0.54664601366395524503440f,
                                -0.13466110473359520655053f };
    xm = estrin(xm, c, 5);
    float f1 = xe >= 0.f ? cbrt2 : 1.f / cbrt2,
          f2 = xe >= 0.f ? cbrt4 : 1.f / cbrt4,
          f  = rem == 1.f ? f1 : f2;
    if (rem != 0.f) xm *= f;
    float r = ldexp_(xm, mulsign_bits(xea1, xe));
    r = mulsign_bits(r, x);
    r -= (r - (x / sqr(r))) * third;
    return (std::isfinite(x) && x != 0.f) ? r : x;
}

// drjit/math.h:366-456 (single precision)
static inline float asin_(float x) {
    float xa = std::fabs(x), x2 = sqr(x), r;
    bool big = xa > 0.5f;
    float x1 = 0.5f * (1.f - xa);
    float x3 = big ? x1 : x2;
    float x4 = big ? std::sqrt(x1) : xa;
    static const float c[5] = { 1.6666752422e-1f, 7.4953002686e-2f, 4.5470025998e-2f,
                                2.4181311049e-2f, 4.2163199048e-2f };
    float z1 = estrin(x3, c, 5);
    z1 = fmadd(z1, x3 * x4, x4);
    r = big ? (.5f * Pi) - (z1 + z1) : z1;
    return copysign_bits(r, x);
}

// ext/drjit/include/drjit/math.h:459-506 (acos, single precision), array_router.h:2032-2042 (safe_acos)
static inline float acos_(float x) {
    float xa = std::fabs(x), x2 = sqr(x);
    bool big = xa > 0.5f;
    float x1 = 0.5f * (1.f - xa);
    float x3 = big ? x1 : x2;
    float x4 = big ? std::sqrt(x1) : xa;
    static const float c[5] = { 1.666675242e-1f, 7.4953002686e-2f, 4.5470025998e-2f, 2.4181311049e-2f, 4.2163199048e-2f };
    float z1 = estrin(x3, c, 5);
    z1 = fmadd(z1, x3 * x4, x4);
    float z2 = z1 + z1;
    z2 = x < 0.f ? Pi - z2 : z2;
    float z3 = (Pi * .5f) - copysign_bits(z1, x);
    return big ? z2 : z3;
}
static inline float safe_acos_(float a) { return acos_(maximum(minimum(a, 1.f), -1.f)); }
// math.h:517-594 (atan2, single precision)
static inline float atan2_(float y, float x) {
    float abs_x = std::fabs(x), abs_y = std::fabs(y), min_val = minimum(abs_y, abs_x), max_val = maximum(abs_x, abs_y),
          scaled_min = min_val / max_val, z = sqr(scaled_min);
    static const float c[7] = { (float) 0.99999934166683966009, (float) -0.33326497518773606976, (float) 0.19881342388439013552,
                                (float) -0.13486708938456973185, (float) 0.083863120428809689910, (float) -0.037006525670417265220,
                                (float) 0.0078613793713198150252 };
    float t = estrin(z, c, 7);
    t = t * scaled_min;
    t = abs_y > abs_x ? Pi * .5f - t : t;
    t = x < 0.f ? Pi - t : t;
    float r = y < 0.f ? -t : t;
    return max_val != 0.f ? r : 0.f;
}

// ------------------------------------------------------------------
// 3-vectors with the arithmetic of Vector<DiffArray<LLVMArray<float>>,3>
// ------------------------------------------------------------------
struct Vec3 {
    float x, y, z;
    Vec3() : x(0), y(0), z(0) {}
    Vec3(float v) : x(v), y(v), z(v) {}
    Vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    float &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
};
static inline Vec3 operator+(Vec3 a, Vec3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
static inline Vec3 operator-(Vec3 a, Vec3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
static inline Vec3 operator-(Vec3 a) { return { -a.x, -a.y, -a.z }; }
static inline Vec3 operator*(Vec3 a, float s) { return { a.x * s, a.y * s, a.z * s }; }
static inline Vec3 operator*(float s, Vec3 a) { return { a.x * s, a.y * s, a.z * s }; }
static inline Vec3 operator*(Vec3 a, Vec3 b) { return { a.x * b.x, a.y * b.y, a.z * b.z }; }
// drjit/array_router.h:219-233: vector / scalar = vector * rcp(scalar)
static inline Vec3 operator/(Vec3 a, float s) { float r = rcp(s); return { a.x * r, a.y * r, a.z * r }; }
static inline Vec3 fmadd(Vec3 a, float b, Vec3 c) { return { fmadd(a.x, b, c.x), fmadd(a.y, b, c.y), fmadd(a.z, b, c.z) }; }
static inline Vec3 fmadd(Vec3 a, Vec3 b, Vec3 c) { return { fmadd(a.x, b.x, c.x), fmadd(a.y, b.y, c.y), fmadd(a.z, b.z, c.z) }; }
// drjit/array_base.h:603-630
static inline float dot(Vec3 a, Vec3 b) { return fmadd(a.z, b.z, fmadd(a.y, b.y, a.x * b.x)); }
// drjit/array_router.h:635-652
static inline float squared_norm(Vec3 v) { return fmadd(v.z, v.z, fmadd(v.y, v.y, sqr(v.x))); }
static inline float norm(Vec3 v) { return std::sqrt(squared_norm(v)); }
static inline Vec3 normalize(Vec3 v) { return v * rsqrt(squared_norm(v)); }
// drjit/array_router.h:654-666 (x86 branch)
static inline Vec3 cross(Vec3 a, Vec3 b) {
    return { fmsub(a.y, b.z, a.z * b.y), fmsub(a.z, b.x, a.x * b.z), fmsub(a.x, b.y, a.y * b.x) };
}
static inline float max3(Vec3 v) { return maximum(maximum(v.x, v.y), v.z); }
static inline float min3(Vec3 v) { return minimum(minimum(v.x, v.y), v.z); }
static inline Vec3 abs3(Vec3 v) { return { std::fabs(v.x), std::fabs(v.y), std::fabs(v.z) }; }

// include/mitsuba/core/vector.h:116-136
static inline void coordinate_system(Vec3 n, Vec3 *s, Vec3 *t) {
    float sign = sign_ad(n.z), a = -rcp(sign + n.z), b = n.x * n.y * a;
    *s = Vec3(mulsign_ad(sqr(n.x) * a, n.z) + 1.f, mulsign_ad(b, n.z), mulsign_neg_ad(n.x, n.z));
    *t = Vec3(b, fmadd(n.y, n.y * a, sign), -n.y);
}

// include/mitsuba/core/frame.h:17-41
struct Frame {
    Vec3 s, t, n;
    Frame() {}
    explicit Frame(Vec3 v) : n(v) { coordinate_system(v, &s, &t); }
    Frame(Vec3 s_, Vec3 t_, Vec3 n_) : s(s_), t(t_), n(n_) {}
    Vec3 to_local(Vec3 v) const { return { dot(v, s), dot(v, t), dot(v, n) }; }
    Vec3 to_world(Vec3 v) const { return fmadd(n, v.z, fmadd(t, v.y, s * v.x)); }
};

// drjit/sphere.h:53-60
static inline float unit_angle(Vec3 a, Vec3 b) {
    float dot_uv = dot(a, b);
    Vec3 am(mulsign_ad(a.x, dot_uv), mulsign_ad(a.y, dot_uv), mulsign_ad(a.z, dot_uv));
    float temp = 2.f * asin_(.5f * norm(b - am));
    return dot_uv >= 0.f ? temp : Pi - temp;
}

// ------------------------------------------------------------------
// 4x4 row-major affine/projective transforms (include/mitsuba/core/transform.h:98-155)
// ------------------------------------------------------------------
struct Mat4 { float m[4][4]; };
static inline Mat4 mat4_from(const float *p) { Mat4 r; std::memcpy(r.m, p, 64); return r; }
static inline Vec3 transform_point_affine(const Mat4 &M, Vec3 p) {
    Vec3 r;
    for (int i = 0; i < 3; ++i)
        r[i] = fmadd(M.m[i][2], p.z, fmadd(M.m[i][1], p.y, fmadd(M.m[i][0], p.x, M.m[i][3])));
    return r;
}
static inline Vec3 transform_point(const Mat4 &M, Vec3 p) {
    float r[4];
    for (int i = 0; i < 4; ++i)
        r[i] = fmadd(M.m[i][2], p.z, fmadd(M.m[i][1], p.y, fmadd(M.m[i][0], p.x, M.m[i][3])));
    float iw = rcp(r[3]);
    return { r[0] * iw, r[1] * iw, r[2] * iw };
}
static inline Vec3 transform_vector(const Mat4 &M, Vec3 v) {
    Vec3 r;
    for (int i = 0; i < 3; ++i)
        r[i] = fmadd(M.m[i][2], v.z, fmadd(M.m[i][1], v.y, M.m[i][0] * v.x));
    return r;
}

} // namespace orc

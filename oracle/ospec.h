// ospec.h -- TEST INFRASTRUCTURE (parity oracle). Not part of the product.
// RNG (PCG32 + TEA), Mueller algebra and the three Spectrum kinds of the
// reference (Color<Float,1>, Color<Float,3>, MuellerMatrix<Color<Float,1>>).
#pragma once
#include "omath.h"

namespace orc {

// include/mitsuba/core/random.h:76-91
static inline void sample_tea_32(uint32_t v0, uint32_t v1, uint32_t *o0, uint32_t *o1, int rounds = 4) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    *o0 = v0; *o1 = v1;
}

// ext/drjit/include/drjit/random.h:25-27,55-120
struct PCG32 {
    uint64_t state, inc;
    void seed(uint64_t initstate = 0x853c49e6748fea9bULL, uint64_t initseq = 0xda3e39cb94b95bdbULL) {
        state = 0;
        inc = (initseq << 1) | 1u;
        next_uint32();
        state += initstate;
        next_uint32();
    }
    uint32_t next_uint32() {
        uint64_t oldstate = state;
        state = oldstate * 0x5851f42d4c957f2dULL + inc;
        uint32_t xorshifted = (uint32_t) (((oldstate >> 18u) ^ oldstate) >> 27u);
        uint32_t rot = (uint32_t) (oldstate >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((-(int32_t) rot) & 31));
    }
    float next_float32() { return u2f((next_uint32() >> 9) | 0x3f800000u) - 1.f; }
};

// src/render/sampler.cpp:115-134 (JIT branch) + src/samplers/independent.cpp:88-97
struct Sampler {
    PCG32 rng;
    // llvm_ad seeding: lane index scrambled with TEA
    void seed_llvm(uint32_t seed_value, uint32_t lane) {
        uint32_t v0, v1;
        sample_tea_32(seed_value, lane, &v0, &v1);
        rng.seed((uint64_t) v0, (uint64_t) v1);
    }
    // scalar_* seeding (sampler.cpp:131-133): PCG32(initstate = seed_value, initseq = PCG32_DEFAULT_STREAM); one stream
    // per pixel, continued over the pixel's samples (integrator.cpp:1110-1123: sampler->seed(seed + i), then spp samples)
    void seed_scalar(uint32_t seed_value) { rng.seed((uint64_t) seed_value, 0xda3e39cb94b95bdbULL); }
    float next_1d() { return rng.next_float32(); }
    void next_2d(float *a, float *b) { *a = next_1d(); *b = next_1d(); }
};

// ------------------------------------------------------------------
// Spectrum kinds
// ------------------------------------------------------------------
template <int N> struct Color {
    float v[N];
    Color() { for (int i = 0; i < N; ++i) v[i] = 0.f; }
    Color(float s) { for (int i = 0; i < N; ++i) v[i] = s; }
    float &operator[](int i) { return v[i]; }
    float operator[](int i) const { return v[i]; }
};
#define ORC_COLOR_OP(op) \
    template <int N> static inline Color<N> operator op(Color<N> a, Color<N> b) { Color<N> r; for (int i = 0; i < N; ++i) r.v[i] = a.v[i] op b.v[i]; return r; } \
    template <int N> static inline Color<N> operator op(Color<N> a, float b) { Color<N> r; for (int i = 0; i < N; ++i) r.v[i] = a.v[i] op b; return r; } \
    template <int N> static inline Color<N> operator op(float a, Color<N> b) { Color<N> r; for (int i = 0; i < N; ++i) r.v[i] = a op b.v[i]; return r; }
ORC_COLOR_OP(+)
ORC_COLOR_OP(-)
ORC_COLOR_OP(*)
#undef ORC_COLOR_OP
// Color / Color is element-wise division; Color / Float is Color * rcp(Float) (array_router.h:219-233)
template <int N> static inline Color<N> operator/(Color<N> a, Color<N> b) { Color<N> r; for (int i = 0; i < N; ++i) r.v[i] = a.v[i] / b.v[i]; return r; }
template <int N> static inline Color<N> div_scalar(Color<N> a, float b) { float q = rcp(b); Color<N> r; for (int i = 0; i < N; ++i) r.v[i] = a.v[i] * q; return r; }
template <int N> static inline Color<N> operator-(Color<N> a) { Color<N> r; for (int i = 0; i < N; ++i) r.v[i] = -a.v[i]; return r; }
template <int N> static inline bool any_neq0(Color<N> a) { for (int i = 0; i < N; ++i) if (a.v[i] != 0.f) return true; return false; }
template <int N> static inline float cmax(Color<N> a) { float m = a.v[0]; for (int i = 1; i < N; ++i) m = maximum(m, a.v[i]); return m; }
template <int N> static inline Color<N> cexp(Color<N> a) { Color<N> r; for (int i = 0; i < N; ++i) r.v[i] = exp_(a.v[i]); return r; }

// MuellerMatrix<Color<Float,1>>: row-major m[r][c]
struct Mueller {
    float m[4][4];
    Mueller() { for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) m[i][j] = 0.f; }
    // drjit Matrix(scalar) = scalar * identity
    Mueller(float s) { for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) m[i][j] = (i == j) ? s : 0.f; }
};
// ext/drjit/include/drjit/matrix.h:102-117: column j = A.col(0)*B(0,j), then fmadd over i=1..3
static inline Mueller matmul(const Mueller &A, const Mueller &B) {
    Mueller R;
    for (int j = 0; j < 4; ++j)
        for (int r = 0; r < 4; ++r) {
            float acc = A.m[r][0] * B.m[0][j];
            for (int i = 1; i < 4; ++i)
                acc = fmadd(A.m[r][i], B.m[i][j], acc);
            R.m[r][j] = acc;
        }
    return R;
}
static inline Mueller transpose(const Mueller &A) {
    Mueller R;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) R.m[i][j] = A.m[j][i];
    return R;
}
static inline Mueller operator*(const Mueller &A, float s) { Mueller R; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) R.m[i][j] = A.m[i][j] * s; return R; }
static inline Mueller operator*(float s, const Mueller &A) { Mueller R; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) R.m[i][j] = s * A.m[i][j]; return R; }
static inline Mueller operator+(const Mueller &A, const Mueller &B) { Mueller R; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) R.m[i][j] = A.m[i][j] + B.m[i][j]; return R; }
static inline Mueller operator-(const Mueller &A, const Mueller &B) { Mueller R; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) R.m[i][j] = A.m[i][j] - B.m[i][j]; return R; }

// include/mitsuba/core/spectrum.h:71-96
static inline Mueller depolarizer(float s) { Mueller R; R.m[0][0] = s; return R; }

// include/mitsuba/render/mueller.h:164-173
static inline Mueller rotator(float theta) {
    float s, c;
    sincos(2.f * theta, &s, &c);
    Mueller R;
    R.m[0][0] = 1.f; R.m[1][1] = c; R.m[1][2] = s; R.m[2][1] = -s; R.m[2][2] = c; R.m[3][3] = 1.f;
    return R;
}
// mueller.h:347-350
static inline Vec3 stokes_basis(Vec3 forward) { Vec3 s, t; coordinate_system(forward, &s, &t); return s; }
// mueller.h:375-386
static inline Mueller rotate_stokes_basis(Vec3 forward, Vec3 basis_current, Vec3 basis_target) {
    float theta = unit_angle(normalize(basis_current), normalize(basis_target));
    if (dot(forward, cross(basis_current, basis_target)) < 0.f)
        theta *= -1.f;
    return rotator(theta);
}
// mueller.h:421-434
static inline Mueller rotate_mueller_basis(const Mueller &M, Vec3 in_forward, Vec3 in_basis_current, Vec3 in_basis_target,
                                           Vec3 out_forward, Vec3 out_basis_current, Vec3 out_basis_target) {
    Mueller R_in  = rotate_stokes_basis(in_forward, in_basis_current, in_basis_target);
    Mueller R_out = rotate_stokes_basis(out_forward, out_basis_current, out_basis_target);
    return matmul(matmul(R_out, M), transpose(R_in));
}
// mueller.h:311-328
static inline Mueller rayleigh_scatter(float cos_theta, float sin_theta, float rho) {
    float r1 = (1.f - rho) / (1.f + rho / 2.f), r2 = (1.f + rho) / (1.f - rho), r3 = (1.f - 2.f * rho) / (1.f - rho);
    float a = r2 + sqr(cos_theta), b = 1.f + sqr(cos_theta), c = -sqr(sin_theta), d = 2.f * cos_theta;
    Mueller M;
    M.m[0][0] = a; M.m[0][1] = c; M.m[1][0] = c; M.m[1][1] = b; M.m[2][2] = d; M.m[3][3] = d * r3;
    // Float(3/16) * InvPi * r1 * Matrix: scalars multiply left-to-right, then scale the matrix
    float k = (3.f / 16.f) * InvPi * r1;
    return k * M;
}

// to_world_mueller for a frame (interaction.h:374-395 and :574-595)
static inline Mueller to_world_mueller(const Frame &fr, const Mueller &M_local, Vec3 in_forward_local, Vec3 out_forward_local) {
    Vec3 in_forward_world = fr.to_world(in_forward_local), out_forward_world = fr.to_world(out_forward_local);
    Vec3 in_basis_current = fr.to_world(stokes_basis(in_forward_local)), in_basis_target = stokes_basis(in_forward_world);
    Vec3 out_basis_current = fr.to_world(stokes_basis(out_forward_local)), out_basis_target = stokes_basis(out_forward_world);
    return rotate_mueller_basis(M_local, in_forward_world, in_basis_current, in_basis_target,
                                out_forward_world, out_basis_current, out_basis_target);
}

// ------------------------------------------------------------------
// Spectrum traits: SPEC = 0 mono, 1 rgb, 2 mono_polarized
// ------------------------------------------------------------------
template <int SPEC> struct SpecTraits;
template <> struct SpecTraits<0> { static constexpr int NC = 1; static constexpr bool POL = false; using USpec = Color<1>; using Spec = Color<1>; };
template <> struct SpecTraits<1> { static constexpr int NC = 3; static constexpr bool POL = false; using USpec = Color<3>; using Spec = Color<3>; };
template <> struct SpecTraits<2> { static constexpr int NC = 1; static constexpr bool POL = true;  using USpec = Color<1>; using Spec = Mueller; };

// Spectrum * Spectrum (Color: element-wise; Mueller: matrix product)
template <int N> static inline Color<N> smul(Color<N> a, Color<N> b) { return a * b; }
static inline Mueller smul(const Mueller &a, const Mueller &b) { return matmul(a, b); }
// Spectrum * UnpolarizedSpectrum
template <int N> static inline Color<N> smul_u(Color<N> a, Color<N> b) { return a * b; }
static inline Mueller smul_u(const Mueller &a, Color<1> b) { return a * b.v[0]; }
// unpolarized_spectrum()
template <int N> static inline Color<N> unpol(Color<N> a) { return a; }
static inline Color<1> unpol(const Mueller &a) { return Color<1>(a.m[0][0]); }
// depolarizer<Spectrum>(UnpolarizedSpectrum)
template <typename Spec, int N> struct Depol;
template <int N> struct Depol<Color<N>, N> { static Color<N> make(Color<N> u) { return u; } };
template <> struct Depol<Mueller, 1> { static Mueller make(Color<1> u) { return depolarizer(u.v[0]); } };
// Spectrum from a scalar (Color: broadcast; Mueller: scalar * identity)
template <typename Spec> static inline Spec spec_from_scalar(float s) { return Spec(s); }
// masked NaN -> 0 (rayleigh.cpp:87-88)
static inline void zero_nans(Mueller &M) { for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) if (M.m[i][j] != M.m[i][j]) M.m[i][j] = 0.f; }

} // namespace orc

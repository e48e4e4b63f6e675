// dspec.cuh -- spectrum kinds in registers: mono (1 float), rgb (3 floats),
// mono_polarized (4x4 Mueller throughput, Stokes 4-vector result).
//
// Exactness notes.  The reference multiplies dense 4x4 matrices with the fma
// chain of ext/drjit/include/drjit/matrix.h:102-117:
//     C(r,j) = fma(A(r,3),B(3,j), fma(A(r,2),B(2,j), fma(A(r,1),B(1,j), A(r,0)*B(0,j)))).
// Rotators (mueller.h:164-173) are [1 0 0 0; 0 c s 0; 0 -s c 0; 0 0 0 1] and
// emitters / diffuse BSDFs are depolarizers (only entry (0,0) set), so for finite
// inputs fma(a, 0, x) == x and 1*x == x hold exactly and the sparse forms below
// return the same floats as the dense chain (up to the sign of zeros).  Radiance
// only ever has its first column populated (emitters are depolarizers), hence
// the result is kept as a Stokes 4-vector.
#pragma once
#include "dscene.cuh"

namespace mib {

template <int SPEC> struct ST {
    static constexpr bool POL = SPEC == MIB_SPEC_MONO_POLARIZED;
    static constexpr int NC = SPEC == MIB_SPEC_RGB ? 3 : 1;
    static constexpr int NT = POL ? 16 : NC;   // throughput registers
    static constexpr int NR = POL ? 4 : NC;    // result registers
};

// dense product, row-major (matrix.h:102-117)
MIB_D void m4_mul(const float *A, const float *B, float *C) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float acc = A[4 * r] * B[j];
            acc = fma_(A[4 * r + 1], B[4 + j], acc);
            acc = fma_(A[4 * r + 2], B[8 + j], acc);
            acc = fma_(A[4 * r + 3], B[12 + j], acc);
            C[4 * r + j] = acc;
        }
}
// M <- rotator(co,so) * M   (rows 1,2 mix)
MIB_D void m4_rot_left(float *M, float c, float s) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float m1 = M[4 + j], m2 = M[8 + j];
        M[4 + j] = fma_(s, m2, c * m1);
        M[8 + j] = fma_(c, m2, (-s) * m1);
    }
}
// M <- M * transpose(rotator(ci,si))   (cols 1,2 mix)
MIB_D void m4_rot_right_t(float *M, float c, float s) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        float m1 = M[4 * r + 1], m2 = M[4 * r + 2];
        M[4 * r + 1] = fma_(m2, s, m1 * c);
        M[4 * r + 2] = fma_(m2, c, m1 * (-s));
    }
}
// T <- T * B where B = [1 0 0 0; 0 b11 b12 0; 0 b21 b22 0; 0 0 0 1]
MIB_D void m4_mul_right_block(float *T, float b11, float b12, float b21, float b22) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        float t1 = T[4 * r + 1], t2 = T[4 * r + 2];
        T[4 * r + 1] = fma_(t2, b21, t1 * b11);
        T[4 * r + 2] = fma_(t2, b22, t1 * b12);
    }
}
MIB_D void m4_identity(float *M, float s) {
#pragma unroll
    for (int i = 0; i < 16; ++i) M[i] = (i % 5 == 0) ? s : 0.f;
}

// mueller.h:375-386: angle of rotate_stokes_basis -> (cos 2theta, sin 2theta)
MIB_SUB_ROT float2 stokes_rotation2(V3 forward, V3 basis_current, V3 basis_target) {       // (cos, sin)
    float theta = unit_angle(normalize(basis_current), normalize(basis_target));
    if (dot(forward, cross(basis_current, basis_target)) < 0.f) theta *= -1.f;
    float s, c;
    sincos_(2.f * theta, s, c);
    return make_float2(c, s);
}
MIB_D void stokes_rotation(V3 forward, V3 basis_current, V3 basis_target, float &c, float &s) {
    float2 r = stokes_rotation2(forward, basis_current, basis_target);
    c = r.x; s = r.y;
}

// to_world_mueller (interaction.h:374-395, :574-595): the two rotations for local directions in/out
MIB_SUB_ROT float2 world_rotation1(Frame3 fr, V3 l) {
    V3 w = to_world(fr, l);
    V3 cur = to_world(fr, stokes_basis(l)), tgt = stokes_basis(w);
    return stokes_rotation2(w, cur, tgt);
}
MIB_D void world_rotations(const Frame3 &fr, V3 in_l, V3 out_l, float &ci, float &si, float &co, float &so) {
    float2 ri = world_rotation1(fr, in_l), ro = world_rotation1(fr, out_l);
    ci = ri.x; si = ri.y; co = ro.x; so = ro.y;
}
// same with the rotation of the outgoing side given (cos, sin)
MIB_D void m4_to_world_with(const Frame3 &fr, float *M, V3 in_l, float2 out_rot) {
    const float2 ri = world_rotation1(fr, in_l);
    m4_rot_left(M, out_rot.x, out_rot.y);
    m4_rot_right_t(M, ri.x, ri.y);
}
MIB_D void m4_to_world(const Frame3 &fr, float *M, V3 in_l, V3 out_l) {
    float ci, si, co, so;
    world_rotations(fr, in_l, out_l, ci, si, co, so);
    m4_rot_left(M, co, so);
    m4_rot_right_t(M, ci, si);
}

// Phase-matrix rotation from the scattering plane (rayleigh.cpp:68-88, tabphase_polarized.cpp:169-197)
MIB_SUB_ROT float4 phase_rotation_angles(V3 wo, V3 wi) {           // (ci, si, co, so)
    V3 mwo = -wo;
    V3 x_hat = cross(mwo, wi);
    V3 p_in = normalize(cross(x_hat, mwo)), p_out = normalize(cross(x_hat, wi));
    float2 a = stokes_rotation2(mwo, p_in, stokes_basis(mwo)), b = stokes_rotation2(wi, p_out, stokes_basis(wi));
    return make_float4(a.x, a.y, b.x, b.y);
}
MIB_D void m4_phase_rotate(float *M, V3 wo, V3 wi) {
    float4 r = phase_rotation_angles(wo, wi);
    m4_rot_left(M, r.z, r.w);
    m4_rot_right_t(M, r.x, r.y);
#pragma unroll
    for (int i = 0; i < 16; ++i) if (M[i] != M[i]) M[i] = 0.f;
}

// ---- throughput helpers -------------------------------------------------
template <int SPEC> MIB_D void T_init(float *T) {
    if constexpr (ST<SPEC>::POL) m4_identity(T, 1.f);
    else { for (int c = 0; c < ST<SPEC>::NC; ++c) T[c] = 1.f; }
}
template <int SPEC> MIB_D void T_mul_f(float *T, float f) {
#pragma unroll
    for (int i = 0; i < ST<SPEC>::NT; ++i) T[i] *= f;
}
template <int SPEC> MIB_D void T_mul_u(float *T, const float *u) {
    if constexpr (ST<SPEC>::POL) { for (int i = 0; i < 16; ++i) T[i] *= u[0]; }
    else { for (int c = 0; c < ST<SPEC>::NC; ++c) T[c] *= u[c]; }
}
template <int SPEC> MIB_D bool T_any_nonzero(const float *T) {
    if constexpr (ST<SPEC>::POL) return T[0] != 0.f;
    else { bool r = false; for (int c = 0; c < ST<SPEC>::NC; ++c) r = r || (T[c] != 0.f); return r; }
}
template <int SPEC> MIB_D float T_max(const float *T) {
    if constexpr (ST<SPEC>::POL) return T[0];
    else { float m = T[0]; for (int c = 1; c < ST<SPEC>::NC; ++c) m = fmaxf(m, T[c]); return m; }
}
// T <- T * W (W same layout as T: Mueller or per-channel)
template <int SPEC> MIB_D void T_mul_W(float *T, const float *W) {
    if constexpr (ST<SPEC>::POL) {
        float C[16];
        m4_mul(T, W, C);
#pragma unroll
        for (int i = 0; i < 16; ++i) T[i] = C[i];
    } else { for (int c = 0; c < ST<SPEC>::NC; ++c) T[c] *= W[c]; }
}
// T <- T * depolarizer(v)  /  T * v
template <int SPEC> MIB_D void T_mul_depol(float *T, const float *v) {
    if constexpr (ST<SPEC>::POL) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            float t0 = T[4 * r];
            T[4 * r] = t0 * v[0];
            T[4 * r + 1] = t0 * 0.f; T[4 * r + 2] = t0 * 0.f; T[4 * r + 3] = t0 * 0.f;
        }
    } else { for (int c = 0; c < ST<SPEC>::NC; ++c) T[c] *= v[c]; }
}
// first column of (T * W):  x[r] = sum_i T(r,i) W(i,0) in the reference's fma order
template <int SPEC> MIB_D void TW_col0(const float *T, const float *W, float *x) {
    if constexpr (ST<SPEC>::POL) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            float acc = T[4 * r] * W[0];
            acc = fma_(T[4 * r + 1], W[4], acc);
            acc = fma_(T[4 * r + 2], W[8], acc);
            acc = fma_(T[4 * r + 3], W[12], acc);
            x[r] = acc;
        }
    } else { for (int c = 0; c < ST<SPEC>::NC; ++c) x[c] = T[c] * W[c]; }
}
// first column of T itself
template <int SPEC> MIB_D void T_col0(const float *T, float *x) {
    if constexpr (ST<SPEC>::POL) { for (int r = 0; r < 4; ++r) x[r] = T[4 * r]; }
    else { for (int c = 0; c < ST<SPEC>::NC; ++c) x[c] = T[c]; }
}
template <int SPEC> MIB_D float emis(const float *e, int r) { return ST<SPEC>::POL ? e[0] : e[r]; }

} // namespace mib

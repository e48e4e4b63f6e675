// omie.h -- TEST INFRASTRUCTURE (parity oracle). Not part of the product.
// Lorenz-Mie phase-matrix tabulation restated from
//   mie()                       include/mitsuba/render/mie.h:45-222   (float, Dr.Jit complex arithmetic)
//   mueller::mie_scatter        include/mitsuba/render/mueller.h:282-295
//   MiePhaseFunction::eval_mie  src/phase/mie.cpp:117-215             (size-distribution average, JIT branch)
// Complex arithmetic follows ext/drjit/include/drjit/complex.h:81-135,172-182 and sincosh math.h:1112-1158.
// Parity status: no reference test pins mie -> validated against an independent float64
// implementation in tests/test_mie.py.
#pragma once
#include "omath.h"
#include <vector>

namespace orc {

struct Cx { float re, im; };
static inline Cx cx(float r, float i = 0.f) { return { r, i }; }
static inline Cx operator+(Cx a, Cx b) { return { a.re + b.re, a.im + b.im }; }
static inline Cx operator-(Cx a, Cx b) { return { a.re - b.re, a.im - b.im }; }
static inline Cx operator-(Cx a) { return { -a.re, -a.im }; }
static inline Cx operator*(Cx a, Cx b) { return { fmsub(a.re, b.re, a.im * b.im), fmadd(a.re, b.im, a.im * b.re) }; }
static inline Cx operator*(Cx a, float s) { return { a.re * s, a.im * s }; }
static inline Cx operator*(float s, Cx a) { return { s * a.re, s * a.im }; }
static inline float cx_sqnorm(Cx z) { return fmadd(z.im, z.im, z.re * z.re); }
static inline Cx cx_rcp(Cx z) { float r = rcp(cx_sqnorm(z)); return { z.re * r, -z.im * r }; }
static inline Cx operator/(Cx a, Cx b) { return a * cx_rcp(b); }

static inline void sincosh_(float x, float *sh, float *ch) {
    float xa = std::fabs(x), exp0 = exp_(x), exp1 = rcp(exp0);
    float r_big = (exp0 - exp1) * 0.5f;
    float x2 = sqr(x);
    const float c[3] = { 1.66667160211e-1f, 8.33028376239e-3f, 2.03721912945e-4f };
    float y = estrin(x2, c, 3);
    float r_small = fmadd(y, x2 * x, x);
    *sh = xa > 1.f ? r_big : r_small;
    *ch = (exp0 + exp1) * .5f;
}
static inline Cx cx_sin(Cx z) { float s, c, sh, ch; sincos(z.re, &s, &c); sincosh_(z.im, &sh, &ch); return { s * ch, c * sh }; }
static inline Cx cx_exp(Cx z) { float e = exp_(z.re), s, c; sincos(z.im, &s, &c); return { e * c, e * s }; }

struct MieSetup {
    Cx m_sq, x, y, rcp_x, rcp_y, kx;
    float x_norm, y_norm;
};
static inline MieSetup mie_setup(float wavelength, float radius, Cx ior_med, Cx ior_sph) {
    MieSetup s;
    Cx m = ior_sph * cx_rcp(ior_med);
    float rw = rcp(wavelength);
    s.kx = (TwoPi * ior_med) * rw;
    Cx ky = (TwoPi * ior_sph) * rw;
    s.x = s.kx * radius; s.y = ky * radius;
    s.m_sq = m * m;
    s.rcp_x = cx_rcp(s.x); s.rcp_y = cx_rcp(s.y);
    s.x_norm = std::sqrt(cx_sqnorm(s.x)); s.y_norm = std::sqrt(cx_sqnorm(s.y));
    return s;
}
static inline int mie_nmax(float norm) { return (int) (8.f + norm + 4.05f * cbrt_(norm)); }
static inline int mie_ndown(int nmax) { return nmax + 8 * (int) std::sqrt((double) nmax) + 3; }

struct MieResult { Cx s1, s2; float ns, cs, ct; };

// One (mu, radius) lane of mie(); x_nmax / y_nmax are the maxima over all lanes (max_nested, mie.h:76-79).
static inline MieResult mie_lane(float wavelength, float mu, float radius, Cx ior_med, Cx ior_sph, int x_nmax, int y_nmax) {
    MieSetup s = mie_setup(wavelength, radius, ior_med, ior_sph);
    const Cx i = cx(0.f, 1.f);
    int x_ndown = mie_ndown(x_nmax), y_ndown = mie_ndown(y_nmax);
    std::vector<Cx> jrx(x_ndown), jry(y_ndown);
    Cx j = s.x * rcp(2.f * (float) x_ndown + 1.f);
    for (int n = x_ndown - 1; n >= 1; --n) { Cx k = (float) (2 * n + 1) * s.rcp_x; j = cx_rcp(k - j); jrx[n] = j; }
    j = s.y * rcp(2.f * (float) y_ndown + 1.f);
    for (int n = y_ndown - 1; n >= 1; --n) { Cx k = (float) (2 * n + 1) * s.rcp_y; j = cx_rcp(k - j); jry[n] = j; }

    Cx jx_0 = cx_sin(s.x) * s.rcp_x, jy_0 = cx_sin(s.y) * s.rcp_y;
    Cx h_exp = cx_exp(i * s.x) * s.rcp_x, hx_0 = (-i) * h_exp;
    Cx one_p = i * s.rcp_x; one_p.re = 1.f + one_p.re;
    Cx hx_1 = (-h_exp) * one_p;
    float pi_0 = 0.f, pi_1 = 1.f;
    MieResult r = { cx(0.f), cx(0.f), 0.f, 0.f, 0.f };
    for (int n = 1; n <= x_nmax; ++n) {
        float fn = (float) n;
        Cx jrx_n = jrx[n], jry_n = jry[n];
        Cx hx_n, hx_dx;
        if (n == 1) { hx_n = hx_1; hx_dx = s.x * hx_0 - fn * hx_1; }
        else { hx_n = ((2.f * fn - 1.f) * s.rcp_x) * hx_1 - hx_0; hx_dx = s.x * hx_1 - fn * hx_n; hx_0 = hx_1; hx_1 = hx_n; }
        Cx jx_n = jrx_n * jx_0, jy_n = jry_n * jy_0;
        Cx jx_dx = s.x * jx_0 - fn * jx_n, jy_dy = s.y * jy_0 - fn * jy_n;
        jx_0 = jx_n; jy_0 = jy_n;
        float pi_n, tau_n;
        if (n == 1) { pi_n = pi_1; tau_n = mu; }
        else {
            pi_n = ((2.f * fn - 1.f) / (fn - 1.f)) * mu * pi_1 - (fn / (fn - 1.f)) * pi_0;
            tau_n = fn * mu * pi_n - (fn + 1.f) * pi_1;
            pi_0 = pi_1; pi_1 = pi_n;
        }
        Cx a_n = (s.m_sq * jy_n * jx_dx - jx_n * jy_dy) / (s.m_sq * jy_n * hx_dx - hx_n * jy_dy);
        Cx b_n = (jy_n * jx_dx - jx_n * jy_dy) / (jy_n * hx_dx - hx_n * jy_dy);
        bool nan = std::isnan(a_n.re) || std::isnan(a_n.im) || std::isnan(b_n.re) || std::isnan(b_n.im);
        if (nan) continue;
        float cn = 2.f * fn + 1.f, kn = cn / (fn * (fn + 1.f));
        r.s1 = r.s1 + kn * (a_n * tau_n + b_n * pi_n);
        r.s2 = r.s2 + kn * (a_n * pi_n + b_n * tau_n);
        float nn = cn * (cx_sqnorm(a_n) + cx_sqnorm(b_n));
        r.ns += nn; r.cs += nn;
        r.ct += (cn * (a_n + b_n)).re;
    }
    Cx f = i * cx_rcp(s.kx);
    r.s1 = r.s1 * f; r.s2 = r.s2 * f;
    float k = TwoPi / cx_sqnorm(s.kx);
    r.ns *= k; r.cs *= k; r.ct *= k;
    return r;
}

// mueller::mie_scatter entries (m11, m12, m33, m34)
static inline void mie_scatter4(const MieResult &r, float *m) {
    float n1 = cx_sqnorm(r.s1), n2 = cx_sqnorm(r.s2);
    float s11 = 0.5f * (n1 + n2), s12 = 0.5f * (n1 - n2);
    Cx p = r.s1 * Cx{ r.s2.re, -r.s2.im };
    float rn = rcp(r.ns);
    m[0] = rn * s11; m[1] = rn * s12; m[2] = rn * p.re; m[3] = rn * p.im;
}

// eval_mie over a table of cos(theta): out = 4 arrays of n_mu (m11, m12, m33, m34).
// n_gauss == 1 with sdf == nullptr: monodisperse (radius[0]).
static inline void mie_tabulate(float wavelength, Cx ior_med, Cx ior_sph, int nmax, const float *radius, const float *weight,
                                const float *sdf, int n_gauss, const float *mu, int n_mu, float *m11, float *m12, float *m33, float *m34) {
    int x_nmax = nmax, y_nmax = nmax;
    if (nmax == -1) {
        float xm = 0.f, ym = 0.f;
        for (int g = 0; g < n_gauss; ++g) {
            MieSetup s = mie_setup(wavelength, radius[g], ior_med, ior_sph);
            xm = maximum(xm, 8.f + s.x_norm + 4.05f * cbrt_(s.x_norm));
            ym = maximum(ym, 8.f + s.y_norm + 4.05f * cbrt_(s.y_norm));
        }
        x_nmax = (int) xm; y_nmax = (int) ym;
    }
    for (int k = 0; k < n_mu; ++k) {
        if (n_gauss == 1 && !sdf) {
            float m[4];
            mie_scatter4(mie_lane(wavelength, mu[k], radius[0], ior_med, ior_sph, x_nmax, y_nmax), m);
            m11[k] = m[0]; m12[k] = m[1]; m33[k] = m[2]; m34[k] = m[3];
            continue;
        }
        float cs_avg = 0.f, acc[4] = { 0.f, 0.f, 0.f, 0.f };
        for (int g = 0; g < n_gauss; ++g) {
            MieResult r = mie_lane(wavelength, mu[k], radius[g], ior_med, ior_sph, x_nmax, y_nmax);
            float m[4];
            mie_scatter4(r, m);
            float wc = (weight[g] * sdf[g]) * r.cs;
            cs_avg += wc;
            for (int q = 0; q < 4; ++q) acc[q] += wc * m[q];
        }
        m11[k] = acc[0] / cs_avg; m12[k] = acc[1] / cs_avg; m33[k] = acc[2] / cs_avg; m34[k] = acc[3] / cs_avg;
    }
}

} // namespace orc

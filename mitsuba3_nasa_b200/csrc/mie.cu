// mie.cu -- Lorenz-Mie phase-matrix tabulation on the device (mib_mie_tabulate).
//   mie()                       include/mitsuba/render/mie.h:45-222
//   mueller::mie_scatter        include/mitsuba/render/mueller.h:282-295
//   MiePhaseFunction::eval_mie  src/phase/mie.cpp:117-215 (size-distribution average)
// The reference evaluates the whole series per (cos theta, radius) lane.  The Bessel-ratio
// recurrences and the coefficients a_n, b_n depend on the radius only, so they are computed
// once per radius (k_mie_coeffs, one thread per quadrature node) and the angular sums run as
// one thread per (cos theta, radius) pair (k_mie_angles) with the same float operations in
// the same order as the per-lane reference; k_mie_reduce does the quadrature sum.
#include "dmath.cuh"
#include "../../include/mi_b200.h"
#include <cuda_runtime.h>
#include <string>
#include <stdexcept>
#include <vector>
#include <cmath>

namespace mib {

struct Cx { float re, im; };
MIB_D Cx cx(float r, float i = 0.f) { Cx z; z.re = r; z.im = i; return z; }
MIB_D Cx operator+(Cx a, Cx b) { return cx(a.re + b.re, a.im + b.im); }
MIB_D Cx operator-(Cx a, Cx b) { return cx(a.re - b.re, a.im - b.im); }
MIB_D Cx operator-(Cx a) { return cx(-a.re, -a.im); }
MIB_D Cx operator*(Cx a, Cx b) { return cx(fms_(a.re, b.re, a.im * b.im), fma_(a.re, b.im, a.im * b.re)); }   // complex.h:81-88
MIB_D Cx operator*(Cx a, float s) { return cx(a.re * s, a.im * s); }
MIB_D Cx operator*(float s, Cx a) { return cx(s * a.re, s * a.im); }
MIB_D float cx_sqnorm(Cx z) { return fma_(z.im, z.im, z.re * z.re); }
MIB_D Cx cx_rcp(Cx z) { float r = rcp_(cx_sqnorm(z)); return cx(z.re * r, -z.im * r); }                       // complex.h:101-103
MIB_D Cx operator/(Cx a, Cx b) { return a * cx_rcp(b); }

MIB_D void sincosh_(float x, float &sh, float &ch) {                                                          // math.h:1112-1158
    float xa = fabsf(x), exp0 = exp_(x), exp1 = rcp_(exp0);
    float r_big = (exp0 - exp1) * 0.5f;
    float x2 = sqr_(x);
    float y = fma_(sqr_(x2), 2.03721912945e-4f, fma_(x2, 8.33028376239e-3f, 1.66667160211e-1f));               // estrin, 3 coefficients
    float r_small = fma_(y, x2 * x, x);
    sh = xa > 1.f ? r_big : r_small;
    ch = (exp0 + exp1) * .5f;
}
MIB_D Cx cx_sin(Cx z) { float s, c, sh, ch; sincos_(z.re, s, c); sincosh_(z.im, sh, ch); return cx(s * ch, c * sh); }
MIB_D Cx cx_exp(Cx z) { float e = exp_(z.re), s, c; sincos_(z.im, s, c); return cx(e * c, e * s); }

struct MieSetup { Cx m_sq, x, y, rcp_x, rcp_y, kx; float x_norm, y_norm; };
MIB_D MieSetup mie_setup(float wavelength, float radius, Cx ior_med, Cx ior_sph) {
    MieSetup s;
    Cx m = ior_sph * cx_rcp(ior_med);
    float rw = rcp_(wavelength);
    s.kx = (kTwoPi * ior_med) * rw;
    Cx ky = (kTwoPi * ior_sph) * rw;
    s.x = s.kx * radius; s.y = ky * radius;
    s.m_sq = m * m;
    s.rcp_x = cx_rcp(s.x); s.rcp_y = cx_rcp(s.y);
    s.x_norm = sqrtf(cx_sqnorm(s.x)); s.y_norm = sqrtf(cx_sqnorm(s.y));
    return s;
}

struct MieParams {
    float wavelength;
    Cx ior_med, ior_sph;
    int x_nmax, y_nmax, x_ndown, y_ndown, n_gauss, n_mu, mono;
};

// per-radius stopping-criterion terms 8 + |x| + 4.05 cbrt(|x|) (mie.h:76-79): out[2*g], out[2*g+1]
static __global__ void k_mie_norms(MieParams p, const float *radius, float *out) {
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= p.n_gauss) return;
    MieSetup s = mie_setup(p.wavelength, radius[g], p.ior_med, p.ior_sph);
    out[2 * g] = 8.f + s.x_norm + 4.05f * cbrt_(s.x_norm);
    out[2 * g + 1] = 8.f + s.y_norm + 4.05f * cbrt_(s.y_norm);
}

// one thread per radius: Bessel ratios (downward), then a_n, b_n (upward); ab[g][n] = (a.re, a.im, b.re, b.im), n = 1..x_nmax
static __global__ void k_mie_coeffs(MieParams p, const float *radius, Cx *scratch_x, Cx *scratch_y, float4 *ab, float *sums /*[g][3]*/) {
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= p.n_gauss) return;
    MieSetup s = mie_setup(p.wavelength, radius[g], p.ior_med, p.ior_sph);
    const Cx i = cx(0.f, 1.f);
    Cx *jrx = scratch_x + (size_t) g * p.x_ndown, *jry = scratch_y + (size_t) g * p.y_ndown;
    Cx j = s.x * rcp_(2.f * (float) p.x_ndown + 1.f);
    for (int n = p.x_ndown - 1; n >= 1; --n) { Cx k = (float) (2 * n + 1) * s.rcp_x; j = cx_rcp(k - j); jrx[n] = j; }
    j = s.y * rcp_(2.f * (float) p.y_ndown + 1.f);
    for (int n = p.y_ndown - 1; n >= 1; --n) { Cx k = (float) (2 * n + 1) * s.rcp_y; j = cx_rcp(k - j); jry[n] = j; }

    Cx jx_0 = cx_sin(s.x) * s.rcp_x, jy_0 = cx_sin(s.y) * s.rcp_y;
    Cx h_exp = cx_exp(i * s.x) * s.rcp_x, hx_0 = (-i) * h_exp;
    Cx one_p = i * s.rcp_x; one_p.re = 1.f + one_p.re;
    Cx hx_1 = (-h_exp) * one_p;
    float ns = 0.f, ct = 0.f;
    float4 *abg = ab + (size_t) g * (p.x_nmax + 1);
    for (int n = 1; n <= p.x_nmax; ++n) {
        float fn = (float) n;
        Cx hx_n, hx_dx;
        if (n == 1) { hx_n = hx_1; hx_dx = s.x * hx_0 - fn * hx_1; }
        else { hx_n = ((2.f * fn - 1.f) * s.rcp_x) * hx_1 - hx_0; hx_dx = s.x * hx_1 - fn * hx_n; hx_0 = hx_1; hx_1 = hx_n; }
        Cx jx_n = jrx[n] * jx_0, jy_n = jry[n] * jy_0;
        Cx jx_dx = s.x * jx_0 - fn * jx_n, jy_dy = s.y * jy_0 - fn * jy_n;
        jx_0 = jx_n; jy_0 = jy_n;
        Cx a_n = (s.m_sq * jy_n * jx_dx - jx_n * jy_dy) / (s.m_sq * jy_n * hx_dx - hx_n * jy_dy);
        Cx b_n = (jy_n * jx_dx - jx_n * jy_dy) / (jy_n * hx_dx - hx_n * jy_dy);
        abg[n] = make_float4(a_n.re, a_n.im, b_n.re, b_n.im);
        bool nan = a_n.re != a_n.re || a_n.im != a_n.im || b_n.re != b_n.re || b_n.im != b_n.im;
        if (nan) continue;
        float cn = 2.f * fn + 1.f;
        ns += cn * (cx_sqnorm(a_n) + cx_sqnorm(b_n));
        ct += (cn * (a_n + b_n)).re;
    }
    float k = __fdiv_rn(kTwoPi, cx_sqnorm(s.kx));
    sums[3 * g] = ns * k;        // Ns == Cs (mie.h:197-200)
    sums[3 * g + 1] = ct * k;
    sums[3 * g + 2] = 0.f;
}

// one thread per (mu, radius): angular sums S1, S2 -> weighted phase-matrix entries part[g][k] = (wc*m11, wc*m12, wc*m33, wc*m34), wcs[g][k] = wc
static __global__ void k_mie_angles(MieParams p, const float *radius, const float *weight, const float *sdf, const float *mu,
                                    const float4 *ab, const float *sums, float4 *part, float *wcs) {
    int k = blockIdx.x * blockDim.x + threadIdx.x, g = blockIdx.y;
    if (k >= p.n_mu) return;
    MieSetup s = mie_setup(p.wavelength, radius[g], p.ior_med, p.ior_sph);
    const float m = mu[k];
    const float4 *abg = ab + (size_t) g * (p.x_nmax + 1);
    float pi_0 = 0.f, pi_1 = 1.f;
    Cx s1 = cx(0.f), s2 = cx(0.f);
    for (int n = 1; n <= p.x_nmax; ++n) {
        float fn = (float) n;
        float pi_n, tau_n;
        if (n == 1) { pi_n = pi_1; tau_n = m; }
        else {
            pi_n = (__fdiv_rn(2.f * fn - 1.f, fn - 1.f) * m) * pi_1 - __fdiv_rn(fn, fn - 1.f) * pi_0;
            tau_n = (fn * m) * pi_n - (fn + 1.f) * pi_1;
            pi_0 = pi_1; pi_1 = pi_n;
        }
        float4 v = __ldg(abg + n);
        Cx a_n = cx(v.x, v.y), b_n = cx(v.z, v.w);
        bool nan = v.x != v.x || v.y != v.y || v.z != v.z || v.w != v.w;
        if (nan) continue;
        float cn = 2.f * fn + 1.f, kn = __fdiv_rn(cn, fn * (fn + 1.f));
        s1 = s1 + kn * (a_n * tau_n + b_n * pi_n);
        s2 = s2 + kn * (a_n * pi_n + b_n * tau_n);
    }
    Cx f = cx(0.f, 1.f) * cx_rcp(s.kx);
    s1 = s1 * f; s2 = s2 * f;
    float ns = sums[3 * g];
    float n1 = cx_sqnorm(s1), n2 = cx_sqnorm(s2);
    float s11 = 0.5f * (n1 + n2), s12 = 0.5f * (n1 - n2);
    Cx pr = s1 * cx(s2.re, -s2.im);
    float rn = rcp_(ns);
    float m11 = rn * s11, m12 = rn * s12, m33 = rn * pr.re, m34 = rn * pr.im;
    size_t o = (size_t) g * p.n_mu + k;
    if (p.mono) { part[o] = make_float4(m11, m12, m33, m34); wcs[o] = 1.f; return; }
    float wc = (weight[g] * sdf[g]) * ns;          // Cs == Ns
    part[o] = make_float4(wc * m11, wc * m12, wc * m33, wc * m34);
    wcs[o] = wc;
}

static __global__ void k_mie_reduce(MieParams p, const float4 *part, const float *wcs, float *m11, float *m12, float *m33, float *m34) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= p.n_mu) return;
    if (p.mono) { float4 v = part[k]; m11[k] = v.x; m12[k] = v.y; m33[k] = v.z; m34[k] = v.w; return; }
    float cs = 0.f, a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    for (int g = 0; g < p.n_gauss; ++g) {
        size_t o = (size_t) g * p.n_mu + k;
        float4 v = part[o];
        cs += wcs[o]; a0 += v.x; a1 += v.y; a2 += v.z; a3 += v.w;
    }
    m11[k] = __fdiv_rn(a0, cs); m12[k] = __fdiv_rn(a1, cs); m33[k] = __fdiv_rn(a2, cs); m34[k] = __fdiv_rn(a3, cs);
}

} // namespace mib

extern std::string &mib_error_slot();

#define MCK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) throw std::runtime_error(std::string(#x) + ": " + cudaGetErrorString(e_)); } while (0)

extern "C" int mib_mie_tabulate(int device, float wavelength, float ior_med_re, float ior_med_im, float ior_sph_re, float ior_sph_im,
                                int nmax, const float *radius, const float *weight, const float *sdf, int n_gauss,
                                const float *mu, int n_mu, float *m11, float *m12, float *m33, float *m34) {
    using namespace mib;
    try {
        if (!radius || !mu || !m11 || !m12 || !m33 || !m34 || n_gauss < 1 || n_mu < 1) throw std::runtime_error("mib_mie_tabulate: bad arguments");
        if (nmax < -1) throw std::runtime_error("The number of series terms must be positive or -1 (automatic)!");       // mie.cpp:88-89
        if (ior_med_re <= 0 || ior_sph_re <= 0) throw std::runtime_error("Indices of refraction must be positive!");   // mie.cpp:91-92
        if (!(wavelength > 0)) throw std::runtime_error("mib_mie_tabulate: wavelength must be positive");
        MCK(cudaSetDevice(device));
        MieParams p;
        p.wavelength = wavelength;
        p.ior_med.re = ior_med_re; p.ior_med.im = ior_med_im; p.ior_sph.re = ior_sph_re; p.ior_sph.im = ior_sph_im;
        p.n_gauss = n_gauss; p.n_mu = n_mu; p.mono = (n_gauss == 1 && !sdf) ? 1 : 0;
        if (!p.mono && (!weight || !sdf)) throw std::runtime_error("mib_mie_tabulate: quadrature weights / size distribution missing");
        float *d_rad = nullptr, *d_w = nullptr, *d_sdf = nullptr, *d_mu = nullptr, *d_norms = nullptr, *d_sums = nullptr, *d_wcs = nullptr, *d_out = nullptr;
        Cx *d_sx = nullptr, *d_sy = nullptr;
        float4 *d_ab = nullptr, *d_part = nullptr;
        auto cleanup = [&]() { cudaFree(d_rad); cudaFree(d_w); cudaFree(d_sdf); cudaFree(d_mu); cudaFree(d_norms); cudaFree(d_sums);
                               cudaFree(d_wcs); cudaFree(d_out); cudaFree(d_sx); cudaFree(d_sy); cudaFree(d_ab); cudaFree(d_part); };
        try {
            MCK(cudaMalloc(&d_rad, n_gauss * sizeof(float)));
            MCK(cudaMemcpy(d_rad, radius, n_gauss * sizeof(float), cudaMemcpyHostToDevice));
            MCK(cudaMalloc(&d_w, n_gauss * sizeof(float)));
            MCK(cudaMalloc(&d_sdf, n_gauss * sizeof(float)));
            if (!p.mono) {
                MCK(cudaMemcpy(d_w, weight, n_gauss * sizeof(float), cudaMemcpyHostToDevice));
                MCK(cudaMemcpy(d_sdf, sdf, n_gauss * sizeof(float), cudaMemcpyHostToDevice));
            }
            MCK(cudaMalloc(&d_mu, n_mu * sizeof(float)));
            MCK(cudaMemcpy(d_mu, mu, n_mu * sizeof(float), cudaMemcpyHostToDevice));
            p.x_nmax = nmax; p.y_nmax = nmax;
            if (nmax == -1) {       // max_nested over all lanes (mie.h:76-79)
                MCK(cudaMalloc(&d_norms, 2 * n_gauss * sizeof(float)));
                k_mie_norms<<<(n_gauss + 63) / 64, 64>>>(p, d_rad, d_norms);
                MCK(cudaGetLastError());
                std::vector<float> h(2 * n_gauss);
                MCK(cudaMemcpy(h.data(), d_norms, h.size() * sizeof(float), cudaMemcpyDeviceToHost));
                float xm = 0.f, ym = 0.f;
                for (int g = 0; g < n_gauss; ++g) { xm = fmaxf(xm, h[2 * g]); ym = fmaxf(ym, h[2 * g + 1]); }
                p.x_nmax = (int) xm; p.y_nmax = (int) ym;
            }
            p.x_ndown = p.x_nmax + 8 * (int) std::sqrt((double) p.x_nmax) + 3;
            p.y_ndown = p.y_nmax + 8 * (int) std::sqrt((double) p.y_nmax) + 3;
            if (p.x_nmax >= p.y_ndown) throw std::runtime_error("mib_mie_tabulate: series length exceeds the Bessel-ratio table of the sphere medium");
            MCK(cudaMalloc(&d_sx, (size_t) n_gauss * p.x_ndown * sizeof(Cx)));
            MCK(cudaMalloc(&d_sy, (size_t) n_gauss * p.y_ndown * sizeof(Cx)));
            MCK(cudaMalloc(&d_ab, (size_t) n_gauss * (p.x_nmax + 1) * sizeof(float4)));
            MCK(cudaMalloc(&d_sums, 3 * n_gauss * sizeof(float)));
            MCK(cudaMalloc(&d_part, (size_t) n_gauss * n_mu * sizeof(float4)));
            MCK(cudaMalloc(&d_wcs, (size_t) n_gauss * n_mu * sizeof(float)));
            MCK(cudaMalloc(&d_out, 4 * (size_t) n_mu * sizeof(float)));
            k_mie_coeffs<<<(n_gauss + 31) / 32, 32>>>(p, d_rad, d_sx, d_sy, d_ab, d_sums);
            MCK(cudaGetLastError());
            k_mie_angles<<<dim3((n_mu + 127) / 128, n_gauss), 128>>>(p, d_rad, d_w, d_sdf, d_mu, d_ab, d_sums, d_part, d_wcs);
            MCK(cudaGetLastError());
            k_mie_reduce<<<(n_mu + 127) / 128, 128>>>(p, d_part, d_wcs, d_out, d_out + n_mu, d_out + 2 * n_mu, d_out + 3 * n_mu);
            MCK(cudaGetLastError());
            MCK(cudaMemcpy(m11, d_out, n_mu * sizeof(float), cudaMemcpyDeviceToHost));
            MCK(cudaMemcpy(m12, d_out + n_mu, n_mu * sizeof(float), cudaMemcpyDeviceToHost));
            MCK(cudaMemcpy(m33, d_out + 2 * n_mu, n_mu * sizeof(float), cudaMemcpyDeviceToHost));
            MCK(cudaMemcpy(m34, d_out + 3 * n_mu, n_mu * sizeof(float), cudaMemcpyDeviceToHost));
        } catch (...) { cleanup(); throw; }
        cleanup();
    } catch (const std::exception &e) { mib_error_slot() = e.what(); return 1; }
    return 0;
}

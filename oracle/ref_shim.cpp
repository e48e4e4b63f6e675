// ref_shim.cpp -- TEST INFRASTRUCTURE.  Thin extern "C" wrappers around the
// REFERENCE'S OWN HEADERS (compiled where they lie under /root/reference; no
// reference source is copied into this repository).  Built by oracle/Makefile
// into oracle/_ref/libmiref.so and used by tests to pin the building blocks of
// the oracle restatement (scalar float instantiation of the same templates the
// llvm_ad variants trace).
#include <mitsuba/core/fwd.h>
#include <mitsuba/core/vector.h>
#include <mitsuba/core/frame.h>
#include <mitsuba/core/random.h>
#include <mitsuba/core/warp.h>
#include <mitsuba/core/spectrum.h>
#include <mitsuba/core/transform.h>
#include <mitsuba/core/distr_1d.h>
#include <mitsuba/render/mueller.h>
#include <mitsuba/render/fresnel.h>
#include <mitsuba/render/microfacet.h>
#include <mitsuba/render/interaction.h>
#include <mitsuba/render/mie.h>
#include <drjit/dynamic.h>
#include <drjit/math.h>
#include <drjit/random.h>
#include <drjit/sphere.h>
#include <drjit/texture.h>
#include <cstring>


// Link stubs for the two non-template symbols the headers reference (libmitsuba is not built here).
#include <stdexcept>
namespace mitsuba {
    const Class *m_class = nullptr;
    namespace detail {
        void (Throw)(LogLevel, const Class *, const char *, int, const std::string &msg) { throw std::runtime_error(msg); }
    }
}

// mie.h switches JitFlag::LoopRecord off around its loops (drjit-core is not linked here; scalar arrays ignore the flag)
extern "C" {
    void jit_set_flag(JitFlag, int) { }
    int jit_flag(JitFlag) { return 0; }
}

using namespace mitsuba;
using Float = float;
using V3 = Vector<float, 3>;
using P3 = Point<float, 3>;
using P2 = Point<float, 2>;
using MM = MuellerMatrix<float>;
using SpecPol = MuellerMatrix<Color<float, 1>>;

static void store(const MM &M, float *o) { for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) o[4 * r + c] = M(r, c); }
static MM load(const float *i) { MM M; for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) M(r, c) = i[4 * r + c]; return M; }

extern "C" {

void ref_sincos(const float *x, float *s, float *c, size_t n) { for (size_t i = 0; i < n; ++i) { auto [a, b] = dr::sincos(x[i]); s[i] = a; c[i] = b; } }
void ref_exp(const float *x, float *y, size_t n) { for (size_t i = 0; i < n; ++i) y[i] = dr::exp(x[i]); }
void ref_log(const float *x, float *y, size_t n) { for (size_t i = 0; i < n; ++i) y[i] = dr::log(x[i]); }
void ref_cbrt(const float *x, float *y, size_t n) { for (size_t i = 0; i < n; ++i) y[i] = dr::cbrt(x[i]); }
void ref_asin(const float *x, float *y, size_t n) { for (size_t i = 0; i < n; ++i) y[i] = dr::asin(x[i]); }
void ref_tea(const uint32_t *v0, const uint32_t *v1, uint32_t *o0, uint32_t *o1, size_t n) {
    for (size_t i = 0; i < n; ++i) { auto [a, b] = sample_tea_32<uint32_t>(v0[i], v1[i]); o0[i] = a; o1[i] = b; }
}
void ref_pcg32_floats(uint64_t initstate, uint64_t initseq, float *out, size_t n) {
    dr::PCG32<uint32_t> r; r.seed(1, initstate, initseq);
    for (size_t i = 0; i < n; ++i) out[i] = r.next_float32();
}
void ref_pcg32_uints(uint64_t initstate, uint64_t initseq, uint32_t *out, size_t n) {
    dr::PCG32<uint32_t> r; r.seed(1, initstate, initseq);
    for (size_t i = 0; i < n; ++i) out[i] = r.next_uint32();
}
void ref_coordinate_system(const float *n3, float *s3, float *t3, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        auto [s, t] = coordinate_system(V3(n3[3*i], n3[3*i+1], n3[3*i+2]));
        for (int k = 0; k < 3; ++k) { s3[3*i+k] = s[k]; t3[3*i+k] = t[k]; }
    }
}
void ref_unit_angle(const float *a3, const float *b3, float *out, size_t n) {
    for (size_t i = 0; i < n; ++i)
        out[i] = dr::unit_angle(V3(a3[3*i], a3[3*i+1], a3[3*i+2]), V3(b3[3*i], b3[3*i+1], b3[3*i+2]));
}
void ref_rotator(const float *theta, float *m16, size_t n) { for (size_t i = 0; i < n; ++i) store(mueller::rotator<float>(theta[i]), m16 + 16*i); }
void ref_rayleigh_scatter(const float *ct, const float *st, float rho, float *m16, size_t n) {
    for (size_t i = 0; i < n; ++i) store(mueller::rayleigh_scatter<float>(ct[i], st[i], rho), m16 + 16*i);
}
void ref_rotate_mueller_basis(const float *m16_in, const float *v18, float *m16_out, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        const float *v = v18 + 18*i;
        MM R = mueller::rotate_mueller_basis(load(m16_in + 16*i), V3(v[0],v[1],v[2]), V3(v[3],v[4],v[5]), V3(v[6],v[7],v[8]),
                                             V3(v[9],v[10],v[11]), V3(v[12],v[13],v[14]), V3(v[15],v[16],v[17]));
        store(R, m16_out + 16*i);
    }
}
void ref_mei_to_world_mueller(const float *m16_in, const float *d3, const float *in3, const float *out3, float *m16_out, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        MediumInteraction<float, SpecPol> mei = dr::zeros<MediumInteraction<float, SpecPol>>();
        mei.sh_frame = Frame<float>(V3(d3[3*i], d3[3*i+1], d3[3*i+2]));
        MM Min = load(m16_in + 16*i);
        SpecPol S;
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) S(r, c) = Color<float, 1>(Min(r, c));
        SpecPol R = mei.to_world_mueller(S, V3(in3[3*i], in3[3*i+1], in3[3*i+2]), V3(out3[3*i], out3[3*i+1], out3[3*i+2]));
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m16_out[16*i + 4*r + c] = R(r, c)[0];
    }
}
void ref_square_to_uniform_sphere(const float *s2, float *o3, size_t n) {
    for (size_t i = 0; i < n; ++i) { V3 v = warp::square_to_uniform_sphere(P2(s2[2*i], s2[2*i+1])); for (int k = 0; k < 3; ++k) o3[3*i+k] = v[k]; }
}
void ref_square_to_cosine_hemisphere(const float *s2, float *o3, size_t n) {
    for (size_t i = 0; i < n; ++i) { V3 v = warp::square_to_cosine_hemisphere(P2(s2[2*i], s2[2*i+1])); for (int k = 0; k < 3; ++k) o3[3*i+k] = v[k]; }
}
void ref_texture3_eval(const float *data, int rx, int ry, int rz, int ch, int filter, int wrap, const float *p3, float *out, size_t n) {
    using Tex = dr::Texture<float, 3>;
    size_t shape[3] = { (size_t) rz, (size_t) ry, (size_t) rx };
    dr::FilterMode fm = filter == 0 ? dr::FilterMode::Nearest : dr::FilterMode::Linear;
    dr::WrapMode wm = wrap == 0 ? dr::WrapMode::Repeat : (wrap == 1 ? dr::WrapMode::Mirror : dr::WrapMode::Clamp);
    Tex tex(shape, (size_t) ch, false, fm, wm);
    size_t total = (size_t) rx * ry * rz * ch;
    dr::DynamicArray<float> arr = dr::load<dr::DynamicArray<float>>(data, total);
    tex.set_value(arr);
    for (size_t i = 0; i < n; ++i) {
        dr::Array<float, 3> p(p3[3*i], p3[3*i+1], p3[3*i+2]);
        float o[3] = { 0, 0, 0 };
        tex.eval_nonaccel(p, o);
        for (int k = 0; k < ch; ++k) out[ch*i + k] = o[k];
    }
}
void ref_irregular(const float *nodes, const float *pdf, uint32_t size, const float *x, float *eval_out, float *evaln_out,
                   const float *u, float *sample_out, size_t n, float *integral) {
    IrregularContinuousDistribution<float> d(nodes, pdf, size);
    *integral = d.integral();
    for (size_t i = 0; i < n; ++i) { eval_out[i] = d.eval_pdf(x[i]); evaln_out[i] = d.eval_pdf_normalized(x[i], true); sample_out[i] = d.sample(u[i]); }
}
void ref_transform_point(const float *m16, const float *p3, float *o3, size_t n) {
    Transform<Point<float, 4>> T; dr::Matrix<float, 4> M;
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) M(r, c) = m16[4*r + c];
    T = Transform<Point<float, 4>>(M);
    for (size_t i = 0; i < n; ++i) { P3 v = T * P3(p3[3*i], p3[3*i+1], p3[3*i+2]); for (int k = 0; k < 3; ++k) o3[3*i+k] = v[k]; }
}
void ref_transform_vector(const float *m16, const float *p3, float *o3, size_t n) {
    dr::Matrix<float, 4> M;
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) M(r, c) = m16[4*r + c];
    Transform<Point<float, 4>> T(M);
    for (size_t i = 0; i < n; ++i) { V3 v = T * V3(p3[3*i], p3[3*i+1], p3[3*i+2]); for (int k = 0; k < 3; ++k) o3[3*i+k] = v[k]; }
}
void ref_fresnel(const float *ci, float eta, float *out4, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        auto [r, ct, eit, eti] = fresnel<float>(ci[i], eta);
        out4[4*i] = r; out4[4*i+1] = ct; out4[4*i+2] = eit; out4[4*i+3] = eti;
    }
}
void ref_fresnel_polarized(const float *ci, float eta, float *out7, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        auto [a_s, a_p, ct, eit, eti] = fresnel_polarized<float>(ci[i], eta); float *o = out7 + 7*i;
        o[0] = dr::real(a_s); o[1] = dr::imag(a_s); o[2] = dr::real(a_p); o[3] = dr::imag(a_p); o[4] = ct; o[5] = eit; o[6] = eti;
    }
}
void ref_specular_reflection(const float *ci, float eta, float *m16, size_t n) {
    for (size_t i = 0; i < n; ++i) store(mueller::specular_reflection<float>(ci[i], eta), m16 + 16*i);
}
void ref_specular_transmission(const float *ci, float eta, float *m16, size_t n) {
    for (size_t i = 0; i < n; ++i) store(mueller::specular_transmission<float>(ci[i], eta), m16 + 16*i);
}
void ref_refract(const float *wi3, const float *m3, float eta, float *o3, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        V3 wi(wi3[3*i], wi3[3*i+1], wi3[3*i+2]); Normal<float, 3> m(m3[3*i], m3[3*i+1], m3[3*i+2]);
        auto [r, ct, eit, eti] = fresnel<float>(dr::dot(wi, m), eta);
        V3 o = ct == 0.f ? V3(0.f) : refract(wi, m, ct, eti);
        for (int k = 0; k < 3; ++k) o3[3*i+k] = o[k];
    }
}
void ref_erf(const float *x, float *y, size_t n) { for (size_t i = 0; i < n; ++i) y[i] = dr::erf(x[i]); }
void ref_erfinv(const float *x, float *y, size_t n) { for (size_t i = 0; i < n; ++i) y[i] = dr::erfinv(x[i]); }
void ref_exp2(const float *x, float *y, size_t n) { for (size_t i = 0; i < n; ++i) y[i] = dr::exp2(x[i]); }
using MD = MicrofacetDistribution<float, Color<float, 1>>;
void ref_microfacet_eval(int type, float alpha_u, float alpha_v, const float *wi3, const float *m3, float *out3, size_t n) {
    MD d(type == 0 ? MicrofacetType::Beckmann : MicrofacetType::GGX, alpha_u, alpha_v, true);
    for (size_t i = 0; i < n; ++i) {
        V3 wi(wi3[3*i], wi3[3*i+1], wi3[3*i+2]), m(m3[3*i], m3[3*i+1], m3[3*i+2]);
        out3[3*i] = d.eval(m); out3[3*i+1] = d.smith_g1(wi, m); out3[3*i+2] = d.pdf(wi, m);
    }
}
void ref_microfacet_sample(int type, float alpha_u, float alpha_v, const float *wi3, const float *s2, float *out4, size_t n) {
    MD d(type == 0 ? MicrofacetType::Beckmann : MicrofacetType::GGX, alpha_u, alpha_v, true);
    for (size_t i = 0; i < n; ++i) {
        auto [m, pdf] = d.sample(V3(wi3[3*i], wi3[3*i+1], wi3[3*i+2]), P2(s2[2*i], s2[2*i+1]));
        out4[4*i] = m[0]; out4[4*i+1] = m[1]; out4[4*i+2] = m[2]; out4[4*i+3] = pdf;
    }
}

// eval_transmittance / eval_reflectance (microfacet.h:464-568): the tables roughplastic precomputes (roughplastic.cpp:231-250);
// n must be a multiple of the packet size.  fresnel_diffuse_reflectance (fresnel.h:328-355).
void ref_rough_transmittance(int type, float alpha, float eta, const float *mu, float *out_t, float *out_r_inv_eta, size_t n) {
    using FloatP = dr::Packet<float>;
    using FloatX = dr::DynamicArray<float>;
    MicrofacetDistribution<FloatP, Color<float, 1>> d(type == 0 ? MicrofacetType::Beckmann : MicrofacetType::GGX, alpha);
    FloatX m = dr::zeros<FloatX>(n), zero = dr::zeros<FloatX>(n);
    for (size_t i = 0; i < n; ++i) m[i] = mu[i];
    Vector<FloatX, 3> wi(dr::sqrt(1 - m * m), zero, m);
    FloatX t = eval_transmittance(d, wi, eta), r = eval_reflectance(d, wi, 1.f / eta);
    for (size_t i = 0; i < n; ++i) { out_t[i] = t[i]; out_r_inv_eta[i] = r[i]; }
}
float ref_fresnel_diffuse_reflectance(float eta) { return fresnel_diffuse_reflectance(eta); }

// Lorenz-Mie amplitudes (include/mitsuba/render/mie.h:45-222), one (mu, radius) pair per call so that the
// automatic series length (max_nested over the lanes, :76-79) is the lane's own: out7 = (re S1, im S1, re S2, im S2, Ns, Cs, Ct)
void ref_mie(float wavelength, const float *mu, const float *radius, float med_re, float med_im, float sph_re, float sph_im,
             int nmax, float *out7, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        auto [s1, s2, ns, cs, ct] = mie<float, float>(wavelength, mu[i], radius[i], dr::Complex<float>(med_re, med_im),
                                                      dr::Complex<float>(sph_re, sph_im), nmax);
        float *o = out7 + 7 * i;
        o[0] = dr::real(s1); o[1] = dr::imag(s1); o[2] = dr::real(s2); o[3] = dr::imag(s2); o[4] = ns; o[5] = cs; o[6] = ct;
    }
}
// mueller::mie_scatter (mueller.h:282-295): in7 as above -> m16
void ref_mie_scatter(const float *in7, float *m16, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        const float *q = in7 + 7 * i;
        store(mueller::mie_scatter<float, dr::Complex<float>>(dr::Complex<float>(q[0], q[1]), dr::Complex<float>(q[2], q[3]), q[4]), m16 + 16 * i);
    }
}
// quad::gauss_legendre (include/mitsuba/core/quad.h:27-90)
void ref_gauss_legendre(int n, float *nodes, float *weights) {
    auto [x, w] = quad::gauss_legendre<dr::DynamicArray<float>>(n);
    for (int i = 0; i < n; ++i) { nodes[i] = x[i]; weights[i] = w[i]; }
}
// mueller::stokes_basis / rotate_stokes_basis (mueller.h:347-350, :375-386): v9 = forward, basis_current, basis_target
void ref_stokes_basis(const float *f3, float *o3, size_t n) {
    for (size_t i = 0; i < n; ++i) { V3 b = mueller::stokes_basis(V3(f3[3*i], f3[3*i+1], f3[3*i+2])); o3[3*i] = b.x(); o3[3*i+1] = b.y(); o3[3*i+2] = b.z(); }
}
void ref_rotate_stokes_basis(const float *v9, float *m16, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        const float *q = v9 + 9 * i;
        store(mueller::rotate_stokes_basis(V3(q[0], q[1], q[2]), V3(q[3], q[4], q[5]), V3(q[6], q[7], q[8])), m16 + 16 * i);
    }
}

} // extern "C"

// mib_api.cu -- C ABI of libmi_b200.so (include/mi_b200.h) and the host-side
// render drivers: passes, seeding, lane ranges, film develop (the C++ side of
// SamplingIntegrator::render / render_1, src/render/integrator.cpp:150-743).
// There is NO CPU fallback: every entry point needs a CUDA device and fails
// with an error message otherwise.
#include "kernels.cuh"
#include "astage.cuh"
#include "dfresnel.cuh"
#include "dmicrofacet.cuh"
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include <string>
#include <vector>
#include <cmath>
#include <stdexcept>
#include <mutex>

using namespace mib;

static_assert(sizeof(DScene) < 32000, "DScene must fit the constant bank next to the other module constants");

// The device scene is a __constant__ symbol per kernel module: launches of different scenes from concurrent host threads
// are serialised from the upload to the end of the kernel (every render entry point synchronises its stream anyway).
static std::mutex g_launch_mutex;

static thread_local std::string g_err;
std::string &mib_error_slot() { return g_err; }   // shared with mie.cu

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            throw std::runtime_error(std::string(#call) + ": " + cudaGetErrorString(e_));         \
    } while (0)

struct mib_scene {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    DScene ds;                                   // host copy handed to kernels by value
    float *buf[MIB_MAX_BUFFERS] = {};
    size_t buf_n[MIB_MAX_BUFFERS] = {};
    float *grad[MIB_MAX_BUFFERS] = {};
    float *cdf[MIB_MAX_PHASES] = {};             // cdf of the m11 table per phase
    float *super[MIB_MAX_MEDIA] = {};            // majorant supergrids (fast mode)
    int super_n[MIB_MAX_MEDIA] = {};
    float *film_raw = nullptr, *film_dev = nullptr;
    size_t film_raw_n = 0, film_dev_n = 0;
    double *sums = nullptr;                      // [8] device
    unsigned long long *counters = nullptr;      // [8] device
    unsigned int *maxbits = nullptr;
    double *grad_scalar = nullptr;               // [MIB_MAX_MEDIA*4] device
    float *lane_L = nullptr;                     // adjoint: per-lane primal radiance
    size_t lane_L_n = 0;
    uint64_t lane_begin = 0, lane_end = 0;
    uint32_t stripe_rank = 0, stripe_world = 1;  // lane stripes (multi-GPU load balance), see mib_scene_set_lane_stripes
    uint64_t stripe_chunk = 0;
    mib_counters last = {};
    bool committed = false;
    uint32_t last_spp = 1;
};

static int raw_channels(int spec) { return spec == MIB_SPEC_MONO ? 2 : 4; }
static int dev_channels(int spec) { return spec == MIB_SPEC_MONO ? 1 : (spec == MIB_SPEC_RGB ? 3 : 4); }

static void ensure_film(mib_scene *s) {
    const mib_sensor &sn = s->ds.d.sensor;
    size_t npix = (size_t) sn.crop_width * sn.crop_height;
    size_t nr = npix * raw_channels(s->ds.d.spectrum), nd = npix * dev_channels(s->ds.d.spectrum);
    if (nr != s->film_raw_n) {
        if (s->film_raw) CK(cudaFree(s->film_raw));
        CK(cudaMalloc(&s->film_raw, nr * sizeof(float)));
        s->film_raw_n = nr;
    }
    if (nd != s->film_dev_n) {
        if (s->film_dev) CK(cudaFree(s->film_dev));
        CK(cudaMalloc(&s->film_dev, nd * sizeof(float)));
        s->film_dev_n = nd;
    }
}

static float device_max(mib_scene *s, const float *data, size_t n) {
    CK(cudaMemsetAsync(s->maxbits, 0, sizeof(unsigned int), s->stream));
    int blocks = (int) std::min<size_t>((n + 255) / 256, 148 * 8);
    k_max<<<blocks, 256, 0, s->stream>>>(data, n, s->maxbits);
    CK(cudaGetLastError());
    unsigned int b = 0;
    CK(cudaMemcpyAsync(&b, s->maxbits, sizeof(b), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    b = (b & 0x80000000u) ? (b & 0x7fffffffu) : ~b;
    float f;
    memcpy(&f, &b, 4);
    return f;
}

// IrregularContinuousDistribution::compute_cdf_scalar (include/mitsuba/core/distr_1d.h:916-975)
static void build_table(mib_scene *s, int phase, int which, int buf_nodes, int buf_pdf, uint32_t n, bool sampling) {
    DTable &t = s->ds.tables[phase][which];
    memset(&t, 0, sizeof(t));
    if (buf_nodes < 0 || buf_pdf < 0 || !s->buf[buf_nodes] || !s->buf[buf_pdf])
        throw std::runtime_error("tabulated phase function: table buffer missing");
    if (n < 2) throw std::runtime_error("IrregularContinuousDistribution: needs at least two entries!");
    if (s->buf_n[buf_nodes] < n || s->buf_n[buf_pdf] < n) throw std::runtime_error("tabulated phase function: table buffer too small");
    std::vector<float> nodes(n), pdf(n), cdf(n, 0.f);
    CK(cudaMemcpy(nodes.data(), s->buf[buf_nodes], n * sizeof(float), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(pdf.data(), s->buf[buf_pdf], n * sizeof(float), cudaMemcpyDeviceToHost));
    uint32_t vx = (uint32_t) -1, vy = (uint32_t) -1;
    float rx = INFINITY, ry = -INFINITY;
    double integral = 0.;
    for (uint32_t i = 0; i + 1 < n; ++i) {
        double x0 = nodes[i], x1 = nodes[i + 1], y0 = pdf[i], y1 = pdf[i + 1];
        double value = 0.5 * (x1 - x0) * (y0 + y1);
        rx = fminf(rx, (float) x0);
        ry = fmaxf(ry, (float) x1);
        integral += value;
        cdf[i] = (float) integral;
        if (!(x1 > x0)) throw std::runtime_error("IrregularContinuousDistribution: node positions must be strictly increasing!");
        if (sampling) {
            if (y0 < 0. || y1 < 0.) throw std::runtime_error("IrregularContinuousDistribution: entries must be non-negative!");
            else if (value > 0.) { if (vx == (uint32_t) -1) vx = i; vy = i; }
        }
    }
    t.nodes = s->buf[buf_nodes];
    t.pdf = s->buf[buf_pdf];
    t.n = n;
    t.range_x = rx; t.range_y = ry;
    if (sampling) {
        if (vx == (uint32_t) -1) throw std::runtime_error("IrregularContinuousDistribution: no probability mass found!");
        t.valid_x = vx; t.valid_y = vy;
        t.integral = cdf[vy];
        t.normalization = 1.f / t.integral;
        if (s->cdf[phase]) CK(cudaFree(s->cdf[phase]));
        CK(cudaMalloc(&s->cdf[phase], n * sizeof(float)));
        CK(cudaMemcpy(s->cdf[phase], cdf.data(), n * sizeof(float), cudaMemcpyHostToDevice));
        t.cdf = s->cdf[phase];
    }
}

static bool has_dielectric(const mib_scene *s) {
    const mib_scene_desc &d = s->ds.d;
    for (int i = 0; i < d.n_bsdfs; ++i)
        if (d.bsdfs[i].type >= MIB_BSDF_DIELECTRIC) return true;       // every BSDF beyond null / diffuse: dbsdf.cuh's general interface
    return false;
}

// triangle meshes / uv textures: walked by the forward one-thread-one-path kernels only (dmesh.cuh)
static bool has_mesh(const mib_scene *s) {
    const mib_scene_desc &d = s->ds.d;
    for (int i = 0; i < d.n_shapes; ++i) if (d.shapes[i].type == MIB_SHAPE_MESH) return true;
    for (int i = 0; i < d.n_bsdfs; ++i) if (d.bsdfs[i].texture != MIB_TEXTURE_UNIFORM) return true;
    for (int i = 0; i < d.n_emitters; ++i) if (d.emitters[i].type == MIB_EMITTER_ENVMAP || d.emitters[i].type == MIB_EMITTER_SPOT) return true;
    return false;
}

// volpathaos.cpp:262-285: ANY non-null BSDF event after which the ray keeps the sign of its z component counts as a
// refraction through "the" interface and switches the refractive next-event estimation on or off -- meant for the ocean surface,
// but a diffuse bounce off a surface that is not horizontal qualifies too.  Scenes where that can happen (a dielectric, or a
// non-null BSDF on anything but a horizontal rectangle) are walked by the kernel that carries this state (docean.cuh).
static bool aos_interface_state_needed(const mib_scene *s) {
    const mib_scene_desc &d = s->ds.d;
    if (has_dielectric(s)) return true;
    for (int i = 0; i < d.n_shapes; ++i) {
        const mib_shape &sh = d.shapes[i];
        if (d.bsdfs[sh.bsdf].type == MIB_BSDF_NULL) continue;
        if (sh.type != MIB_SHAPE_RECTANGLE) return true;
        if (!(sh.frame_n[0] == 0.f && sh.frame_n[1] == 0.f && fabsf(sh.frame_n[2]) == 1.f)) return true;
    }
    return false;
}

static void commit(mib_scene *s) {
    mib_scene_desc &d = s->ds.d;
    const bool forward_integrator = d.integrator.type == MIB_INTEG_VOLPATH || d.integrator.type == MIB_INTEG_VOLPATHAOS || d.integrator.type == MIB_INTEG_VOLPATHMIS;
    if (has_mesh(s) && !forward_integrator)
        throw std::runtime_error("triangle meshes, textured BSDFs, environment maps and spot lights are walked by the forward integrators (volpath, volpathaos, volpathmis) only");
    if (d.integrator.type == MIB_INTEG_VOLPATHMIS && d.spectrum == MIB_SPEC_MONO_POLARIZED)
        throw std::runtime_error("This integrator currently does not support polarized mode!");                  // volpathmis.cpp:134-136
    for (int i = 0; i < d.n_shapes; ++i) {
        const mib_shape &sh = d.shapes[i];
        if (sh.type != MIB_SHAPE_MESH) continue;
        const size_t V = (size_t) sh.mesh_vertices, F = (size_t) sh.mesh_faces, N = (size_t) sh.mesh_nodes;
        const size_t need = 3 * V + ((sh.mesh_flags & MIB_MESH_HAS_NORMALS) ? 3 * V : 0) + ((sh.mesh_flags & MIB_MESH_HAS_TEXCOORDS) ? 2 * V : 0) + 3 * F + 8 * N + F +
                            ((sh.mesh_flags & MIB_MESH_HAS_AREA_TABLE) ? 2 * F : 0);
        if (sh.mesh_buffer < 0 || sh.mesh_buffer >= MIB_MAX_BUFFERS || !s->buf[sh.mesh_buffer] || s->buf_n[sh.mesh_buffer] < need || V == 0 || F == 0 || N == 0)
            throw std::runtime_error("mesh: data buffer missing or too small for its vertex / face / node counts");
        if (sh.emitter >= 0 && !(sh.mesh_flags & MIB_MESH_HAS_AREA_TABLE)) throw std::runtime_error("mesh: an area emitter needs the mesh's area table");
    }
    for (int i = 0; i < d.n_emitters; ++i) {
        const mib_emitter &e = d.emitters[i];
        if (e.type != MIB_EMITTER_ENVMAP) continue;
        const size_t texels = (size_t) 3 * (size_t) e.env_res[0] * (size_t) e.env_res[1];
        if (e.env_buffer < 0 || e.env_buffer >= MIB_MAX_BUFFERS || !s->buf[e.env_buffer] || e.env_res[0] < 3 || e.env_res[1] < 3 ||
            e.env_levels < 2 || e.env_levels > MIB_MAX_ENV_LEVELS || (size_t) e.env_level_offset[0] != texels ||
            s->buf_n[e.env_buffer] < texels + (size_t) e.env_res[0] * (size_t) e.env_res[1])
            throw std::runtime_error("envmap: data buffer missing or inconsistent with its resolution / warp levels");
    }
    // the ocean interface: walked by the forward volpathaos kernel of docean.cuh only
    if (has_dielectric(s) && !forward_integrator)
        throw std::runtime_error("dielectric / roughdielectric / thindielectric / plastic / roughplastic BSDFs are walked by the forward integrators only (volpath: general BSDF "
                                 "interface of dwalk.cuh; volpathaos: the interface kernel of docean.cuh with refractive next-event estimation); "
                                 "prbvolpath and prb_volpathaos refuse them");
    if (d.integrator.type == MIB_INTEG_PRBVOLPATHAOS && aos_interface_state_needed(s))
        throw std::runtime_error("prb_volpathaos: scenes with a non-null BSDF on anything but a horizontal rectangle need the interface state of "
                                 "volpathaos (a bounce that keeps the sign of d.z counts as a refraction, volpathaos.cpp:262-285); the adjoint "
                                 "kernels do not carry it");
    for (int i = 0; i < MIB_MAX_BUFFERS; ++i) { s->ds.buffers[i] = s->buf[i]; s->ds.grad[i] = s->grad[i]; }
    for (int i = 0; i < d.n_media; ++i) {
        mib_medium &m = d.media[i];
        mib_volume *vols[2] = { &m.sigma_t, &m.albedo };
        for (mib_volume *v : vols) {
            if (v->type == MIB_VOLUME_GRID) {
                size_t n = (size_t) v->res[0] * v->res[1] * v->res[2] * v->channels;
                if (v->buffer < 0 || v->buffer >= MIB_MAX_BUFFERS || !s->buf[v->buffer] || s->buf_n[v->buffer] < n)
                    throw std::runtime_error("gridvolume: data buffer missing or too small");
                if (!v->fixed_max) v->max_value = device_max(s, s->buf[v->buffer], n);      // grid.cpp:296-297: a fixed max_value is kept
            } else {
                v->max_value = v->value[0];
                for (int c = 1; c < v->channels; ++c) v->max_value = fmaxf(v->max_value, v->value[c]);
            }
        }
        m.max_density = m.scale * m.sigma_t.max_value;           // heterogeneous.cpp:163,176-178
        // fast mode: supergrid of local majorants (rebuilt at every commit, i.e. after every parameter update)
        s->ds.super[i] = nullptr;
        const int M = d.integrator.majorant_res;
        const mib_volume &v = m.sigma_t;
        const bool affine = v.to_local[12] == 0.f && v.to_local[13] == 0.f && v.to_local[14] == 0.f && v.to_local[15] == 1.f;
        if (M > 0 && m.type != MIB_MEDIUM_HOMOGENEOUS && v.type == MIB_VOLUME_GRID && v.wrap == MIB_WRAP_CLAMP && affine) {
            if (s->super_n[i] != M * M * M) {
                if (s->super[i]) CK(cudaFree(s->super[i]));
                CK(cudaMalloc(&s->super[i], (size_t) M * M * M * sizeof(float)));
                s->super_n[i] = M * M * M;
            }
            // prb integrators: 25 % head room, so that the null-collision derivative -1 / (1 - sigma_t / majorant) stays bounded
            // (a tight local majorant would make it heavy-tailed wherever the density is locally flat)
            const bool prb = d.integrator.type == MIB_INTEG_PRBVOLPATH || d.integrator.type == MIB_INTEG_PRBVOLPATHAOS;
            k_supergrid<<<M * M * M, 128, 0, s->stream>>>(s->buf[v.buffer], v.res[0], v.res[1], v.res[2], v.channels, M,
                                                         m.scale * (prb ? 1.25f : 1.f), s->super[i]);
            CK(cudaGetLastError());
            s->ds.super[i] = s->super[i];
        }
    }
    s->ds.super_res = d.integrator.majorant_res;
    s->ds.has_area_emitter = 0;
    for (int i = 0; i < d.n_emitters; ++i) if (d.emitters[i].type == MIB_EMITTER_AREA) s->ds.has_area_emitter = 1;
    for (int i = 0; i < d.n_phases; ++i) {
        const mib_phase &p = d.phases[i];
        if (p.type == MIB_PHASE_TABPHASE || p.type == MIB_PHASE_TABPHASE_POLARIZED) {
            build_table(s, i, 0, p.buf_nodes, p.buf_m11, p.n_nodes, true);
            if (p.type == MIB_PHASE_TABPHASE_POLARIZED) {
                build_table(s, i, 1, p.buf_nodes, p.buf_m12, p.n_nodes, false);
                build_table(s, i, 2, p.buf_nodes, p.buf_m33, p.n_nodes, false);
                build_table(s, i, 3, p.buf_nodes, p.buf_m34, p.n_nodes, false);
            }
        }
    }
    for (int i = 0; i < d.n_shapes; ++i) {
        if (d.shapes[i].type != MIB_SHAPE_CUBE) continue;
        const float *T = d.shapes[i].to_object;                   // normals: transpose(to_object) (transform.h:146-155)
        for (int f = 0; f < 6; ++f) {
            float nl[3] = { 0, 0, 0 };
            nl[f / 2] = (f & 1) ? 1.f : -1.f;
            float nw[3];
            for (int r = 0; r < 3; ++r)
                nw[r] = fmaf(T[8 + r], nl[2], fmaf(T[4 + r], nl[1], T[r] * nl[0]));
            float sq = fmaf(nw[2], nw[2], fmaf(nw[1], nw[1], nw[0] * nw[0]));
            float inv = sqrtf(1.f / sq);
            for (int r = 0; r < 3; ++r) s->ds.cube_normals[i][f][r] = nw[r] * inv;
        }
    }
    // prepare_scene (prbvolpath.py:76-89): which special cases the prb walks need
    s->ds.prb_null = 0; s->ds.prb_homog = 0;
    for (int i = 0; i < d.n_shapes; ++i)
        for (int m : { d.shapes[i].interior_medium, d.shapes[i].exterior_medium })
            if (m >= 0) {
                bool hom = d.media[m].type == MIB_MEDIUM_HOMOGENEOUS;
                s->ds.prb_homog |= hom ? 1 : 0;
                s->ds.prb_null |= hom ? 0 : 1;
            }
    s->ds.grad_scalar = s->grad_scalar;
    {   // reconstruction filter constants (gaussian.cpp:50-84: Remez fit to exp(-x/2), scaled, zero at the radius)
        const mib_sensor &sn = d.sensor;
        for (int i = 0; i < 10; ++i) s->ds.rf_coeff[i] = 0.f;
        s->ds.rf_inv_radius = 1.f;
        if (sn.rfilter == MIB_RFILTER_GAUSSIAN) {
            const double coeff[10] = { 9.992604880e-1, -4.977025247e-1, 1.222248550e-1, -1.932406282e-2, 2.136713061e-3,
                                       -1.679873860e-4, 9.202145248e-6, -3.329417433e-7, 7.128382794e-9, -6.821193280e-11 };
            double scale = 1;
            float *c = s->ds.rf_coeff;
            for (int i = 0; i < 10; ++i) { c[i] = (float) (coeff[i] * scale); scale /= (double) sn.rfilter_stddev * (double) sn.rfilter_stddev; }
            float x1 = sn.rfilter_radius * sn.rfilter_radius, x2 = x1 * x1, x4 = x2 * x2, x8 = x4 * x4;
            float a0 = fmaf(x1, c[1], c[0]), a1 = fmaf(x1, c[3], c[2]), a2 = fmaf(x1, c[5], c[4]), a3 = fmaf(x1, c[7], c[6]), a4 = fmaf(x1, c[9], c[8]);
            float b0 = fmaf(x2, a1, a0), b1 = fmaf(x2, a3, a2);
            c[0] -= fmaf(x8, a4, fmaf(x4, b1, b0));
        } else if (sn.rfilter == MIB_RFILTER_TENT) {
            s->ds.rf_inv_radius = 1.f / sn.rfilter_radius;
        }
    }
    // planar shading frames, computed by the device (identical floats to the per-hit evaluation they replace)
    {
        std::lock_guard<std::mutex> launch_lock(g_launch_mutex);
        s->ds.frames_ready = 0;
        float *dframes = nullptr;
        const size_t nb = sizeof(s->ds.face_frame), np = sizeof(s->ds.phase_const);
        CK(cudaMalloc(&dframes, nb + np));
        CK(cudaMemcpyToSymbolAsync(sc, &s->ds, sizeof(DScene), 0, cudaMemcpyHostToDevice, s->stream));
        k_shape_frames<<<MIB_MAX_SHAPES, 8, 0, s->stream>>>(dframes, dframes + nb / sizeof(float));
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(s->ds.face_frame, dframes, nb, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaMemcpyAsync(s->ds.phase_const, dframes + nb / sizeof(float), np, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        CK(cudaFree(dframes));
        s->ds.frames_ready = 1;
    }
    s->ds.weight_film = nullptr;
    ensure_film(s);
    s->committed = true;
}

// ---------------------------------------------------------------------------
// Launch helpers
// ---------------------------------------------------------------------------
static uint32_t pick_spt(uint64_t lanes, uint32_t spp) {
    // Aim for >= 8 waves of 148 SMs x 16 resident CTAs x 128 threads while keeping all
    // lanes of a thread inside one pixel whenever spp allows (one film flush per thread).
    const uint64_t target_threads = 148ull * 16 * 128 * 8;
    uint64_t spt = lanes / target_threads;
    if (spt < 1) spt = 1;
    if (spt > 256) spt = 256;
    if (spp >= spt) { while (spp % spt) --spt; }   // largest divisor of spp not above the target
    return (uint32_t) spt;
}

// The walk kernels are instantiated one per translation unit (inst.cu compiled with -DINST_*), so that the
// library builds in parallel; this file only sees their launchers.
namespace mib {
template <int SPEC, int MODE> void launch_k_forward(const DScene &sc, const LaunchParams &lp, unsigned blocks, cudaStream_t st);
template <int SPEC, int MODE, int V> void launch_k_forward_pool(const DScene &sc, const LaunchParams &lp, unsigned blocks, unsigned threads, cudaStream_t st);
template <int SPEC, int MODE> int pool_slots();
template <int SPEC> void launch_k_adjoint(const DScene &sc, const LaunchParams &lp, const AdjParams &ap, unsigned blocks, cudaStream_t st);
template <int SPEC, int V> void launch_k_adjoint_pool(const DScene &sc, const LaunchParams &lp, const AdjParams &ap, unsigned blocks, unsigned threads, cudaStream_t st);
template <int SPEC> int adjoint_pool_slots();
}

// ---------------------------------------------------------------------------
// Scene-specialised instantiations of the pool kernels (dmath.cuh MIB_FEAT; csrc/Makefile VARIANTS): the feature bits a scene
// needs, and the most specific compiled set that covers them.  MIB_VARIANT=<id> forces one (A/B runs; 0 = every feature).
// ---------------------------------------------------------------------------
static const unsigned kVariantMask[] = { 0xffffffffu, 417u, 146u, 150u };      // id -> MIB_FEAT of the object files
static const int kVariantOrder[] = { 1, 2, 3, 0 };                              // most specific first
static unsigned scene_features(const mib_scene *s) {
    const mib_scene_desc &d = s->ds.d;
    unsigned f = 0;
    for (int i = 0; i < d.n_media; ++i) {
        const mib_medium &m = d.media[i];
        f |= m.type == MIB_MEDIUM_HOMOGENEOUS ? FT_HOMOG : FT_GRID;
        if (m.sigma_t.type == MIB_VOLUME_GRID || m.albedo.type == MIB_VOLUME_GRID) f |= FT_GRID;
        if (m.has_spectral_extinction) f |= FT_SPECEXT;
    }
    if (d.integrator.majorant_res > 0) f |= FT_FAST;
    for (int i = 0; i < d.n_emitters; ++i) if (d.emitters[i].type == MIB_EMITTER_AREA) f |= FT_AREA;
    for (int i = 0; i < d.n_phases; ++i)
        f |= (d.phases[i].type == MIB_PHASE_TABPHASE || d.phases[i].type == MIB_PHASE_TABPHASE_POLARIZED) ? FT_TAB : FT_ANALYTIC;
    for (int i = 0; i < d.n_shapes; ++i)
        f |= d.shapes[i].type == MIB_SHAPE_RECTANGLE ? FT_RECT : (d.shapes[i].type == MIB_SHAPE_CUBE ? FT_CUBE : FT_SPHERE);
    if (d.sensor.rfilter != MIB_RFILTER_BOX) f |= FT_FILTER;
    return f;
}
static int pick_variant(const mib_scene *s) {
    const char *v = getenv("MIB_VARIANT");
    const unsigned need = scene_features(s);
    if (v && *v) {
        const int id = atoi(v);
        if (id < 0 || id >= (int) (sizeof(kVariantMask) / sizeof(kVariantMask[0])))
            throw std::runtime_error("MIB_VARIANT: no such instantiation");
        if ((kVariantMask[id] & need) != need)
            throw std::runtime_error("MIB_VARIANT: the scene needs features this instantiation was compiled without");
        return id;
    }
    for (int id : kVariantOrder) if ((kVariantMask[id] & need) == need) return id;
    return 0;
}

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// Forward kernel choice: the persistent path-pool kernel (dpool.cuh) for volpath / volpathaos; MIB_FORWARD_KERNEL=thread
// selects the one-thread-one-path kernel (A/B runs, profiles/).  Pool geometry: one CTA per SM, MIB_POOL_NT threads,
// as many slots as fit in the 227 KB of shared memory (compile-time, PoolL<SPEC>::P), fewer CTAs for tiny wavefronts.
template <int SPEC, int MODE>
static void launch_forward_pool(mib_scene *s, LaunchParams lp) {
    uint64_t lanes = lp.lane_end - lp.lane_begin;
    if (lanes == 0) return;
    int n_sm = 0;
    CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, s->device));
    int threads = env_int("MIB_POOL_NT", MIB_POOL_NT);
    threads = std::max(32, std::min(threads, MIB_POOL_NT)) & ~31;
    const int slots = pool_slots<SPEC, MODE>();
    uint64_t blocks = (lanes + 2ull * slots - 1) / (2ull * slots);
    blocks = std::max<uint64_t>(1, std::min<uint64_t>(blocks, (uint64_t) n_sm * env_int("MIB_POOL_CTAS_PER_SM", 1)));
    switch (pick_variant(s)) {
        case 1: launch_k_forward_pool<SPEC, MODE, 1>(s->ds, lp, (unsigned) blocks, (unsigned) threads, s->stream); break;
        case 2: launch_k_forward_pool<SPEC, MODE, 2>(s->ds, lp, (unsigned) blocks, (unsigned) threads, s->stream); break;
        case 3: launch_k_forward_pool<SPEC, MODE, 3>(s->ds, lp, (unsigned) blocks, (unsigned) threads, s->stream); break;
        default: launch_k_forward_pool<SPEC, MODE, 0>(s->ds, lp, (unsigned) blocks, (unsigned) threads, s->stream); break;
    }
    CK(cudaGetLastError());
}

// Forward kernel family for a scene (volpath / volpathaos): the path-pool kernel; MIB_FORWARD_KERNEL = thread | pool overrides
// (A/B runs, tests).  Round 1 kept the one-thread-one-path kernel for homogeneous scenes (C2 433 vs 178 Msamples/s); with
// bank-owned slots the pool kernel is ahead there too (profiles/r02_ab_pool.md).
static bool forward_uses_pool(const mib_scene *s) {
    const int integ = s->ds.d.integrator.type;
    if (integ != MIB_INTEG_VOLPATH && integ != MIB_INTEG_VOLPATHAOS) return false;     // prb primal passes: per-thread kernel
    if (integ == MIB_INTEG_VOLPATHAOS && aos_interface_state_needed(s)) return false;  // docean.cuh: per-thread kernel
    if (has_mesh(s)) return false;                                                     // dmesh.cuh: per-thread kernel
    if (has_dielectric(s)) return false;                                               // dbsdf.cuh: per-thread kernel
    const char *k = getenv("MIB_FORWARD_KERNEL");
    bool pool = true;
    if (k && strcmp(k, "thread") == 0) pool = false;
    if (k && strcmp(k, "pool") == 0) pool = true;
    return pool;
}

template <int SPEC> static void launch_adjoint(mib_scene *s, const LaunchParams &lp, const AdjParams &ap);   // mib_api_adjoint.inc

// Primal rendering with a prb integrator (mi.render / render_1 before backward(), the per-lane dump of the parity tests): the
// adjoint pool kernel in its primal-only mode (apool.cuh) -- the same stages the backward launch replays, about twice as fast as
// the one-thread-one-path walk.  MIB_PRB_PRIMAL_KERNEL=thread (or MIB_ADJOINT_KERNEL=thread) keeps the per-thread kernel.
static bool prb_primal_uses_pool() {
    const char *k = getenv("MIB_PRB_PRIMAL_KERNEL"), *a = getenv("MIB_ADJOINT_KERNEL");
    return !(k && strcmp(k, "thread") == 0) && !(a && strcmp(a, "thread") == 0);
}

template <int SPEC, int MODE>
static void launch_forward(mib_scene *s, const LaunchParams &lp) {
    if constexpr (MODE <= 1) {
        if (forward_uses_pool(s)) { launch_forward_pool<SPEC, MODE>(s, lp); return; }
    }
    if constexpr (MODE == 2 || MODE == 3) {
        if (prb_primal_uses_pool()) {
            AdjParams ap;
            ap.grad_image = nullptr;
            ap.render_1 = 0;
            ap.primal_only = 1;
            for (int i = 0; i < 4; ++i) ap.grad1[i] = 0.f;
            launch_adjoint<SPEC>(s, lp, ap);
            return;
        }
    }
    uint64_t lanes = lp.lane_end - lp.lane_begin;
    uint64_t threads = (lanes + lp.spt - 1) / lp.spt;
    uint64_t blocks = (threads + 127) / 128;
    if (blocks == 0) return;
    if (blocks > 0x7fffffffull) throw std::runtime_error("render: too many blocks");
    launch_k_forward<SPEC, MODE>(s->ds, lp, (unsigned) blocks, s->stream);
    CK(cudaGetLastError());
}

static void dispatch_forward(mib_scene *s, const LaunchParams &lp) {
    int spec = s->ds.d.spectrum, integ = s->ds.d.integrator.type;
    if (spec < 0 || spec > 2) throw std::runtime_error("unknown spectrum kind");
    switch (integ) {
        case MIB_INTEG_VOLPATH:
            if (spec == 0) launch_forward<0, 0>(s, lp); else if (spec == 1) launch_forward<1, 0>(s, lp); else launch_forward<2, 0>(s, lp);
            break;
        case MIB_INTEG_VOLPATHAOS:
            if (aos_interface_state_needed(s)) {
                if (spec == 0) launch_forward<0, 4>(s, lp); else if (spec == 1) launch_forward<1, 4>(s, lp); else launch_forward<2, 4>(s, lp);
                break;
            }
            if (spec == 0) launch_forward<0, 1>(s, lp); else if (spec == 1) launch_forward<1, 1>(s, lp); else launch_forward<2, 1>(s, lp);
            break;
        case MIB_INTEG_PRBVOLPATH:
            if (spec == 2) throw std::runtime_error("prbvolpath does not support polarized variants; use prb_volpathaos");
            if (spec == 0) launch_forward<0, 2>(s, lp); else launch_forward<1, 2>(s, lp);
            break;
        case MIB_INTEG_PRBVOLPATHAOS:
            if (spec != 2) throw std::runtime_error("prb_volpathaos requires a monochromatic polarized variant");
            launch_forward<2, 3>(s, lp);
            break;
        case MIB_INTEG_VOLPATHMIS:
            if (spec == 2) throw std::runtime_error("This integrator currently does not support polarized mode!");   // volpathmis.cpp:134-136
            if (spec == 0) launch_forward<0, 5>(s, lp); else launch_forward<1, 5>(s, lp);
            break;
        default: throw std::runtime_error("unknown integrator type");
    }
}

static void lane_range(mib_scene *s, uint32_t spp, uint64_t &b, uint64_t &e) {
    const mib_sensor &sn = s->ds.d.sensor;
    uint64_t total = (uint64_t) sn.crop_width * sn.crop_height * spp;
    // integrator.cpp:566-578: the reference's wavefront is limited to 2^32 lanes (32-bit lane index in the seed)
    if (total > 0xffffffffull)
        throw std::runtime_error("Tried to perform a rendering with a total sample count that exceeds 2^32 "
                                 "(the upper limit for one pass). Please use fewer samples per pixel or render using multiple passes.");
    if (s->stripe_world > 1) {
        // chunks are dealt in groups of `world`; in group g this rank owns chunk g * world + ((rank - g) mod world)
        // (kernels.cuh stripe_map).  Only the last chunk of the wavefront can be partial and only the last group incomplete,
        // so the rank's lanes form the contiguous VIRTUAL range [0, n_local).
        const uint64_t ch = s->stripe_chunk, chunks = (total + ch - 1) / ch, w = s->stripe_world;
        const uint64_t full_groups = chunks / w, rem = chunks % w;
        auto my_chunk = [&](uint64_t g) { return g * w + ((s->stripe_rank + w - g % w) % w); };
        const uint64_t owned = full_groups + ((rem && my_chunk(full_groups) < chunks) ? 1 : 0);
        uint64_t n_local = owned * ch;
        if (owned && my_chunk(owned - 1) == chunks - 1) n_local -= ch - (total - (chunks - 1) * ch);   // the wavefront's last chunk
        b = 0; e = n_local;
        return;
    }
    if (s->lane_begin == 0 && s->lane_end == 0) { b = 0; e = total; }
    else { b = std::min(s->lane_begin, total); e = std::min(s->lane_end, total); }
}

static void fetch_counters(mib_scene *s, float ms, uint64_t launches) {
    unsigned long long c[24];
    CK(cudaMemcpy(c, s->counters, sizeof(c), cudaMemcpyDeviceToHost));
#ifdef MIB_POOL_STATS
    fprintf(stderr, "pool batches per class (batches, mean lanes):");
    for (int k = 0; k < 8; ++k) fprintf(stderr, " %d: %llu, %.1f |", k, c[8 + 2 * k], c[8 + 2 * k] ? (double) c[9 + 2 * k] / (double) c[8 + 2 * k] : 0.0);
    fprintf(stderr, "\n");
#endif
    s->last.n_lookups = c[0]; s->last.n_real = c[1]; s->last.n_null = c[2];
    s->last.n_nee_steps = c[3]; s->last.n_surface = c[4]; s->last.n_samples = c[5];
    s->last.kernel_ms = ms; s->last.n_launches = launches;
    if (c[7]) throw std::runtime_error("path-pool kernel: queue watchdog fired (code " + std::to_string(c[7]) + ": 1 = no slot became available within the poll limit; ring build only: 2 = ring entry never arrived, 3 = ring cell reused too early); results are invalid");
}

// mode 0: film, mode 1: sums, mode 2: per-sample dump into per_sample_dev
static void run_forward(mib_scene *s, uint32_t seed, uint32_t spp, int mode, float *per_sample_dev) {
    if (!s->committed) throw std::runtime_error("mib_scene_commit() must be called before rendering");
    if (spp == 0) throw std::runtime_error("spp must be > 0");
    std::lock_guard<std::mutex> launch_lock(g_launch_mutex);
    if (s->ds.d.sensor.rfilter != MIB_RFILTER_BOX && s->ds.d.spectrum == MIB_SPEC_MONO_POLARIZED && mode == 0)
        throw std::runtime_error("the polarized Stokes film supports only the box filter");
    CK(cudaSetDevice(s->device));
    LaunchParams lp;
    lp.seed_value = s->ds.d.sampler_base_seed + seed;
    lp.spp = spp;
    lane_range(s, spp, lp.lane_begin, lp.lane_end);
    lp.stripe_rank = s->stripe_rank; lp.stripe_world = s->stripe_world; lp.stripe_chunk = s->stripe_chunk;
    lp.spt = pick_spt(lp.lane_end - lp.lane_begin, spp);
    if (s->stripe_world > 1) while (s->stripe_chunk % lp.spt) --lp.spt;       // a thread's lanes stay inside one chunk
    lp.mode = mode;
    lp.film = s->film_raw;
    lp.sums = s->sums;
    lp.per_sample = per_sample_dev;
    lp.counters = s->counters;
    CK(cudaMemsetAsync(s->counters, 0, 24 * sizeof(unsigned long long), s->stream));
    if (mode == 0) CK(cudaMemsetAsync(s->film_raw, 0, s->film_raw_n * sizeof(float), s->stream));
    if (mode == 1) CK(cudaMemsetAsync(s->sums, 0, 8 * sizeof(double), s->stream));
    CK(cudaEventRecord(s->ev0, s->stream));
    dispatch_forward(s, lp);
    CK(cudaEventRecord(s->ev1, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
    fetch_counters(s, ms, 1);
    s->last_spp = spp;
}

static void develop(mib_scene *s) {
    const mib_sensor &sn = s->ds.d.sensor;
    size_t npix = (size_t) sn.crop_width * sn.crop_height;
    unsigned blocks = (unsigned) ((npix + 255) / 256);
    switch (s->ds.d.spectrum) {
        case MIB_SPEC_MONO: k_develop<0><<<blocks, 256, 0, s->stream>>>(s->film_raw, s->film_dev, npix, s->last_spp); break;
        case MIB_SPEC_RGB: k_develop<1><<<blocks, 256, 0, s->stream>>>(s->film_raw, s->film_dev, npix, s->last_spp); break;
        default: k_develop<2><<<blocks, 256, 0, s->stream>>>(s->film_raw, s->film_dev, npix, s->last_spp); break;
    }
    CK(cudaGetLastError());
    s->last.n_launches += 1;
}

// ---------------------------------------------------------------------------
// extern "C" entry points (include/mi_b200.h)
// ---------------------------------------------------------------------------
#define MIB_TRY try {
#define MIB_CATCH                                                       \
    } catch (const std::exception &e) { g_err = e.what(); return 1; }   \
    return 0;

extern "C" {

const char *mib_last_error(void) { return g_err.c_str(); }
int mib_abi_version(void) { return MIB_ABI_VERSION; }
size_t mib_sizeof_scene_desc(void) { return sizeof(mib_scene_desc); }
size_t mib_sizeof_device_scene(void) { return sizeof(DScene); }

int mib_scene_create(const mib_scene_desc *desc, int device, mib_scene **out) {
    MIB_TRY
    if (!desc || !out) throw std::runtime_error("mib_scene_create: null argument");
    if (desc->abi_version != MIB_ABI_VERSION) throw std::runtime_error("mib_scene_create: ABI version mismatch");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        throw std::runtime_error("mib_scene_create: no CUDA device available (this library has no CPU fallback)");
    if (device < 0 || device >= ndev) throw std::runtime_error("mib_scene_create: invalid device ordinal");
    CK(cudaSetDevice(device));
    mib_scene *s = new mib_scene();
    s->device = device;
    memset(&s->ds, 0, sizeof(s->ds));
    s->ds.d = *desc;
    CK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&s->ev0));
    CK(cudaEventCreate(&s->ev1));
    CK(cudaMalloc(&s->sums, 8 * sizeof(double)));
    CK(cudaMalloc(&s->counters, 24 * sizeof(unsigned long long)));      // 8 counters + 16 words of pool statistics (MIB_POOL_STATS builds)
    CK(cudaMalloc(&s->maxbits, sizeof(unsigned int)));
    CK(cudaMalloc(&s->grad_scalar, MIB_MAX_MEDIA * 4 * sizeof(double)));
    CK(cudaMemset(s->grad_scalar, 0, MIB_MAX_MEDIA * 4 * sizeof(double)));
    *out = s;
    MIB_CATCH
}

int mib_scene_destroy(mib_scene *s) {
    MIB_TRY
    if (!s) return 0;
    cudaSetDevice(s->device);
    for (int i = 0; i < MIB_MAX_BUFFERS; ++i) { if (s->buf[i]) cudaFree(s->buf[i]); if (s->grad[i]) cudaFree(s->grad[i]); }
    for (int i = 0; i < MIB_MAX_PHASES; ++i) if (s->cdf[i]) cudaFree(s->cdf[i]);
    for (int i = 0; i < MIB_MAX_MEDIA; ++i) if (s->super[i]) cudaFree(s->super[i]);
    if (s->film_raw) cudaFree(s->film_raw);
    if (s->film_dev) cudaFree(s->film_dev);
    if (s->lane_L) cudaFree(s->lane_L);
    cudaFree(s->sums); cudaFree(s->counters); cudaFree(s->maxbits); cudaFree(s->grad_scalar);
    cudaEventDestroy(s->ev0); cudaEventDestroy(s->ev1);
    cudaStreamDestroy(s->stream);
    delete s;
    MIB_CATCH
}

int mib_scene_set_buffer(mib_scene *s, int idx, const float *host, size_t count) {
    MIB_TRY
    if (idx < 0 || idx >= MIB_MAX_BUFFERS) throw std::runtime_error("mib_scene_set_buffer: slot out of range");
    if (!host || count == 0) throw std::runtime_error("mib_scene_set_buffer: empty buffer");
    CK(cudaSetDevice(s->device));
    if (s->buf_n[idx] != count) {
        if (s->buf[idx]) CK(cudaFree(s->buf[idx]));
        if (s->grad[idx]) { CK(cudaFree(s->grad[idx])); s->grad[idx] = nullptr; }
        CK(cudaMalloc(&s->buf[idx], count * sizeof(float)));
        s->buf_n[idx] = count;
    }
    CK(cudaMemcpyAsync(s->buf[idx], host, count * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    s->committed = false;
    MIB_CATCH
}

int mib_scene_get_buffer(mib_scene *s, int idx, float *host, size_t count) {
    MIB_TRY
    if (idx < 0 || idx >= MIB_MAX_BUFFERS || !s->buf[idx] || count > s->buf_n[idx]) throw std::runtime_error("mib_scene_get_buffer: bad slot/size");
    CK(cudaMemcpy(host, s->buf[idx], count * sizeof(float), cudaMemcpyDeviceToHost));
    MIB_CATCH
}

int mib_scene_buffer_dev(mib_scene *s, int idx, float **dev, size_t *count) {
    MIB_TRY
    if (idx < 0 || idx >= MIB_MAX_BUFFERS || !s->buf[idx]) throw std::runtime_error("mib_scene_buffer_dev: bad slot");
    *dev = s->buf[idx]; *count = s->buf_n[idx];
    s->committed = false;   // the caller may write through the pointer: derived data must be recomputed
    MIB_CATCH
}

int mib_scene_set_desc(mib_scene *s, const mib_scene_desc *desc) {
    MIB_TRY
    if (desc->abi_version != MIB_ABI_VERSION) throw std::runtime_error("mib_scene_set_desc: ABI version mismatch");
    s->ds.d = *desc;
    s->committed = false;
    MIB_CATCH
}

int mib_scene_commit(mib_scene *s) {
    MIB_TRY
    CK(cudaSetDevice(s->device));
    commit(s);
    MIB_CATCH
}

int mib_scene_get_desc(mib_scene *s, mib_scene_desc *out) { *out = s->ds.d; return 0; }

int mib_scene_set_lane_range(mib_scene *s, uint64_t b, uint64_t e) {
    MIB_TRY
    if (e < b) throw std::runtime_error("mib_scene_set_lane_range: end < begin");
    s->lane_begin = b; s->lane_end = e;
    MIB_CATCH
}

int mib_scene_set_lane_stripes(mib_scene *s, uint32_t rank, uint32_t world, uint64_t chunk_lanes) {
    MIB_TRY
    if (world <= 1) { s->stripe_rank = 0; s->stripe_world = 1; s->stripe_chunk = 0; return 0; }
    if (rank >= world) throw std::runtime_error("mib_scene_set_lane_stripes: rank >= world");
    if (chunk_lanes == 0) throw std::runtime_error("mib_scene_set_lane_stripes: chunk must be > 0");
    s->stripe_rank = rank; s->stripe_world = world; s->stripe_chunk = chunk_lanes;
    s->lane_begin = s->lane_end = 0;
    MIB_CATCH
}

int mib_render(mib_scene *s, uint32_t seed, uint32_t spp, int do_develop, float *out_host) {
    MIB_TRY
    run_forward(s, seed, spp, 0, nullptr);
    if (do_develop) develop(s);
    if (out_host) {
        if (do_develop) CK(cudaMemcpyAsync(out_host, s->film_dev, s->film_dev_n * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
        else CK(cudaMemcpyAsync(out_host, s->film_raw, s->film_raw_n * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    }
    CK(cudaStreamSynchronize(s->stream));
    MIB_CATCH
}

int mib_film_dev(mib_scene *s, float **raw_dev, size_t *raw_count, float **developed_dev, size_t *developed_count) {
    MIB_TRY
    if (!s->committed) throw std::runtime_error("mib_film_dev: scene not committed");
    if (raw_dev) *raw_dev = s->film_raw;
    if (raw_count) *raw_count = s->film_raw_n;
    if (developed_dev) *developed_dev = s->film_dev;
    if (developed_count) *developed_count = s->film_dev_n;
    MIB_CATCH
}

int mib_film_develop(mib_scene *s) {
    MIB_TRY
    CK(cudaSetDevice(s->device));
    develop(s);
    CK(cudaStreamSynchronize(s->stream));
    MIB_CATCH
}

int mib_film_read(mib_scene *s, int developed, float *host) {
    MIB_TRY
    if (!s->committed || !host) throw std::runtime_error("mib_film_read: scene not committed or null output");
    CK(cudaSetDevice(s->device));
    CK(cudaMemcpyAsync(host, developed ? s->film_dev : s->film_raw, (developed ? s->film_dev_n : s->film_raw_n) * sizeof(float),
                       cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    MIB_CATCH
}

static int render_1_impl(mib_scene *s, uint32_t seed, uint32_t spp, double *out4, bool normalize) {
    MIB_TRY
    if (s->ds.d.spectrum == MIB_SPEC_RGB)
        throw std::runtime_error("This render loop only supports monochromatic and spectral modes!");   // integrator.cpp:409-410
    run_forward(s, seed, spp, 1, nullptr);
    double h[8];
    CK(cudaMemcpy(h, s->sums, sizeof(h), cudaMemcpyDeviceToHost));
    int nch = s->ds.d.spectrum == MIB_SPEC_MONO_POLARIZED ? 4 : 1;
    const mib_sensor &sn = s->ds.d.sensor;
    // integrator.cpp:702-705: nf = rcp((float) n_pixels * total_samples_done)
    float nf = 1.f / ((float) ((size_t) sn.crop_width * sn.crop_height) * (float) spp);
    for (int c = 0; c < 4; ++c) out4[c] = c < nch ? (normalize ? h[c] * (double) nf : h[c]) : 0.0;
    MIB_CATCH
}
int mib_render_1(mib_scene *s, uint32_t seed, uint32_t spp, double *out4) { return render_1_impl(s, seed, spp, out4, true); }
int mib_render_1_raw(mib_scene *s, uint32_t seed, uint32_t spp, double *out4) { return render_1_impl(s, seed, spp, out4, false); }

int mib_counters_get(mib_scene *s, mib_counters *out) { *out = s->last; return 0; }
int mib_scene_forward_kernel(mib_scene *s, int *is_pool) { *is_pool = forward_uses_pool(s) ? 1 : 0; return 0; }
int mib_scene_stream(mib_scene *s, void **stream) { *stream = (void *) s->stream; return 0; }

// A per-lane request must name lanes of the wavefront: (0, 0) is the "whole wavefront" sentinel of lane_range() and a
// range past the end would be clamped silently, leaving rows of the caller's buffer unwritten.
static void check_lane_slice(const mib_scene *s, uint32_t spp, uint64_t lane_begin, uint64_t count, const char *who) {
    const mib_sensor &sn = s->ds.d.sensor;
    const uint64_t total = (uint64_t) sn.crop_width * sn.crop_height * spp;
    if (count == 0) throw std::runtime_error(std::string(who) + ": count must be > 0");
    if (lane_begin > total || count > total - lane_begin)
        throw std::runtime_error(std::string(who) + ": lanes [" + std::to_string(lane_begin) + ", " + std::to_string(lane_begin) + " + " +
                                 std::to_string(count) + ") exceed the wavefront of " + std::to_string(total) + " lanes (crop_width * crop_height * spp)");
}

// Debug / test hook: per-lane aovs (C floats per lane) for lanes [lane_begin, lane_begin+count)
int mib_debug_sample_lanes(mib_scene *s, uint32_t seed, uint32_t spp, uint64_t lane_begin, uint64_t count, float *out_host) {
    MIB_TRY
    if (!s || !out_host) throw std::runtime_error("mib_debug_sample_lanes: null argument");
    check_lane_slice(s, spp, lane_begin, count, "mib_debug_sample_lanes");
    CK(cudaSetDevice(s->device));
    int C = raw_channels(s->ds.d.spectrum);
    float *dev = nullptr;
    CK(cudaMalloc(&dev, count * C * sizeof(float)));
    CK(cudaMemsetAsync(dev, 0, count * C * sizeof(float), s->stream));
    uint64_t sb = s->lane_begin, se = s->lane_end;
    const uint32_t sw = s->stripe_world;
    s->lane_begin = lane_begin; s->lane_end = lane_begin + count; s->stripe_world = 1;
    try { run_forward(s, seed, spp, 2, dev); } catch (...) { s->lane_begin = sb; s->lane_end = se; s->stripe_world = sw; cudaFree(dev); throw; }
    s->lane_begin = sb; s->lane_end = se; s->stripe_world = sw;
    CK(cudaMemcpy(out_host, dev, count * C * sizeof(float), cudaMemcpyDeviceToHost));
    CK(cudaFree(dev));
    MIB_CATCH
}

} // extern "C"

// ---------------------------------------------------------------------------
// Allocation-free per-lane sampling (no counterpart in the reference).
//
// mib_debug_sample_lanes() above cudaMalloc()s its per-lane scratch on every call, which tight serving loops and harnesses
// that forbid allocation cannot afford.  This variant keeps the scratch in a file-scope device array of
// MIB_NOALLOC_MAX_FLOATS floats (4 MiB, held by every GPU that loads the library until the process exits).  In host code the
// array's NAME is only a host-side stand-in, not a device address, so the address is resolved once per device with
// cudaGetSymbolAddress() (that device current) and that pointer goes to the kernel.  The call allocates nothing, but it is
// NOT capturable into a CUDA graph: like every render entry point it synchronises the scene's stream and reads the event
// counters back.  All calls of a process share the array and it keeps the last call's lanes until the next call: the part
// a call uses is cleared first, and g_noalloc_mutex serialises the calls (two calls overlapping on one GPU would race).
// ---------------------------------------------------------------------------
static __device__ __align__(256) float g_noalloc_lanes[MIB_NOALLOC_MAX_FLOATS];
static constexpr int kNoallocMaxDevices = 64;
static float *g_noalloc_ptr[kNoallocMaxDevices] = {};
static std::mutex g_noalloc_mutex;

// Device address of g_noalloc_lanes on `device`, which must be the current device.
static float *noalloc_scratch(int device) {
    if (device < 0 || device >= kNoallocMaxDevices) throw std::runtime_error("mib_sample_lanes_noalloc: device index out of range");
    if (!g_noalloc_ptr[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_lanes);
        if (e != cudaSuccess || !p)
            throw std::runtime_error(std::string("mib_sample_lanes_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
        g_noalloc_ptr[device] = static_cast<float *>(p);
    }
    return g_noalloc_ptr[device];
}

extern "C" int mib_sample_lanes_noalloc(mib_scene *s, uint32_t seed, uint32_t spp, uint64_t lane_begin, uint64_t count, float *out_host) {
    MIB_TRY
    if (!s || !out_host) throw std::runtime_error("mib_sample_lanes_noalloc: null argument");
    const uint64_t C = (uint64_t) raw_channels(s->ds.d.spectrum);
    if (count > MIB_NOALLOC_MAX_FLOATS / C)
        throw std::runtime_error("mib_sample_lanes_noalloc: " + std::to_string(count) + " lanes x " + std::to_string(C) +
                                 " channels exceed the static scratch of " + std::to_string((unsigned long long) MIB_NOALLOC_MAX_FLOATS) +
                                 " floats (MIB_NOALLOC_MAX_FLOATS); split the lane range or use mib_debug_sample_lanes");
    check_lane_slice(s, spp, lane_begin, count, "mib_sample_lanes_noalloc");
    std::lock_guard<std::mutex> call_lock(g_noalloc_mutex);
    CK(cudaSetDevice(s->device));
    float *dev = noalloc_scratch(s->device);
    const size_t bytes = (size_t) (count * C) * sizeof(float);
    CK(cudaMemsetAsync(dev, 0, bytes, s->stream));               // the array keeps the previous call's lanes
    const uint64_t sb = s->lane_begin, se = s->lane_end;
    const uint32_t sw = s->stripe_world;
    s->lane_begin = lane_begin; s->lane_end = lane_begin + count; s->stripe_world = 1;
    try { run_forward(s, seed, spp, 2, dev); } catch (...) { s->lane_begin = sb; s->lane_end = se; s->stripe_world = sw; throw; }
    s->lane_begin = sb; s->lane_end = se; s->stripe_world = sw;
    CK(cudaMemcpyAsync(out_host, dev, bytes, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    MIB_CATCH
}

#include "mib_api_adjoint.inc"
#include "mib_api_debug.inc"

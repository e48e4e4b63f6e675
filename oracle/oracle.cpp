// oracle.cpp -- TEST INFRASTRUCTURE (parity oracle). Not part of the product.
//
// CPU restatement of the mitsuba3-nasa volumetric path tracing hot path with the
// lane -> pixel -> PCG32-seed mapping of the reference's llvm_ad_* variants:
//   SamplingIntegrator::render        src/render/integrator.cpp:276-390
//   SamplingIntegrator::render_1      src/render/integrator.cpp:553-743
//   SamplingIntegrator::render_sample src/render/integrator.cpp:1139-1240
//   to_sensor_mueller                 src/render/integrator.cpp:1245-1267
//   ImageBlock::put (box)             src/render/imageblock.cpp:206-232
//   HDRFilm::develop                  src/films/hdrfilm.cpp:304-418
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may load this library.  Parity status: building blocks are
// pinned bit-for-bit against the reference headers compiled in oracle/_ref
// (tests/test_oracle_vs_ref.py); the integrator level has no fixed-seed golden
// in the reference -> "parity unpinned" above the building blocks.
#include "owalk.h"
#include "oadjoint.h"
#include "omie.h"
#include "ofresnel.h"
#include "omicrofacet.h"
#include "oroughdielectric.h"
#include <thread>
#include <mutex>
#include <atomic>
#include <cstdio>
#include <string>

using namespace orc;

namespace {

struct OScene {
    Scene sc;
    std::vector<float> storage[MIB_MAX_BUFFERS];
    std::vector<double> grad[MIB_MAX_BUFFERS];
    double grad_scalar[MIB_MAX_MEDIA][4] = {};
    Counters counters;
    std::string error;
};

thread_local std::string g_error;

// to_sensor_mueller (integrator.cpp:1245-1267)
Mueller to_sensor_mueller(const Ray &ray, const Mueller &spec) {
    Vec3 md = -ray.d;
    Vec3 current_basis = stokes_basis(md);
    Vec3 vertical(0.f, 0.f, 1.f);
    Vec3 tmp = cross(md, vertical);
    bool ray_is_vertical = norm(tmp) < RayEpsilon;
    Vec3 target_basis = ray_is_vertical ? Vec3(1.f, 0.f, 0.f) : cross(md, normalize(tmp));
    Mueller spec2sensor = rotate_stokes_basis(md, current_basis, target_basis);
    return matmul(spec2sensor, spec);
}

struct LaneSetup { uint32_t px, py; float sample_x, sample_y; Ray ray; };

// integrator.cpp:323-340 + :1139-1170
// reseed = false: the sampler keeps its stream (scalar variants: one stream per pixel, integrator.cpp:1110-1123)
LaneSetup setup_lane(const Scene &sc, Sampler &sampler, uint32_t seed, uint64_t lane, uint32_t spp, bool reseed = true) {
    const mib_sensor &s = sc.d.sensor;
    if (reseed) sampler.seed_llvm(sc.d.sampler_base_seed + seed, (uint32_t) lane);
    uint32_t idx = (uint32_t) (lane / spp);
    LaneSetup L;
    L.py = idx / (uint32_t) s.crop_width;
    L.px = idx - (uint32_t) s.crop_width * L.py;
    float posx = (float) ((int) L.px + s.crop_offset_x), posy = (float) ((int) L.py + s.crop_offset_y);
    float scale_x = 1.f / (float) s.crop_width, scale_y = 1.f / (float) s.crop_height;
    float offset_x = -(float) s.crop_offset_x * scale_x, offset_y = -(float) s.crop_offset_y * scale_y;
    float jx, jy; sampler.next_2d(&jx, &jy);
    L.sample_x = posx + jx; L.sample_y = posy + jy;
    float ax = fmadd(L.sample_x, scale_x, offset_x), ay = fmadd(L.sample_y, scale_y, offset_y);
    float apx = .5f, apy = .5f;
    if (s.needs_aperture_sample) sampler.next_2d(&apx, &apy);
    L.ray = sensor_sample_ray(sc, ax, ay, apx, apy);
    return L;
}

template <int SPEC>
typename SpecTraits<SPEC>::Spec trace(const Scene &sc, Sampler &sampler, Counters &cnt, const Ray &ray, bool *valid) {
    Walk<SPEC> w(sc, sampler, cnt);
    cnt.n_samples++;
    int t = sc.d.integrator.type;
    if (t == MIB_INTEG_VOLPATH) return w.sample_volpath(ray, sc.d.sensor.medium, valid);
    if (t == MIB_INTEG_VOLPATHAOS) return w.sample_volpathaos(ray, sc.d.sensor.medium, valid);
    if (t == MIB_INTEG_VOLPATHMIS) return w.sample_volpathmis(ray, sc.d.sensor.medium, valid);
    Adjoint<SPEC> a(sc, sampler, cnt);
    if (t == MIB_INTEG_PRBVOLPATH) return a.primal_prbvolpath(ray, sc.d.sensor.medium, valid);
    return a.primal_prbvolpathaos(ray, sc.d.sensor.medium, valid);
}

// aovs of one sample: rgb -> r,g,b,1 ; mono -> I,1 ; polarized -> I,Q,U,V (after to_sensor_mueller)
template <int SPEC>
int sample_aovs(const Scene &sc, Sampler &sampler, Counters &cnt, const Ray &ray, float *aovs) {
    bool valid;
    auto spec = trace<SPEC>(sc, sampler, cnt, ray, &valid);
    if constexpr (SPEC == MIB_SPEC_RGB) { aovs[0] = spec.v[0]; aovs[1] = spec.v[1]; aovs[2] = spec.v[2]; aovs[3] = 1.f; return 4; }
    else if constexpr (SPEC == MIB_SPEC_MONO) { aovs[0] = spec.v[0]; aovs[1] = 1.f; return 2; }
    else {
        Mueller m = to_sensor_mueller(ray, spec);
        for (int i = 0; i < 4; ++i) aovs[i] = m.m[i][0];
        return 4;
    }
}

// ImageBlock::put for a reconstruction filter (imageblock.cpp:261-530, the JIT recorded-loop forms, border 0):
// calls fn(pixel, weight) for every pixel of the footprint.  Coalesced form when spp >= 4 (integrator.cpp:320,
// common.py:93), the general form otherwise.
template <typename Fn>
void filter_footprint(const Scene &sc, uint32_t spp, float pos_x, float pos_y, Fn &&fn) {
    const mib_sensor &s = sc.d.sensor;
    const float radius = s.rfilter_radius;
    const uint32_t size_x = (uint32_t) s.crop_width, size_y = (uint32_t) s.crop_height;
    if (spp >= 4) {
        uint32_t n = (uint32_t) (int) std::ceil(radius - .5f), count = 2 * n + 1;
        int pix = (int) std::floor(pos_x) - (int) n, piy = (int) std::floor(pos_y) - (int) n;
        uint32_t x = (uint32_t) (pix - s.crop_offset_x), y = (uint32_t) (piy - s.crop_offset_y);
        float rel_x = ((float) pix + .5f) - pos_x, rel_y = ((float) piy + .5f) - pos_y;
        for (uint32_t ys = 0; ys < count; ++ys) {
            float weight_y = sc.rfilter_eval(rel_y + (float) ys);
            bool active_1 = y + ys < size_y;
            for (uint32_t xs = 0; xs < count; ++xs) {
                float weight = sc.rfilter_eval(rel_x + (float) xs) * weight_y;
                if (active_1 && x + xs < size_x) fn((size_t) (y + ys) * size_x + (x + xs), weight);
            }
        }
    } else {
        float off_x = (float) (-s.crop_offset_x) - .5f, off_y = (float) (-s.crop_offset_y) - .5f;
        float pfx = pos_x + off_x, pfy = pos_y + off_y;
        float p0x = pfx - radius, p0y = pfy - radius, p1x = pfx + radius, p1y = pfy + radius;
        uint32_t x0 = (uint32_t) std::max((int) std::ceil(p0x), 0), y0 = (uint32_t) std::max((int) std::ceil(p0y), 0);
        uint32_t x1 = (uint32_t) std::min((int) std::floor(p1x), (int) size_x - 1), y1 = (uint32_t) std::min((int) std::floor(p1y), (int) size_y - 1);
        uint32_t count = (uint32_t) (int) std::ceil(2.f * radius);
        bool active = x0 <= x1 && y0 <= y1;
        float rel_x = (float) x0 - pfx, rel_y = (float) y0 - pfy;
        for (uint32_t ys = 0; ys < count; ++ys) {
            float weight_y = sc.rfilter_eval(rel_y + (float) ys);
            bool active_1 = active && (y0 + ys <= y1);
            for (uint32_t xs = 0; xs < count; ++xs) {
                float weight = sc.rfilter_eval(rel_x + (float) xs) * weight_y;
                if (active_1 && (x0 + xs <= x1)) fn((size_t) (y0 + ys) * size_x + (x0 + xs), weight);
            }
        }
    }
}

int n_raw_channels(int spec) { return spec == MIB_SPEC_MONO ? 2 : 4; }
int n_dev_channels(int spec) { return spec == MIB_SPEC_MONO ? 1 : (spec == MIB_SPEC_RGB ? 3 : 4); }

template <int SPEC>
void render_range(OScene *os, uint32_t seed, uint32_t spp, uint64_t lane_begin, uint64_t lane_end,
                  double *film, std::mutex *mtx, Counters *cnt_out, float *per_sample) {
    const Scene &sc = os->sc;
    const mib_sensor &s = sc.d.sensor;
    const int C = n_raw_channels(SPEC);
    Counters cnt;
    Sampler sampler;
    double acc[4] = { 0, 0, 0, 0 };
    int64_t cur_pixel = -1;
    auto flush = [&]() {
        if (cur_pixel >= 0) {
            std::lock_guard<std::mutex> g(*mtx);
            for (int c = 0; c < C; ++c) film[cur_pixel * C + c] += acc[c];
        }
        for (int c = 0; c < 4; ++c) acc[c] = 0;
    };
    for (uint64_t lane = lane_begin; lane < lane_end; ++lane) {
        LaneSetup L = setup_lane(sc, sampler, seed, lane, spp);
        float aovs[4];
        sample_aovs<SPEC>(sc, sampler, cnt, L.ray, aovs);
        if (per_sample) for (int c = 0; c < C; ++c) per_sample[(lane - lane_begin) * C + c] = aovs[c];
        if (s.rfilter != MIB_RFILTER_BOX && !per_sample) {
            std::lock_guard<std::mutex> g(*mtx);
            filter_footprint(sc, spp, L.sample_x, L.sample_y, [&](size_t pixel, float w) {
                for (int c = 0; c < C; ++c) film[pixel * C + c] += (double) (aovs[c] * w);
            });
            continue;
        }
        // box filter put (imageblock.cpp:206-232): p = floor2int(pos) - offset; active &= all(p < size)
        int ix = (int) std::floor(L.sample_x) - s.crop_offset_x, iy = (int) std::floor(L.sample_y) - s.crop_offset_y;
        if ((uint32_t) ix >= (uint32_t) s.crop_width || (uint32_t) iy >= (uint32_t) s.crop_height) continue;
        int64_t pixel = (int64_t) iy * s.crop_width + ix;
        if (pixel != cur_pixel) { flush(); cur_pixel = pixel; }
        for (int c = 0; c < C; ++c) acc[c] += (double) aovs[c];
    }
    flush();
    std::lock_guard<std::mutex> g(*mtx);
    cnt_out->add(cnt);
}

template <int SPEC>
void render_all(OScene *os, uint32_t seed, uint32_t spp, uint64_t lane_begin, uint64_t lane_end, int nthreads,
                double *film, float *per_sample) {
    if (nthreads < 1) nthreads = 1;
    uint64_t total = lane_end - lane_begin;
    if ((uint64_t) nthreads > total) nthreads = (int) std::max<uint64_t>(1, total);
    std::mutex mtx;
    Counters cnt;
    std::vector<std::thread> th;
    uint64_t chunk = (total + nthreads - 1) / nthreads;
    const int C = n_raw_channels(SPEC);
    for (int t = 0; t < nthreads; ++t) {
        uint64_t b = lane_begin + t * chunk, e = std::min(lane_end, b + chunk);
        if (b >= e) break;
        th.emplace_back([=, &mtx, &cnt]() {
            render_range<SPEC>(os, seed, spp, b, e, film, &mtx, &cnt, per_sample ? per_sample + (b - lane_begin) * C : nullptr);
        });
    }
    for (auto &x : th) x.join();
    os->counters = cnt;
}

void dispatch_render(OScene *os, uint32_t seed, uint32_t spp, uint64_t lb, uint64_t le, int nthreads, double *film, float *per_sample) {
    if (os->sc.d.sensor.rfilter != MIB_RFILTER_BOX && os->sc.d.spectrum == MIB_SPEC_MONO_POLARIZED && !per_sample)
        throw std::runtime_error("the polarized Stokes film supports only the box filter");
    switch (os->sc.d.spectrum) {
        case MIB_SPEC_MONO: render_all<0>(os, seed, spp, lb, le, nthreads, film, per_sample); break;
        case MIB_SPEC_RGB: render_all<1>(os, seed, spp, lb, le, nthreads, film, per_sample); break;
        default: render_all<2>(os, seed, spp, lb, le, nthreads, film, per_sample); break;
    }
}

} // namespace

extern "C" {

const char *orc_last_error() { return g_error.c_str(); }

int orc_scene_create(const mib_scene_desc *desc, void **out) {
    if (!desc || desc->abi_version != MIB_ABI_VERSION) { g_error = "orc_scene_create: ABI version mismatch"; return 1; }
    OScene *os = new OScene();
    os->sc.d = *desc;
    *out = os;
    return 0;
}
int orc_scene_destroy(void *h) { delete (OScene *) h; return 0; }
int orc_scene_set_desc(void *h, const mib_scene_desc *desc) { ((OScene *) h)->sc.d = *desc; return 0; }
int orc_scene_get_desc(void *h, mib_scene_desc *out) { *out = ((OScene *) h)->sc.d; return 0; }
int orc_scene_set_buffer(void *h, int idx, const float *data, size_t n) {
    OScene *os = (OScene *) h;
    if (idx < 0 || idx >= MIB_MAX_BUFFERS) { g_error = "buffer index out of range"; return 1; }
    os->storage[idx].assign(data, data + n);
    os->sc.buffers[idx] = os->storage[idx].data();
    os->sc.buffer_count[idx] = n;
    os->grad[idx].assign(n, 0.0);
    return 0;
}
int orc_scene_commit(void *h) {
    OScene *os = (OScene *) h;
    try { os->sc.prepare(); } catch (const std::exception &e) { g_error = e.what(); return 1; }
    return 0;
}

// raw: [H*W*C] float sums, developed: [H*W*Cout]; lane range (0,0) = everything
int orc_render(void *h, uint32_t seed, uint32_t spp, uint64_t lane_begin, uint64_t lane_end, int nthreads,
               float *raw, float *developed) {
    OScene *os = (OScene *) h;
    try {
        const mib_sensor &s = os->sc.d.sensor;
        size_t npix = (size_t) s.crop_width * s.crop_height;
        if (lane_begin == 0 && lane_end == 0) lane_end = (uint64_t) npix * spp;
        const int C = n_raw_channels(os->sc.d.spectrum), Co = n_dev_channels(os->sc.d.spectrum);
        std::vector<double> film(npix * C, 0.0);
        dispatch_render(os, seed, spp, lane_begin, lane_end, nthreads, film.data(), nullptr);
        if (raw) for (size_t i = 0; i < npix * C; ++i) raw[i] = (float) film[i];
        if (developed) {
            for (size_t p = 0; p < npix; ++p) {
                const double *f = &film[p * C];
                if (os->sc.d.spectrum == MIB_SPEC_RGB) {
                    float w = (float) f[3]; float dv = (w == 0.f) ? 1.f : w;
                    for (int c = 0; c < 3; ++c) developed[p * Co + c] = (float) f[c] / dv;
                } else if (os->sc.d.spectrum == MIB_SPEC_MONO) {
                    float I = (float) f[0], w = (float) f[1]; float dv = (w == 0.f) ? 1.f : w;
                    float lum = I * 0.212671f + I * 0.715160f + I * 0.072169f;   // hdrfilm.cpp:388-392 luminance(rgb)
                    developed[p] = lum / dv;
                } else {
                    // Stokes film (extension): sums / spp
                    for (int c = 0; c < 4; ++c) developed[p * Co + c] = (float) (f[c] / (double) spp);
                }
            }
        }
    } catch (const std::exception &e) { g_error = e.what(); return 1; }
    return 0;
}

// Per-sample aovs for lanes [lane_begin, lane_begin+count): out[count*C]
int orc_sample_lanes(void *h, uint32_t seed, uint32_t spp, uint64_t lane_begin, uint64_t count, int nthreads, float *out) {
    OScene *os = (OScene *) h;
    try {
        const mib_sensor &s = os->sc.d.sensor;
        size_t npix = (size_t) s.crop_width * s.crop_height;
        const int C = n_raw_channels(os->sc.d.spectrum);
        std::vector<double> film(npix * C, 0.0);
        dispatch_render(os, seed, spp, lane_begin, lane_begin + count, nthreads, film.data(), out);
    } catch (const std::exception &e) { g_error = e.what(); return 1; }
    return 0;
}

// Scalar-variant seeding (the scheme of the reference's scalar_* variants, which oracle/_ref/scalar can run):
// pixel k = pixel_index[k] (row-major) draws its spp samples from ONE PCG32 stream seeded with
// base_seed + pixel_seed[k] (sampler.cpp:131-133, integrator.cpp:1110-1123).  out[(k * spp + j) * C + c].
int orc_sample_pixels_scalar(void *h, const uint32_t *pixel_index, const uint32_t *pixel_seed, uint32_t n_pixels, uint32_t spp, float *out) {
    OScene *os = (OScene *) h;
    try {
        const Scene &sc = os->sc;
        const int spec = sc.d.spectrum, C = n_raw_channels(spec);
        Counters cnt;
        Sampler sampler;
        for (uint32_t k = 0; k < n_pixels; ++k) {
            sampler.seed_scalar(sc.d.sampler_base_seed + pixel_seed[k]);
            for (uint32_t j = 0; j < spp; ++j) {
                LaneSetup L = setup_lane(sc, sampler, 0, (uint64_t) pixel_index[k] * spp + j, spp, false);
                float aovs[4];
                if (spec == MIB_SPEC_MONO) sample_aovs<0>(sc, sampler, cnt, L.ray, aovs);
                else if (spec == MIB_SPEC_RGB) sample_aovs<1>(sc, sampler, cnt, L.ray, aovs);
                else sample_aovs<2>(sc, sampler, cnt, L.ray, aovs);
                for (int c = 0; c < C; ++c) out[((size_t) k * spp + j) * C + c] = aovs[c];
            }
        }
        os->counters = cnt;
    } catch (const std::exception &e) { g_error = e.what(); return 1; }
    return 0;
}

// integrator.cpp:665-705: sum all pixels in double, normalise by n_pixels*spp
int orc_render_1(void *h, uint32_t seed, uint32_t spp, uint64_t lane_begin, uint64_t lane_end, int nthreads, int normalize, double *out4) {
    OScene *os = (OScene *) h;
    try {
        const mib_sensor &s = os->sc.d.sensor;
        size_t npix = (size_t) s.crop_width * s.crop_height;
        if (lane_begin == 0 && lane_end == 0) lane_end = (uint64_t) npix * spp;
        const int C = n_raw_channels(os->sc.d.spectrum);
        std::vector<double> film(npix * C, 0.0);
        dispatch_render(os, seed, spp, lane_begin, lane_end, nthreads, film.data(), nullptr);
        int nch = os->sc.d.spectrum == MIB_SPEC_MONO_POLARIZED ? 4 : (os->sc.d.spectrum == MIB_SPEC_MONO ? 1 : 3);
        for (int c = 0; c < 4; ++c) out4[c] = 0.0;
        for (size_t p = 0; p < npix; ++p) for (int c = 0; c < nch; ++c) out4[c] += film[p * C + c];
        if (normalize) {
            float nf = rcp((float) npix * (float) spp);
            for (int c = 0; c < nch; ++c) out4[c] *= (double) nf;
        }
    } catch (const std::exception &e) { g_error = e.what(); return 1; }
    return 0;
}

int orc_counters_get(void *h, mib_counters *out) {
    OScene *os = (OScene *) h;
    out->n_samples = os->counters.n_samples; out->n_lookups = os->counters.n_lookups; out->n_real = os->counters.n_real;
    out->n_null = os->counters.n_null; out->n_nee_steps = os->counters.n_nee_steps; out->n_surface = os->counters.n_surface;
    out->kernel_ms = 0; out->n_launches = 0;
    return 0;
}

#include "oracle_adjoint_api.inc"
#include "oracle_test_api.inc"

} // extern "C"

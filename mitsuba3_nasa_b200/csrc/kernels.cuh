// kernels.cuh -- sm_100a kernels of the forward pass.
//
// Mapping: global lane l = (y*W + x)*spp + s, seeded TEA(base_seed+seed, l).
// Each thread owns `spt` CONSECUTIVE lanes, i.e. (part of) the samples of one pixel:
// the film contribution is accumulated in registers and leaves the thread with one
// red.global.add per channel when the pixel changes -- no per-sample atomics, no
// inter-thread exchange.  Grid = ceil(lanes / spt / 128) CTAs of 128 threads; with
// spt chosen by the host the grid is many multiples of 148 SMs x resident CTAs.
#pragma once
#include "astage.cuh"
#include <type_traits>
#include "docean.cuh"
#include "dmis.cuh"

#ifndef MIB_MINB
#define MIB_MINB 4          // __launch_bounds__(128, MIB_MINB): min resident CTAs per SM (register cap experiments)
#endif
#ifndef MIB_WARPSYNC
#define MIB_WARPSYNC 1      // 1: warp vote at the head of the flat loop (forced reconvergence every iteration)
#endif
#ifndef MIB_MINB_ADJ
#define MIB_MINB_ADJ 3
#endif

namespace mib {

struct LaunchParams {
    uint32_t seed_value;       // sampler base seed + render seed
    uint32_t spp;
    uint64_t lane_begin, lane_end;   // lanes of this launch; with stripes: VIRTUAL lanes [0, n_local) of this rank
    uint32_t stripe_rank, stripe_world;      // multi-GPU load balance: chunk c of stripe_chunk lanes belongs to rank c % world
    uint64_t stripe_chunk;                   // (stripe_world <= 1: no striping, virtual lane = lane)
    uint32_t spt;              // lanes per thread
    int mode;                  // 0 = film accumulate, 1 = render_1 sums (double), 2 = per-sample dump
    float *film;               // [H*W*C]
    double *sums;              // [4]
    float *per_sample;         // [lanes*C] (mode 2)
    unsigned long long *counters; // [8]: lookups, real, null, nee_steps, surface, samples, -, pool error code
};

// virtual lane of a rank -> lane of the wavefront (mib_scene_set_lane_stripes).  Chunks are dealt in groups of `world`:
// in group g rank r owns chunk g * world + ((r - g) mod world) -- the rotation by g keeps a rank from always getting the
// same column band when the row length is a multiple of world * chunk (512-pixel rows, 64-pixel chunks, 8 ranks).
MIB_D uint64_t stripe_map(const LaunchParams &lp, uint64_t v) {
    if (lp.stripe_world <= 1u) return v;
    const uint64_t g = v / lp.stripe_chunk, w = lp.stripe_world;
    const uint64_t c = g * w + ((lp.stripe_rank + w - g % w) % w);
    return c * lp.stripe_chunk + (v - g * lp.stripe_chunk);
}

// launchers defined in inst.cu (one per object file)
template <int SPEC, int MODE> void launch_k_forward(const DScene &scene, const LaunchParams &lp, unsigned blocks, cudaStream_t st);

template <int SPEC> struct RawChannels { static constexpr int C = SPEC == MIB_SPEC_MONO ? 2 : 4; };

// integrator.cpp:1245-1267 applied to a Stokes vector (first column of spec2sensor * spec)
MIB_D void to_sensor_stokes(const Ray &ray, float *S) {
    V3 md = -ray.d;
    V3 cur = stokes_basis(md);
    V3 tmp = cross(md, v3(0.f, 0.f, 1.f));
    bool vertical = norm(tmp) < kRayEps;
    V3 tgt = vertical ? v3(1.f, 0.f, 0.f) : cross(md, normalize(tmp));
    float c, s;
    stokes_rotation(md, cur, tgt, c, s);
    float s1 = S[1], s2 = S[2];
    S[1] = fma_(s, s2, c * s1);
    S[2] = fma_(c, s2, (-s) * s1);
}

// Reconstruction filters (src/rfilters/gaussian.cpp:86-98 non-CUDA branch: estrin in x^2; tent.cpp:41-43)
MIB_D float rfilter_eval(float x) {
    if (sc.d.sensor.rfilter == MIB_RFILTER_GAUSSIAN) {
        // drjit estrin_impl (math.h:26-41) for 10 coefficients
        const float *c = sc.rf_coeff;
        float x1 = sqr_(x), x2 = sqr_(x1), x4 = sqr_(x2), x8 = sqr_(x4);
        float a0 = fma_(x1, c[1], c[0]), a1 = fma_(x1, c[3], c[2]), a2 = fma_(x1, c[5], c[4]), a3 = fma_(x1, c[7], c[6]),
              a4 = fma_(x1, c[9], c[8]);
        float b0 = fma_(x2, a1, a0), b1 = fma_(x2, a3, a2), b2 = a4;
        float d0 = fma_(x4, b1, b0), d1 = b2;
        return fmaxf(fma_(x8, d1, d0), 0.f);
    }
    return fmaxf(0.f, 1.f - fabsf(x * sc.rf_inv_radius));
}

// ImageBlock::put for a reconstruction filter (imageblock.cpp:261-530, JIT recorded-loop forms, border 0): fn(pixel, weight)
// for every pixel of the footprint.  Coalesced form when spp >= 4 (integrator.cpp:320), general form otherwise.
template <typename Fn>
MIB_D void filter_footprint(uint32_t spp, float pos_x, float pos_y, Fn &&fn) {
    const mib_sensor &s = sc.d.sensor;
    const float radius = s.rfilter_radius;
    const uint32_t size_x = (uint32_t) s.crop_width, size_y = (uint32_t) s.crop_height;
    if (spp >= 4) {
        uint32_t n = (uint32_t) (int) ceilf(radius - .5f), count = 2 * n + 1;
        int pix = (int) floorf(pos_x) - (int) n, piy = (int) floorf(pos_y) - (int) n;
        uint32_t x = (uint32_t) (pix - s.crop_offset_x), y = (uint32_t) (piy - s.crop_offset_y);
        float rel_x = ((float) pix + .5f) - pos_x, rel_y = ((float) piy + .5f) - pos_y;
        for (uint32_t ys = 0; ys < count; ++ys) {
            float weight_y = rfilter_eval(rel_y + (float) ys);
            bool active_1 = y + ys < size_y;
            for (uint32_t xs = 0; xs < count; ++xs) {
                float weight = rfilter_eval(rel_x + (float) xs) * weight_y;
                if (active_1 && x + xs < size_x) fn((size_t) (y + ys) * size_x + (x + xs), weight);
            }
        }
    } else {
        float off_x = (float) (-s.crop_offset_x) - .5f, off_y = (float) (-s.crop_offset_y) - .5f;
        float pfx = pos_x + off_x, pfy = pos_y + off_y;
        float p0x = pfx - radius, p0y = pfy - radius, p1x = pfx + radius, p1y = pfy + radius;
        uint32_t x0 = (uint32_t) max((int) ceilf(p0x), 0), y0 = (uint32_t) max((int) ceilf(p0y), 0);
        uint32_t x1 = (uint32_t) min((int) floorf(p1x), (int) size_x - 1), y1 = (uint32_t) min((int) floorf(p1y), (int) size_y - 1);
        uint32_t count = (uint32_t) (int) ceilf(2.f * radius);
        bool active = x0 <= x1 && y0 <= y1;
        float rel_x = (float) x0 - pfx, rel_y = (float) y0 - pfy;
        for (uint32_t ys = 0; ys < count; ++ys) {
            float weight_y = rfilter_eval(rel_y + (float) ys);
            bool active_1 = active && (y0 + ys <= y1);
            for (uint32_t xs = 0; xs < count; ++xs) {
                float weight = rfilter_eval(rel_x + (float) xs) * weight_y;
                if (active_1 && (x0 + xs <= x1)) fn((size_t) (y0 + ys) * size_x + (x0 + xs), weight);
            }
        }
    }
}

// MODE: 0 volpath, 1 volpathaos, 2 prbvolpath (primal), 3 prb_volpathaos (primal), 4 volpathaos with a dielectric interface (docean.cuh),
//       5 volpathmis (dmis.cuh)
template <int SPEC, int MODE>
__global__ void __launch_bounds__(128, MODE >= 4 ? 2 : MIB_MINB) k_forward(const __grid_constant__ LaunchParams lp) {
    constexpr int C = RawChannels<SPEC>::C, NR = ST<SPEC>::NR;
    const mib_sensor &sn = sc.d.sensor;
    uint64_t tid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t l0 = lp.lane_begin + tid * lp.spt;
    uint64_t l1 = min(l0 + (uint64_t) lp.spt, lp.lane_end);
    Counters cnt = { 0, 0, 0, 0, 0 };
    unsigned long long n_samples = 0;
    float acc[C];
    double dacc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) { acc[c] = 0.f; dacc[c] = 0.0; }
    long long cur_pixel = -1;
    const float scale_x = 1.f / (float) sn.crop_width, scale_y = 1.f / (float) sn.crop_height;
    const float offset_x = -(float) sn.crop_offset_x * scale_x, offset_y = -(float) sn.crop_offset_y * scale_y;

    // Film / render_1 accumulation of one finished sample (imageblock.cpp:206-232 box put).
    auto finish = [&](uint64_t lane, const Ray &ray0, float sx, float sy, float *R) {
        n_samples++;
        float aov[C];
        if constexpr (SPEC == MIB_SPEC_RGB) { aov[0] = R[0]; aov[1] = R[1]; aov[2] = R[2]; aov[3] = 1.f; }
        else if constexpr (SPEC == MIB_SPEC_MONO) { aov[0] = R[0]; aov[1] = 1.f; }
        else { to_sensor_stokes(ray0, R); aov[0] = R[0]; aov[1] = R[1]; aov[2] = R[2]; aov[3] = R[3]; }
        if (lp.mode == 2) {
#pragma unroll
            for (int c = 0; c < C; ++c) lp.per_sample[(lane - lp.lane_begin) * C + c] = aov[c];
            return;
        }
        if (lp.mode == 0 && (MIB_FT(FT_FILTER) && sn.rfilter != MIB_RFILTER_BOX)) {
            filter_footprint(lp.spp, sx, sy, [&](size_t pixel, float w) {
#pragma unroll
                for (int c = 0; c < C; ++c) atomicAdd(lp.film + pixel * C + c, aov[c] * w);
            });
            return;
        }
        int ix = (int) floorf(sx) - sn.crop_offset_x, iy = (int) floorf(sy) - sn.crop_offset_y;
        if ((uint32_t) ix >= (uint32_t) sn.crop_width || (uint32_t) iy >= (uint32_t) sn.crop_height) return;
        if (lp.mode == 1) {
#pragma unroll
            for (int c = 0; c < C; ++c) dacc[c] += (double) aov[c];
        } else {
            long long pixel = (long long) iy * sn.crop_width + ix;
            if (pixel != cur_pixel) {
                if (cur_pixel >= 0) {
#pragma unroll
                    for (int c = 0; c < C; ++c) atomicAdd(lp.film + cur_pixel * C + c, acc[c]);
                }
#pragma unroll
                for (int c = 0; c < C; ++c) acc[c] = 0.f;
                cur_pixel = pixel;
            }
#pragma unroll
            for (int c = 0; c < C; ++c) acc[c] += aov[c];
        }
    };
    // integrator.cpp:323-340 (lane -> pixel) + render_sample :1139-1170 (film position, sensor ray)
    auto start = [&](uint64_t lane, Pcg32 &rng, float &sx, float &sy) -> Ray {
        lane = stripe_map(lp, lane);
        rng.seed_lane(lp.seed_value, (uint32_t) lane);
        uint32_t idx = (uint32_t) (lane / lp.spp);
        uint32_t py = idx / (uint32_t) sn.crop_width, px = idx - (uint32_t) sn.crop_width * py;
        float posx = (float) ((int) px + sn.crop_offset_x), posy = (float) ((int) py + sn.crop_offset_y);
        float jx = rng.next(), jy = rng.next();
        sx = posx + jx; sy = posy + jy;
        float ax = fma_(sx, scale_x, offset_x), ay = fma_(sy, scale_y, offset_y);
        float apx = .5f, apy = .5f;
        if (sn.needs_aperture_sample) { apx = rng.next(); apy = rng.next(); }
        return sensor_ray(ax, ay, apx, apy);
    };

    if constexpr (MODE <= 1 || MODE >= 4) {
        // Flat loop with lane regeneration: a thread starts its next lane as soon as its current
        // path ends instead of idling until the longest path of the warp is finished.
        typename std::conditional<MODE == 5, MisState<SPEC>, PathState<SPEC>>::type st;
        st.active = false;
        Pcg32 rng;
        Ray ray0;
        float sx = 0.f, sy = 0.f;
        uint64_t lane = l0, cur = l0;
#if MIB_WARPSYNC
        // All lanes of the warp stay in the loop until the last one is out of work; the vote at the loop head is a
        // reconvergence point, so the lanes execute each iteration's common code together instead of drifting apart.
        while (__any_sync(0xffffffffu, st.active || lane < l1)) {
            if (!st.active && lane < l1) {
                cur = lane++;
                ray0 = start(cur, rng, sx, sy);
                if constexpr (MODE == 5) mis_init<SPEC>(rng, ray0, st); else path_init<SPEC>(rng, ray0, st);
            }
            if (st.active) {
                if constexpr (MODE == 5) mis_step<SPEC>(rng, cnt, st);
                else if constexpr (MODE == 4) ocean_step<SPEC>(rng, cnt, st);
                else path_step<SPEC, MODE == 1>(rng, cnt, st);
                if (!st.active) finish(cur, ray0, sx, sy, st.R);
            }
        }
#else
        while (true) {
            if (!st.active) {
                if (lane >= l1) break;
                cur = lane++;
                ray0 = start(cur, rng, sx, sy);
                if constexpr (MODE == 5) mis_init<SPEC>(rng, ray0, st); else path_init<SPEC>(rng, ray0, st);
            }
            if constexpr (MODE == 5) mis_step<SPEC>(rng, cnt, st);
            else if constexpr (MODE == 4) ocean_step<SPEC>(rng, cnt, st);
            else path_step<SPEC, MODE == 1>(rng, cnt, st);
            if (!st.active) finish(cur, ray0, sx, sy, st.R);
        }
#endif
    } else {
        for (uint64_t lane = l0; lane < l1; ++lane) {
            Pcg32 rng;
            float sx, sy, R[NR];
            Ray ray = start(lane, rng, sx, sy);
            Sink sink;
            sink.init();
            prb_walk<SPEC, MODE == 3>(rng, cnt, ray, false, nullptr, nullptr, sink, R);
            finish(lane, ray, sx, sy, R);
        }
    }
    if (lp.mode == 0 && cur_pixel >= 0) {
#pragma unroll
        for (int c = 0; c < C; ++c) atomicAdd(lp.film + cur_pixel * C + c, acc[c]);
    }
    if (lp.mode == 1) {
        // warp-level pre-aggregation, then one atomic per warp and channel
#pragma unroll
        for (int c = 0; c < C; ++c) {
            double v = dacc[c];
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if ((threadIdx.x & 31) == 0) atomicAdd(lp.sums + c, v);
        }
    }
    unsigned long long cv[6] = { cnt.n_lookups, cnt.n_real, cnt.n_null, cnt.n_nee_steps, cnt.n_surface, n_samples };
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        unsigned long long v = cv[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(lp.counters + k, v);
    }
}

// ---------------------------------------------------------------------------
// Adjoint kernel (common.py:828-1102): per lane, primal walk -> L, then the replay with
// the same PCG32 state that scatters parameter gradients.  Same lane -> thread mapping as
// k_forward so that neighbouring threads walk neighbouring samples of one pixel.
// ---------------------------------------------------------------------------
template <int SPEC>
__global__ void __launch_bounds__(128, MIB_MINB_ADJ) k_adjoint(const __grid_constant__ LaunchParams lp,
                                                 const __grid_constant__ AdjParams ap) {
    constexpr int NR = ST<SPEC>::NR, Cd = SPEC == MIB_SPEC_MONO ? 1 : (SPEC == MIB_SPEC_RGB ? 3 : 4);
    constexpr bool AOS = ST<SPEC>::POL;
    const mib_sensor &sn = sc.d.sensor;
    uint64_t tid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t l0 = lp.lane_begin + tid * lp.spt;
    uint64_t l1 = min(l0 + (uint64_t) lp.spt, lp.lane_end);
    Counters cnt = { 0, 0, 0, 0, 0 };
    unsigned long long n_samples = 0;
    Sink sink;
    sink.init();
    const float scale_x = 1.f / (float) sn.crop_width, scale_y = 1.f / (float) sn.crop_height;
    const float offset_x = -(float) sn.crop_offset_x * scale_x, offset_y = -(float) sn.crop_offset_y * scale_y;
    const float nf = rcp_((float) ((uint32_t) sn.crop_width * (uint32_t) sn.crop_height * lp.spp));   // common.py:1046

    // Flat loop with lane regeneration (as k_forward): a thread alternates primal pass -> backward pass -> next lane
    // and never waits for the longest path of its warp; the vote at the loop head is the reconvergence point.
    PrbState<SPEC> st;
    st.active = false;
    int pass = 1;                                   // 0: primal pass in flight, 1: backward pass in flight / idle
    Pcg32 rng, start_rng;
    Ray ray0;
    float dl[4] = { 0.f, 0.f, 0.f, 0.f };
    uint64_t lane = l0;
    while (__any_sync(0xffffffffu, st.active || pass == 0 || lane < l1)) {
        if (!st.active) {
            if (pass == 0) {
                float Lp[NR];
#pragma unroll
                for (int i = 0; i < NR; ++i) Lp[i] = st.L[i];
                rng = start_rng;
                prb_init<SPEC, AOS>(rng, ray0, true, Lp, dl, st);
                pass = 1;
            } else if (lane < l1) {
                uint64_t cur = stripe_map(lp, lane++);
                rng.seed_lane(lp.seed_value, (uint32_t) cur);
                uint32_t idx = (uint32_t) (cur / lp.spp);
                uint32_t py = idx / (uint32_t) sn.crop_width, px = idx - (uint32_t) sn.crop_width * py;
                float posx = (float) ((int) px + sn.crop_offset_x), posy = (float) ((int) py + sn.crop_offset_y);
                float jx = rng.next(), jy = rng.next();
                float sx = posx + jx, sy = posy + jy;
                float ax = fma_(sx, scale_x, offset_x), ay = fma_(sy, scale_y, offset_y);
                float apx = .5f, apy = .5f;
                if (sn.needs_aperture_sample) { apx = rng.next(); apy = rng.next(); }
                ray0 = sensor_ray(ax, ay, apx, apy);
                n_samples++;
                dl[0] = dl[1] = dl[2] = dl[3] = 0.f;
                if (ap.render_1) {
#pragma unroll
                    for (int c = 0; c < Cd; ++c) dl[c] = ap.grad1[c] * nf;
                } else {
                    int ix = (int) floorf(sx) - sn.crop_offset_x, iy = (int) floorf(sy) - sn.crop_offset_y;
                    if ((MIB_FT(FT_FILTER) && sn.rfilter != MIB_RFILTER_BOX)) {
                        // adjoint of splat + weight division (hdrfilm.cpp:304-418): dL_c = sum_p grad[p,c] / W[p] * w_p
                        filter_footprint(lp.spp, sx, sy, [&](size_t pixel, float w) {
                            float W = __ldg(sc.weight_film + pixel), dv = W == 0.f ? 1.f : W;
#pragma unroll
                            for (int c = 0; c < Cd; ++c) dl[c] += div_(__ldg(ap.grad_image + pixel * Cd + c), dv) * w;
                        });
                    } else if ((uint32_t) ix < (uint32_t) sn.crop_width && (uint32_t) iy < (uint32_t) sn.crop_height) {
                        size_t pixel = (size_t) iy * sn.crop_width + ix;
#pragma unroll
                        for (int c = 0; c < Cd; ++c) dl[c] = div_(__ldg(ap.grad_image + pixel * Cd + c), (float) lp.spp);
                    }
                }
                start_rng = rng;
                prb_init<SPEC, AOS>(rng, ray0, false, nullptr, nullptr, st);
                pass = 0;
            }
        }
        if (st.active) prb_step<SPEC, AOS>(rng, cnt, sink, st);
    }
    sink.flush();
    unsigned long long cv[6] = { cnt.n_lookups, cnt.n_real, cnt.n_null, cnt.n_nee_steps, cnt.n_surface, n_samples };
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        unsigned long long v = cv[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(lp.counters + k, v);
    }
}

template <int SPEC> void launch_k_adjoint(const DScene &scene, const LaunchParams &lp, const AdjParams &ap, unsigned blocks, cudaStream_t st);

// Filter-weight film of the whole wavefront (positions only): the denominator HDRFilm::develop divides by.
static __global__ void k_weight_film(uint32_t seed_value, uint32_t spp, uint64_t total, float *wfilm) {
    const mib_sensor &sn = sc.d.sensor;
    for (uint64_t lane = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; lane < total; lane += (uint64_t) gridDim.x * blockDim.x) {
        Pcg32 rng;
        rng.seed_lane(seed_value, (uint32_t) lane);
        uint32_t idx = (uint32_t) (lane / spp);
        uint32_t py = idx / (uint32_t) sn.crop_width, px = idx - (uint32_t) sn.crop_width * py;
        float posx = (float) ((int) px + sn.crop_offset_x), posy = (float) ((int) py + sn.crop_offset_y);
        float jx = rng.next(), jy = rng.next();
        filter_footprint(spp, posx + jx, posy + jy, [&](size_t pixel, float w) { atomicAdd(wfilm + pixel, w); });
    }
}

// HDRFilm::develop (hdrfilm.cpp:304-418): value / (weight == 0 ? 1 : weight)
template <int SPEC>
__global__ void k_develop(const float *raw, float *out, size_t npix, uint32_t spp) {
    size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    if constexpr (SPEC == MIB_SPEC_RGB) {
        float w = raw[4 * p + 3], dv = (w == 0.f) ? 1.f : w;
        for (int c = 0; c < 3; ++c) out[3 * p + c] = div_(raw[4 * p + c], dv);
    } else if constexpr (SPEC == MIB_SPEC_MONO) {
        float I = raw[2 * p], w = raw[2 * p + 1], dv = (w == 0.f) ? 1.f : w;
        float lum = I * 0.212671f + I * 0.715160f + I * 0.072169f;
        out[p] = div_(lum, dv);
    } else {
        for (int c = 0; c < 4; ++c) out[4 * p + c] = (float) ((double) raw[4 * p + c] / (double) spp);
    }
}

// Majorant supergrid ("fast" mode): one CTA per supercell.  A point of cell i along an axis has local coordinate in
// [i/M, (i+1)/M], i.e. texel coordinate f = p*res - 0.5 in [i*res/M - 0.5, (i+1)*res/M - 0.5]; a trilinear look-up
// touches texels floor(f) and floor(f)+1 (clamped), a nearest look-up floor(p*res): the cell's majorant is the maximum over
// that index box, times scale, times (1 + 2^-20) against the rounding of the interpolation.
static __global__ void k_supergrid(const float *data, int rx, int ry, int rz, int ch, int M, float scale, float *out) {
    const int cx = blockIdx.x % M, cy = (blockIdx.x / M) % M, cz = blockIdx.x / (M * M);
    auto range = [M](int c, int r, int &lo, int &hi) {
        float a = (float) c * (float) r / (float) M - .5f, b = (float) (c + 1) * (float) r / (float) M - .5f;
        lo = max((int) floorf(a) - 1, 0);                 // one texel of slack on both sides (float rounding of p*res)
        hi = min((int) floorf(b) + 2, r - 1);
    };
    int x0, x1, y0, y1, z0, z1;
    range(cx, rx, x0, x1); range(cy, ry, y0, y1); range(cz, rz, z0, z1);
    const int nx = x1 - x0 + 1, ny = y1 - y0 + 1, nz = z1 - z0 + 1;
    float mval = 0.f;
    for (int i = threadIdx.x; i < nx * ny * nz; i += blockDim.x) {
        int x = x0 + i % nx, y = y0 + (i / nx) % ny, z = z0 + i / (nx * ny);
        size_t idx = (((size_t) z * ry + y) * rx + x) * ch;
        for (int c = 0; c < ch; ++c) mval = fmaxf(mval, data[idx + c]);
    }
    for (int o = 16; o > 0; o >>= 1) mval = fmaxf(mval, __shfl_xor_sync(0xffffffffu, mval, o));
    __shared__ float wm[32];
    if ((threadIdx.x & 31) == 0) wm[threadIdx.x >> 5] = mval;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int) (blockDim.x >> 5); ++w) mval = fmaxf(mval, wm[w]);
        out[blockIdx.x] = mval * scale * (1.f + 9.5367431640625e-07f);
    }
}

// Shading frames of the planar faces, evaluated once per commit with the device's own arithmetic (dscene.cuh face_frame)
static __global__ void k_shape_frames(float *out, float *phase_out) {
    const int shape = blockIdx.x, face = threadIdx.x;
    if (face == 7 && shape < MIB_MAX_PHASES) {                      // one spare thread per block: phase-function constants
        float r1 = 0.f, r2 = 0.f, r3 = 0.f;
        if (shape < sc.d.n_phases && sc.d.phases[shape].type == MIB_PHASE_RAYLEIGH)
            rayleigh_factors_compute(sc.d.phases[shape].depolarization, r1, r2, r3);
        phase_out[shape * 4 + 0] = r1; phase_out[shape * 4 + 1] = r2; phase_out[shape * 4 + 2] = r3; phase_out[shape * 4 + 3] = 0.f;
    }
    if (shape >= sc.d.n_shapes || face >= 6) return;
    float *o = out + (shape * 6 + face) * 6;
    for (int i = 0; i < 6; ++i) o[i] = 0.f;
    if (sc.d.shapes[shape].type == MIB_SHAPE_SPHERE) return;
    V3 n, dp_du, fs, ft;
    face_normal_dpdu(shape, face, n, dp_du);
    frame_from_normal(n, dp_du, fs, ft);
    o[0] = fs.x; o[1] = fs.y; o[2] = fs.z; o[3] = ft.x; o[4] = ft.y; o[5] = ft.z;
}

// grid maximum (gridvolume m_max, grid.cpp:306-319): grid-stride max + one atomic per block
static __global__ void k_max(const float *data, size_t n, unsigned int *out_bits) {
    float m = -inf_f();
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x)
        m = fmaxf(m, data[i]);
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) {
        // order-preserving float -> uint key
        unsigned int b = (unsigned int) __float_as_int(m);
        b = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
        atomicMax(out_bits, b);
    }
}

} // namespace mib

#include "dpool.cuh"

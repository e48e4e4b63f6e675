// apool.cuh -- k_adjoint_pool: the path-replay adjoint as a persistent CTA-resident path pool.
//
// Same machinery as k_forward_pool (dpool.cuh: slots in shared memory, one ring of slot ids per stage, warps
// claim 32 slots of one stage at a time) with the stages of astage.cuh.  A slot carries a lane through its
// primal pass (-> L) and then through the replay; START turns a finished primal pass into the replay by
// re-deriving the lane's sampler state and sensor ray from the lane index (nothing but L, dl and the lane index
// survives between the passes, as in the reference: common.py:828-983).  Gradients leave the kernel as
// red.global.add.f32 on the 8 trilinear taps (grids) or through the per-thread Sink (constant volumes).
#pragma once
#include "dpool.cuh"

namespace mib {

template <int SPEC> struct APoolL {
    static constexpr int NC = ST<SPEC>::NC, NT = ST<SPEC>::NT, NR = ST<SPEC>::NR;
    static constexpr int F_RNG = 0;                    // state lo, state hi, inc lo, inc hi
    static constexpr int F_LANE = 4;
    static constexpr int F_DEPTH = 5;
    static constexpr int F_FLAGS = 6;
    static constexpr int F_ETA = 7;
    static constexpr int F_NRNG = 8;                   // 2 words: sampler state at the start of the shadow walk / of the gradient walk
    static constexpr int F_RAY = 10;                   // o(3) d(3) maxt
    static constexpr int F_SI = 17;                    // t, p(3), shape | face << 16
    static constexpr int F_SCP = 22;                   // scatter_p(3)
    static constexpr int F_T = 25;
    static constexpr int F_L = F_T + NT;
    static constexpr int F_DL = F_L + NR;              // adjoint radiance (4)
    static constexpr int F_NRAY = F_DL + 4;            // shadow ray o(3) d(3)
    static constexpr int F_NSI = F_NRAY + 6;
    static constexpr int F_NED = F_NSI + 5;
    static constexpr int F_NTR = F_NED + 3;
    static constexpr int F_NEV = F_NTR + NC;
    static constexpr int F_NDIST = F_NEV + NC;         // total_dist, max_dist
    static constexpr int F_NA = F_NDIST + 2;           // contribution per unit emitted radiance, then the contribution
    static constexpr int F_COUNT = F_NA + NR;
    static constexpr int P = pool_slots_for(F_COUNT);
};
constexpr uint32_t FL_BACKWARD = 0x4;                  // (the forward kernel's FL_SPEC_CHAIN bit: prb walks have no specular chain)

template <int SPEC> __host__ __device__ constexpr size_t apool_smem_bytes() {
    return pool_state_offset(APoolL<SPEC>::P) + (size_t) APoolL<SPEC>::F_COUNT * APoolL<SPEC>::P * sizeof(uint32_t) + kPoolDebugBytes;
}

// PRIMAL: the primal-only instantiation (AdjParams::primal_only) -- its own kernel, so that the backward kernel carries none of its code
template <int SPEC, int V, bool PRIMAL>
__global__ void __launch_bounds__(MIB_POOL_NT, 1) k_adjoint_pool(const __grid_constant__ LaunchParams lp, const __grid_constant__ AdjParams ap) {
    using L = APoolL<SPEC>;
    constexpr int NR = ST<SPEC>::NR, NC = ST<SPEC>::NC, NT = ST<SPEC>::NT, Cd = SPEC == MIB_SPEC_MONO ? 1 : (SPEC == MIB_SPEC_RGB ? 3 : 4);
    constexpr bool AOS = ST<SPEC>::POL;
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int P = L::P;
    static_assert(apool_smem_bytes<SPEC>() <= 232448, "pool does not fit the shared memory of sm_100");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    PoolMem<P> pm;
    pm.pw = reinterpret_cast<uint32_t *>(smem_raw + pool_state_offset(P));
#ifdef MIB_POOL_DEBUG
    pm.n_fields = L::F_COUNT;
#endif
    const mib_sensor &sn = sc.d.sensor;
    const unsigned lane_id = threadIdx.x & 31;
    PoolQ<P, kAdjGroupMap> q;
    q.init(smem_raw, lp.lane_end - lp.lane_begin, lp.counters + 6, C_START);
#ifdef MIB_POOL_DEBUG
    q.dbg_init(smem_raw + pool_state_offset(P) + (size_t) L::F_COUNT * P * sizeof(uint32_t), C_START);
#endif
    PoolShared *ps = q.ps;
    for (int s = threadIdx.x; s < P; s += blockDim.x) { pm.su(L::F_LANE, s, kNoLane); pm.su(L::F_FLAGS, s, 0u); }
    __syncthreads();

    Counters cnt = { 0, 0, 0, 0, 0 };
    unsigned long long n_samples = 0;
    Sink sink;
    sink.init();
    const float scale_x = 1.f / (float) sn.crop_width, scale_y = 1.f / (float) sn.crop_height;
    const float offset_x = -(float) sn.crop_offset_x * scale_x, offset_y = -(float) sn.crop_offset_y * scale_y;
    const float nf = rcp_((float) ((uint32_t) sn.crop_width * (uint32_t) sn.crop_height * lp.spp));   // common.py:1046

    // shared-memory images of the walk state
    auto store_nee = [&](int slot, const PrbNee<SPEC> &ns) {
        pm.sv3(L::F_NRAY, slot, ns.ray.o); pm.sv3(L::F_NRAY + 3, slot, ns.ray.d);
        pm.store_si(L::F_NSI, slot, ns.si);
        pm.sv3(L::F_NED, slot, ns.emitter_d);
#pragma unroll
        for (int c = 0; c < NC; ++c) { pm.sf(L::F_NTR + c, slot, ns.tr[c]); pm.sf(L::F_NEV + c, slot, ns.ev[c]); }
        pm.sf(L::F_NDIST, slot, ns.total_dist); pm.sf(L::F_NDIST + 1, slot, ns.max_dist);
    };
    auto nee_flags = [&](const PrbNee<SPEC> &ns) -> uint32_t {
        return (ns.active ? FL_NEE_ACTIVE : 0u) | (ns.needs_intersection ? FL_NEE_NEEDS_INT : 0u) | ((uint32_t) (ns.medium + 1) << 16);
    };
    auto load_nee = [&](int slot, uint32_t fl, PrbNee<SPEC> &ns) {
        ns.active = fl & FL_NEE_ACTIVE;
        ns.needs_intersection = fl & FL_NEE_NEEDS_INT;
        ns.medium = (int) fl_nee_medium(fl) - 1;
        ns.ray.o = pm.v3f(L::F_NRAY, slot); ns.ray.d = pm.v3f(L::F_NRAY + 3, slot); ns.ray.maxt = 0.f;
        pm.load_si(L::F_NSI, slot, ns.si);
        if (!ns.needs_intersection && ns.si.shape >= 0) si_finalize(ns.ray.d, ns.si);
        ns.emitter_d = pm.v3f(L::F_NED, slot);
#pragma unroll
        for (int c = 0; c < NC; ++c) { ns.tr[c] = pm.f(L::F_NTR + c, slot); ns.ev[c] = pm.f(L::F_NEV + c, slot); }
        ns.total_dist = pm.f(L::F_NDIST, slot); ns.max_dist = pm.f(L::F_NDIST + 1, slot);
    };
    constexpr uint32_t kNeeBits = FL_NEE_ACTIVE | FL_NEE_NEEDS_INT | FL_NEE_SURFACE | 0xff0000u;

    while (true) {
        int cls, slot;
        bool has;
        if (!q.claim(cls, has, slot)) break;
        int next = -1;
        switch (cls) {
        case C_START: {
            // a finished primal pass becomes the replay of the same lane; a finished replay (or an empty slot) takes the next lane
            uint32_t j = kNoLane;
            bool replay = false;
            if (has) {
                j = pm.u(L::F_LANE, slot);
                replay = j != kNoLane && !(pm.u(L::F_FLAGS, slot) & FL_BACKWARD);
            }
            if constexpr (PRIMAL) {
                // primal rendering with the prb integrators (mi.render / render_1 before backward()): a finished primal pass is a
                // finished sample -- L goes to the film / the render_1 sums / the per-sample dump like in the forward kernels
                // (kernels.cuh finish), and the slot takes the next lane
                constexpr int C = RawChannels<SPEC>::C;
                float aov[C];
#pragma unroll
                for (int c = 0; c < C; ++c) aov[c] = 0.f;
                bool inside = false;
                if (replay) {
                    const uint64_t lane = stripe_map(lp, lp.lane_begin + q.lane_of(j));
                    Pcg32 rng;
                    rng.seed_lane(lp.seed_value, (uint32_t) lane);
                    const uint32_t idx = (uint32_t) (lane / lp.spp);
                    const uint32_t py = idx / (uint32_t) sn.crop_width, px = idx - (uint32_t) sn.crop_width * py;
                    const float sx = (float) ((int) px + sn.crop_offset_x) + rng.next(), sy = (float) ((int) py + sn.crop_offset_y) + rng.next();
                    float R[NR];
#pragma unroll
                    for (int r = 0; r < NR; ++r) R[r] = pm.f(L::F_L + r, slot);
                    if constexpr (SPEC == MIB_SPEC_RGB) { aov[0] = R[0]; aov[1] = R[1]; aov[2] = R[2]; aov[3] = 1.f; }
                    else if constexpr (SPEC == MIB_SPEC_MONO) { aov[0] = R[0]; aov[1] = 1.f; }
                    else {
                        const float ax = fma_(sx, scale_x, offset_x), ay = fma_(sy, scale_y, offset_y);
                        float apx = .5f, apy = .5f;
                        if (sn.needs_aperture_sample) { apx = rng.next(); apy = rng.next(); }
                        const Ray ray0 = sensor_ray(ax, ay, apx, apy);
                        to_sensor_stokes(ray0, R);
                        aov[0] = R[0]; aov[1] = R[1]; aov[2] = R[2]; aov[3] = R[3];
                    }
                    const int ix = (int) floorf(sx) - sn.crop_offset_x, iy = (int) floorf(sy) - sn.crop_offset_y;
                    inside = (uint32_t) ix < (uint32_t) sn.crop_width && (uint32_t) iy < (uint32_t) sn.crop_height;
                    if (lp.mode == 2) {
#pragma unroll
                        for (int c = 0; c < C; ++c) lp.per_sample[(lane - lp.lane_begin) * C + c] = aov[c];
                    } else if (lp.mode == 0 && (MIB_FT(FT_FILTER) && sn.rfilter != MIB_RFILTER_BOX)) {
                        filter_footprint(lp.spp, sx, sy, [&](size_t pixel, float w) {
#pragma unroll
                            for (int c = 0; c < C; ++c) atomicAdd(lp.film + pixel * C + c, aov[c] * w);
                        });
                    } else if (lp.mode == 0 && inside) {
                        const size_t pixel = (size_t) iy * sn.crop_width + ix;
#pragma unroll
                        for (int c = 0; c < C; ++c) atomicAdd(lp.film + pixel * C + c, aov[c]);
                    }
                }
                if (lp.mode == 1) {
                    // render_1: one double atomic per batch and channel (the batch is summed over the warp first)
#pragma unroll
                    for (int c = 0; c < C; ++c) {
                        double v = (replay && inside) ? (double) aov[c] : 0.0;
                        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
                        if (lane_id == 0 && v != 0.0) atomicAdd(lp.sums + c, v);
                    }
                }
                replay = false;
            }
            const bool need = has && !replay;
            const uint32_t jn = q.take_lanes(need);
            bool fresh = false;
            if (need) {
                if (jn != kNoLane) { j = jn; fresh = true; }
                else { pm.su(L::F_LANE, slot, kNoLane); q.retire(slot); }
            }
            if (replay || fresh) {
                const uint64_t lane = stripe_map(lp, lp.lane_begin + q.lane_of(j));
                // integrator.cpp:323-340 (lane -> pixel) + render_sample :1139-1170 (film position, sensor ray)
                Pcg32 rng;
                rng.seed_lane(lp.seed_value, (uint32_t) lane);
                uint32_t idx = (uint32_t) (lane / lp.spp);
                uint32_t py = idx / (uint32_t) sn.crop_width, px = idx - (uint32_t) sn.crop_width * py;
                float posx = (float) ((int) px + sn.crop_offset_x), posy = (float) ((int) py + sn.crop_offset_y);
                float jx = rng.next(), jy = rng.next();
                float sx = posx + jx, sy = posy + jy;
                float ax = fma_(sx, scale_x, offset_x), ay = fma_(sy, scale_y, offset_y);
                float apx = .5f, apy = .5f;
                if (sn.needs_aperture_sample) { apx = rng.next(); apy = rng.next(); }
                Ray ray0 = sensor_ray(ax, ay, apx, apy);
                float dl[4] = { 0.f, 0.f, 0.f, 0.f }, Lin[NR];
#pragma unroll
                for (int r = 0; r < NR; ++r) Lin[r] = 0.f;
                if (fresh) {
                    n_samples++;
                    if (PRIMAL) {
                        // no adjoint radiance: the replay never runs
                    } else if (ap.render_1) {
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
#pragma unroll
                    for (int c = 0; c < 4; ++c) pm.sf(L::F_DL + c, slot, dl[c]);
                } else {
#pragma unroll
                    for (int c = 0; c < 4; ++c) dl[c] = pm.f(L::F_DL + c, slot);
#pragma unroll
                    for (int r = 0; r < NR; ++r) Lin[r] = pm.f(L::F_L + r, slot);
                }
                PrbState<SPEC> st;
                prb_init<SPEC, AOS>(rng, ray0, replay, Lin, dl, st);
                pm.store_rng(slot, rng);
                pm.su(L::F_LANE, slot, j);
                pm.su(L::F_DEPTH, slot, 0u);
                pm.su(L::F_FLAGS, slot, st.channel | FL_NEEDS_INT | (replay ? FL_BACKWARD : 0u) | ((uint32_t) (st.medium + 1) << 8));
                pm.sf(L::F_ETA, slot, 1.f);
                pm.sv3(L::F_RAY, slot, ray0.o); pm.sv3(L::F_RAY + 3, slot, ray0.d); pm.sf(L::F_RAY + 6, slot, ray0.maxt);
                pm.store_si(L::F_SI, slot, st.si);
#pragma unroll
                for (int i = 0; i < NT; ++i) pm.sf(L::F_T + i, slot, st.T[i]);
#pragma unroll
                for (int r = 0; r < NR; ++r) pm.sf(L::F_L + r, slot, st.L[r]);
                next = replay ? C_MAINB : C_MAIN;
            }
            break;
        }
        case C_MAINB:            // MAIN of the replay in its own queue: a batch either scatters gradients or it does not
        case C_MAIN: {
            if (has) {
                PrbState<SPEC> st;
                Pcg32 rng;
                pm.load_rng(slot, rng);
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                st.depth = pm.u(L::F_DEPTH, slot);
                st.channel = fl & FL_CHANNEL;
                st.needs_intersection = fl & FL_NEEDS_INT;
                st.backward = fl & FL_BACKWARD;
                st.medium = (int) fl_medium(fl) - 1;
                st.eta = pm.f(L::F_ETA, slot);
                st.ray.o = pm.v3f(L::F_RAY, slot); st.ray.d = pm.v3f(L::F_RAY + 3, slot); st.ray.maxt = pm.f(L::F_RAY + 6, slot);
                pm.load_si(L::F_SI, slot, st.si);
                if (!st.needs_intersection && st.si.shape >= 0) si_finalize(st.ray.d, st.si);
#pragma unroll
                for (int i = 0; i < NT; ++i) st.T[i] = pm.f(L::F_T + i, slot);
#pragma unroll
                for (int r = 0; r < NR; ++r) st.L[r] = pm.f(L::F_L + r, slot);
#pragma unroll
                for (int c = 0; c < 4; ++c) st.dl[c] = pm.f(L::F_DL + c, slot);
                st.dl_sum = (AOS && st.backward) ? ((st.dl[0] + st.dl[1]) + st.dl[2]) + st.dl[3] : 0.f;
                int ev = EV_NULL;
                for (int it = 0; it < kAdjMainIters && ev == EV_NULL; ++it) ev = prb_main<SPEC, AOS>(rng, cnt, sink, st);
                if (ev == EV_DONE) next = C_START;
                else {
                    pm.store_rng_state(slot, rng);
#pragma unroll
                    for (int i = 0; i < NT; ++i) pm.sf(L::F_T + i, slot, st.T[i]);
                    pm.su(L::F_DEPTH, slot, st.depth);
                    fl = (fl & ~(FL_NEEDS_INT | kNeeBits)) | (st.needs_intersection ? FL_NEEDS_INT : 0u);
                    pm.sv3(L::F_RAY, slot, st.ray.o); pm.sf(L::F_RAY + 6, slot, st.ray.maxt);
                    pm.store_si(L::F_SI, slot, st.si);
                    if (ev == EV_NULL) next = cls;
                    else if (ev == EV_SCATTER) { pm.sv3(L::F_SCP, slot, st.scatter_p); next = C_SCAT; }
                    else if (ev == EV_NEE) {
                        store_nee(slot, st.nee);
#pragma unroll
                        for (int r = 0; r < NR; ++r) pm.sf(L::F_NA + r, slot, st.a[r]);
                        pm.su(L::F_NRNG, slot, (uint32_t) st.nee_rng.state); pm.su(L::F_NRNG + 1, slot, (uint32_t) (st.nee_rng.state >> 32));
                        fl |= nee_flags(st.nee) | FL_NEE_SURFACE;
                        next = C_NEE;
                    } else next = C_BSDF;
                    pm.su(L::F_FLAGS, slot, fl);
                }
            }
            break;
        }
        case C_SCAT: {
            if (has) {
                PrbState<SPEC> st;
                Pcg32 rng;
                pm.load_rng(slot, rng);
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                st.medium = (int) fl_medium(fl) - 1;
                st.channel = fl & FL_CHANNEL;
                st.ray.d = pm.v3f(L::F_RAY + 3, slot);
                st.scatter_p = pm.v3f(L::F_SCP, slot);
#pragma unroll
                for (int i = 0; i < NT; ++i) st.T[i] = pm.f(L::F_T + i, slot);
                if (prb_scatter<SPEC, AOS>(rng, st)) {
                    pm.store_rng_state(slot, rng);
                    store_nee(slot, st.nee);
#pragma unroll
                    for (int r = 0; r < NR; ++r) pm.sf(L::F_NA + r, slot, st.a[r]);
                    pm.su(L::F_NRNG, slot, (uint32_t) st.nee_rng.state); pm.su(L::F_NRNG + 1, slot, (uint32_t) (st.nee_rng.state >> 32));
                    pm.su(L::F_FLAGS, slot, (fl & ~kNeeBits) | nee_flags(st.nee));
                    next = C_NEE;
                } else next = C_PHASE;
            }
            break;
        }
        case C_NEE: {
            // value walk; at its end the contribution goes into L and, in the replay, the walk restarts as a gradient walk
            if (has) {
                PrbState<SPEC> st;
                Pcg32 rng;
                pm.load_rng(slot, rng);
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                st.channel = fl & FL_CHANNEL;
                load_nee(slot, fl, st.nee);
                for (int it = 0; it < kAdjNeeIters && st.nee.active; ++it)
                    prb_nee_step<SPEC, AOS>(rng, cnt, sink, st.channel, st.nee, false, nullptr, nullptr);
                pm.store_rng_state(slot, rng);
                if (st.nee.active) {
                    store_nee(slot, st.nee);
                    pm.su(L::F_FLAGS, slot, (fl & ~(FL_NEE_ACTIVE | FL_NEE_NEEDS_INT | 0xff0000u)) | nee_flags(st.nee));
                    next = C_NEE;
                } else {
                    st.backward = fl & FL_BACKWARD;
                    st.nee_surface = fl & FL_NEE_SURFACE;
                    st.medium = (int) fl_medium(fl) - 1;
#pragma unroll
                    for (int r = 0; r < NR; ++r) { st.a[r] = pm.f(L::F_NA + r, slot); st.L[r] = pm.f(L::F_L + r, slot); }
                    if (st.backward) {
                        st.nee_rng.state = (uint64_t) pm.u(L::F_NRNG, slot) | ((uint64_t) pm.u(L::F_NRNG + 1, slot) << 32);
                        st.nee_rng.inc = rng.inc;
                        if (st.nee_surface) {
                            st.ray.d = pm.v3f(L::F_RAY + 3, slot);
                            pm.load_si(L::F_SI, slot, st.si);
                            si_finalize(st.ray.d, st.si);
                        } else st.scatter_p = pm.v3f(L::F_SCP, slot);
                    }
                    Pcg32 grad_rng;
                    const bool again = prb_nee_end<SPEC, AOS>(st, grad_rng);
#pragma unroll
                    for (int r = 0; r < NR; ++r) { pm.sf(L::F_NA + r, slot, st.a[r]); pm.sf(L::F_L + r, slot, st.L[r]); }
                    if (again) {
                        store_nee(slot, st.nee);
                        pm.su(L::F_NRNG, slot, (uint32_t) grad_rng.state); pm.su(L::F_NRNG + 1, slot, (uint32_t) (grad_rng.state >> 32));
                        pm.su(L::F_FLAGS, slot, (fl & ~(FL_NEE_ACTIVE | FL_NEE_NEEDS_INT | 0xff0000u)) | nee_flags(st.nee));
                        next = C_NEEG;
                    } else next = st.nee_surface ? C_BSDF : C_PHASE;
                }
            }
            break;
        }
        case C_NEEG: {
            // gradient walk of the replay: same decisions as the value walk, scatters d(transmittance)/d(sigma_t)
            if (has) {
                PrbNee<SPEC> ns;
                Pcg32 grng;
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                grng.state = (uint64_t) pm.u(L::F_NRNG, slot) | ((uint64_t) pm.u(L::F_NRNG + 1, slot) << 32);
                grng.inc = (uint64_t) pm.u(L::F_RNG + 2, slot) | ((uint64_t) pm.u(L::F_RNG + 3, slot) << 32);
                load_nee(slot, fl, ns);
                float adj[NR], dl[4];
#pragma unroll
                for (int r = 0; r < NR; ++r) adj[r] = pm.f(L::F_NA + r, slot);
#pragma unroll
                for (int c = 0; c < 4; ++c) dl[c] = pm.f(L::F_DL + c, slot);
                for (int it = 0; it < kAdjNeeIters && ns.active; ++it)
                    prb_nee_step<SPEC, AOS>(grng, cnt, sink, fl & FL_CHANNEL, ns, true, adj, dl);
                if (ns.active) {
                    pm.su(L::F_NRNG, slot, (uint32_t) grng.state); pm.su(L::F_NRNG + 1, slot, (uint32_t) (grng.state >> 32));
                    store_nee(slot, ns);
                    pm.su(L::F_FLAGS, slot, (fl & ~(FL_NEE_ACTIVE | FL_NEE_NEEDS_INT | 0xff0000u)) | nee_flags(ns));
                    next = C_NEEG;
                } else next = (fl & FL_NEE_SURFACE) ? C_BSDF : C_PHASE;
            }
            break;
        }
        case C_PHASE: {
            if (has) {
                PrbState<SPEC> st;
                Pcg32 rng;
                pm.load_rng(slot, rng);
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                st.medium = (int) fl_medium(fl) - 1;
                st.needs_intersection = fl & FL_NEEDS_INT;
                st.ray.o = pm.v3f(L::F_RAY, slot); st.ray.d = pm.v3f(L::F_RAY + 3, slot); st.ray.maxt = pm.f(L::F_RAY + 6, slot);
                st.scatter_p = pm.v3f(L::F_SCP, slot);
#pragma unroll
                for (int i = 0; i < NT; ++i) st.T[i] = pm.f(L::F_T + i, slot);
                prb_phase<SPEC, AOS>(rng, cnt, st);
                pm.store_rng_state(slot, rng);
#pragma unroll
                for (int i = 0; i < NT; ++i) pm.sf(L::F_T + i, slot, st.T[i]);
                pm.sv3(L::F_RAY, slot, st.ray.o); pm.sv3(L::F_RAY + 3, slot, st.ray.d); pm.sf(L::F_RAY + 6, slot, st.ray.maxt);
                pm.su(L::F_FLAGS, slot, (fl & ~FL_NEEDS_INT) | (st.needs_intersection ? FL_NEEDS_INT : 0u));
                next = (fl & FL_BACKWARD) ? C_MAINB : C_MAIN;
            }
            break;
        }
        default: {   // C_BSDF
            if (has) {
                PrbState<SPEC> st;
                Pcg32 rng;
                pm.load_rng(slot, rng);
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                st.medium = (int) fl_medium(fl) - 1;
                st.depth = pm.u(L::F_DEPTH, slot);
                st.ray.o = pm.v3f(L::F_RAY, slot); st.ray.d = pm.v3f(L::F_RAY + 3, slot); st.ray.maxt = pm.f(L::F_RAY + 6, slot);
                pm.load_si(L::F_SI, slot, st.si);
                si_finalize(st.ray.d, st.si);
#pragma unroll
                for (int i = 0; i < NT; ++i) st.T[i] = pm.f(L::F_T + i, slot);
                if (prb_bsdf<SPEC, AOS>(rng, st)) {
                    pm.store_rng_state(slot, rng);
#pragma unroll
                    for (int i = 0; i < NT; ++i) pm.sf(L::F_T + i, slot, st.T[i]);
                    pm.sv3(L::F_RAY, slot, st.ray.o); pm.sv3(L::F_RAY + 3, slot, st.ray.d); pm.sf(L::F_RAY + 6, slot, st.ray.maxt);
                    pm.su(L::F_DEPTH, slot, st.depth);
                    pm.su(L::F_FLAGS, slot, (fl & ~(FL_NEEDS_INT | 0xff00u)) | FL_NEEDS_INT | ((uint32_t) (st.medium + 1) << 8));
                    next = (fl & FL_BACKWARD) ? C_MAINB : C_MAIN;
                } else next = C_START;
            }
            break;
        }
        }
        q.push(next, slot);
    }

    sink.flush();
    unsigned long long cv[6] = { cnt.n_lookups, cnt.n_real, cnt.n_null, cnt.n_nee_steps, cnt.n_surface, n_samples };
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        unsigned long long v = cv[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
        if (lane_id == 0 && v) atomicAdd(lp.counters + k, v);
    }
    if (threadIdx.x == 0 && ps->error) atomicExch((unsigned int *) (lp.counters + 7), ps->error);
}

template <int SPEC, int V> void launch_k_adjoint_pool(const DScene &scene, const LaunchParams &lp, const AdjParams &ap, unsigned blocks, unsigned threads, cudaStream_t st);
template <int SPEC> int adjoint_pool_slots();

} // namespace mib

// adjoint.cuh -- path-replay backpropagation on the device.
//   prbvolpath      src/python/python/ad/integrators/prbvolpath.py:91-431      (mono, rgb)
//   prb_volpathaos  src/python/python/ad/integrators/prb_volpathaos.py:95-503  (mono_polarized)
// The reference records nothing during the primal pass: it runs the walk once to get
// L, then replays it with the same sampler state and differentiates each step locally
// (dr.backward on `weight` / `tr_multiplier`).  Here both passes run back to back in the
// same thread: pass 0 produces L in registers, pass 1 re-seeds the lane's PCG32 and
// scatters d(loss)/d(sigma_t grid), d/d(albedo grid) through the 8 trilinear taps with
// red.global.add.f32, and d/d(constant sigma_t | albedo) into per-thread doubles.
//
// Reference quirks kept on purpose (tests pin them against the oracle):
//  * prb_volpathaos back-propagates through I = replace_grad(1, weight) without
//    dividing by the weight (prb_volpathaos.py:214-217, :483-486).
//  * render_1_backward hands the sensor-frame adjoint Stokes vector to sample()
//    without rotating it back (common.py:1024-1072).
#pragma once
#include "dwalk.cuh"

namespace mib {

struct AdjParams {
    const float *grad_image;   // [H*W*Cd] image gradient (render_backward) or nullptr
    float grad1[4];            // d loss / d (I,Q,U,V)  (render_1_backward)
    int render_1;
    int primal_only;           // pool kernel: stop after the primal pass and accumulate L as the forward kernels do (LaunchParams::mode)
};

// Per-thread gradient sink.  Grid gradients go straight to global memory; scalar
// (constant-volume) gradients are kept in registers for one medium at a time.
struct Sink {
    double acc[4];             // [0..2] albedo (per channel), [3] sigma_t
    int medium;
    MIB_D void init() { acc[0] = acc[1] = acc[2] = acc[3] = 0.0; medium = -1; }
    MIB_D void flush() {
        if (medium >= 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) if (acc[k] != 0.0) atomicAdd(sc.grad_scalar + medium * 4 + k, acc[k]);
        }
        acc[0] = acc[1] = acc[2] = acc[3] = 0.0;
    }
    MIB_D void scalar(int m, int k, float v) {
        if (m != medium) { flush(); medium = m; }
        acc[k] += (double) v;
    }
};

// One gradient contribution to a grid voxel.  Shipped: one red.global.add.f32 per lane and tap.
// Variant MIB_GRAD_WARP_AGG (A/B only, `make EXTRA=-DMIB_GRAD_WARP_AGG B=build_wagg OUT=../libmi_b200_wagg.so`, see
// profiles/ab.sh): lanes of the converged group that hit the same voxel are found with match.any, their values are
// summed by shuffles and the lowest lane issues one atomic.  Not the default: on C4 the ncu capture of k_adjoint_pool shows
// the atomics at 7 of 32 lanes and < 2 % of issued instructions, and consecutive lanes (same pixel) only share
// voxels up to their first scatter, so the match.any round trip (8 per collision) costs more than the atomics it saves.
MIB_D void grad_add(float *addr, float v) {
#ifdef MIB_GRAD_WARP_AGG
    const unsigned act = __activemask();
    const unsigned peers = __match_any_sync(act, (unsigned long long) addr);
    const unsigned lane = threadIdx.x & 31u;
    if (peers != (1u << lane)) {
        float sum = 0.f;
        for (unsigned m = peers; m; m &= m - 1) sum += __shfl_sync(peers, v, __ffs(m) - 1);
        if (lane != (unsigned) (__ffs(peers) - 1)) return;
        v = sum;
    }
#endif
    atomicAdd(addr, v);
}

// d(loss) += coeff[c] * d(volume value at the taps, channel c)
template <int NC>
MIB_D void sink_volume(Sink &sink, const mib_volume &v, int medium, bool is_albedo, const GridTaps &taps, const float *coeff) {
    if (v.type == MIB_VOLUME_GRID) {
        float *g = sc.grad[v.buffer];
        if (!g) return;
        const int n = v.filter == MIB_FILTER_NEAREST ? 1 : 8;
        if (v.channels == 1) {
            float s = coeff[0];
#pragma unroll
            for (int c = 1; c < NC; ++c) s += coeff[c];
            if (s == 0.f) return;
#ifndef MIB_NO_GRAD_ATOMICS
            for (int k = 0; k < n; ++k) grad_add(g + taps.idx[k], s * taps.w[k]);
#endif
        } else if constexpr (NC == 1) {
            // a 3-channel grid seen through a 1-channel variant is its luminance (dscene.cuh eval_grid): d value / d texel
            // channel = the luminance weight
            for (int k = 0; k < n; ++k) {
                const float cw = coeff[0] * taps.w[k];
                grad_add(g + taps.idx[k], cw * 0.212671f);
                grad_add(g + taps.idx[k] + 1, cw * 0.715160f);
                grad_add(g + taps.idx[k] + 2, cw * 0.072169f);
            }
        } else {
            for (int k = 0; k < n; ++k)
#pragma unroll
                for (int c = 0; c < NC; ++c) grad_add(g + taps.idx[k] + c, coeff[c] * taps.w[k]);
        }
    } else {
#pragma unroll
        for (int c = 0; c < NC; ++c) sink.scalar(medium, is_albedo ? (v.channels == 1 ? 0 : c) : 3, coeff[c]);
    }
}

} // namespace mib

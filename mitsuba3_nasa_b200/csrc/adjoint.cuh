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
            for (int k = 0; k < n; ++k) atomicAdd(g + taps.idx[k], s * taps.w[k]);
#endif
        } else {
            for (int k = 0; k < n; ++k)
#pragma unroll
                for (int c = 0; c < NC; ++c) atomicAdd(g + taps.idx[k] + c, coeff[c] * taps.w[k]);
        }
    } else {
#pragma unroll
        for (int c = 0; c < NC; ++c) sink.scalar(medium, is_albedo ? (v.channels == 1 ? 0 : c) : 3, coeff[c]);
    }
}

} // namespace mib

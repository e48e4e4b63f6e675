// oadjoint.h -- TEST INFRASTRUCTURE (parity oracle). Not part of the product.
// Path-replay adjoint restatement (prbvolpath.py / prb_volpathaos.py). Filled in below.
#pragma once
#include "owalk.h"
namespace orc {
template <int SPEC> struct Adjoint {
    using Tr = SpecTraits<SPEC>;
    using Spec = typename Tr::Spec;
    const Scene &sc; Sampler &sampler; Counters &cnt;
    Adjoint(const Scene &s, Sampler &sm, Counters &c) : sc(s), sampler(sm), cnt(c) {}
    Spec primal_prbvolpath(Ray, int, bool *) { throw std::runtime_error("oracle: prbvolpath not implemented yet"); }
    Spec primal_prbvolpathaos(Ray, int, bool *) { throw std::runtime_error("oracle: prb_volpathaos not implemented yet"); }
};
}

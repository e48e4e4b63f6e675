// inst.cu -- one explicit instantiation of a walk kernel per object file.
// Compiled as:  nvcc -DINST_FWD|-DINST_POOL -DINST_SPEC=s -DINST_MODE=m   or   nvcc -DINST_ADJ|-DINST_APOOL -DINST_SPEC=s   (see Makefile).
#include "kernels.cuh"
#include "apool.cuh"

namespace mib {
#if defined(INST_FWD)
template <> void launch_k_forward<INST_SPEC, INST_MODE>(const DScene &scene, const LaunchParams &lp, unsigned blocks, cudaStream_t st) {
    cudaMemcpyToSymbolAsync(sc, &scene, sizeof(DScene), 0, cudaMemcpyHostToDevice, st);
    k_forward<INST_SPEC, INST_MODE><<<blocks, 128, 0, st>>>(lp);
}
#elif defined(INST_POOL)
#ifndef INST_VARIANT
#define INST_VARIANT 0
#endif
template <> void launch_k_forward_pool<INST_SPEC, INST_MODE, INST_VARIANT>(const DScene &scene, const LaunchParams &lp, unsigned blocks, unsigned threads, cudaStream_t st) {
    const size_t smem = pool_smem_bytes<INST_SPEC>();
    // per device and cheap: set at every launch (a process may hold scenes on several devices)
    cudaFuncSetAttribute(k_forward_pool<INST_SPEC, INST_MODE, INST_VARIANT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
    cudaMemcpyToSymbolAsync(sc, &scene, sizeof(DScene), 0, cudaMemcpyHostToDevice, st);
    k_forward_pool<INST_SPEC, INST_MODE, INST_VARIANT><<<blocks, threads, smem, st>>>(lp);
}
#if INST_VARIANT == 0
template <> int pool_slots<INST_SPEC, INST_MODE>() { return PoolL<INST_SPEC>::P; }
#endif
#elif defined(INST_APOOL)
#ifndef INST_VARIANT
#define INST_VARIANT 0
#endif
template <> void launch_k_adjoint_pool<INST_SPEC, INST_VARIANT>(const DScene &scene, const LaunchParams &lp, const AdjParams &ap, unsigned blocks, unsigned threads, cudaStream_t st) {
    cudaMemcpyToSymbolAsync(sc, &scene, sizeof(DScene), 0, cudaMemcpyHostToDevice, st);
    if (ap.primal_only) {
        cudaFuncSetAttribute(k_adjoint_pool<INST_SPEC, INST_VARIANT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        k_adjoint_pool<INST_SPEC, INST_VARIANT, true><<<blocks, threads, apool_smem_bytes<INST_SPEC>(), st>>>(lp, ap);
    } else {
        cudaFuncSetAttribute(k_adjoint_pool<INST_SPEC, INST_VARIANT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        k_adjoint_pool<INST_SPEC, INST_VARIANT, false><<<blocks, threads, apool_smem_bytes<INST_SPEC>(), st>>>(lp, ap);
    }
}
#if INST_VARIANT == 0
template <> int adjoint_pool_slots<INST_SPEC>() { return APoolL<INST_SPEC>::P; }
#endif
#elif defined(INST_ADJ)
template <> void launch_k_adjoint<INST_SPEC>(const DScene &scene, const LaunchParams &lp, const AdjParams &ap, unsigned blocks, cudaStream_t st) {
    cudaMemcpyToSymbolAsync(sc, &scene, sizeof(DScene), 0, cudaMemcpyHostToDevice, st);
    k_adjoint<INST_SPEC><<<blocks, 128, 0, st>>>(lp, ap);
}
#endif
}

// fast_kernels.cuh -- specialised N = 2048 kernels (placeholder: not enabled yet; the generic kernels serve every set).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "generic_kernels.cuh"

namespace mkfhe {
struct FastState {};
inline bool fast_supported(const DevCtx&) { return false; }
inline int fast_init(FastState&, const DevCtx&, uint32_t, cudaStream_t, uint64_t*) { return 0; }
inline void fast_destroy(FastState&) {}
inline cudaError_t fast_load_bk(FastState&, const DevCtx&, cudaStream_t, uint64_t*, uint64_t*) { return cudaSuccess; }
inline cudaError_t fast_load_pkey(FastState&, const DevCtx&, cudaStream_t, uint64_t*, uint64_t*) { return cudaSuccess; }
inline cudaError_t fast_blind_rotate(FastState&, const DevCtx&, bool, const GateIO&, uint32_t, cudaStream_t, uint64_t*) { return cudaErrorNotSupported; }
inline cudaError_t fast_ntt(FastState&, const DevCtx&, uint32_t*, uint64_t, int, cudaStream_t, uint64_t*) { return cudaErrorNotSupported; }
}  // namespace mkfhe

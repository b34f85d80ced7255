// bfly_bench.cu -- micro-benchmark: can the FP64 pipe of sm_100a take the quotient estimate of the lazy butterfly off the
// fma-heavy pipe?  (profiles/r1_s35_mklwe4096.txt: the gate kernel is bound by that pipe -- IMAD.HI costs 4 of the 8 pipe
// cycles of a butterfly.)
//
//   Shoup      q = umulhi(v, w')                  t = v w - q Q in [0, 2Q)        IMAD.HI + 2 IMAD            (8 fma-heavy cycles)
//   FP64       d = (2^52 | v) - 2^52              q = low32(fma(d, w / Q, 2^52))  DADD + DFMA + 2 IMAD (+ MOV) (4 fma-heavy, 4 fp64)
//              round to nearest: t = v w - q Q in [-Q/2 - 1, Q/2 + 1]; the butterfly adds Q, so t + Q in [0, 2Q) like Shoup's
//
// MIX = how many of every 8 butterflies use the FP64 quotient.  Prints butterflies/s per MIX and checks that every variant
// leaves the same residues mod Q as MIX = 0.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -o tools/_build/bfly_bench tools/bfly_bench.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

struct Tw {
    uint32_t w, ws;  // twiddle, floor(w 2^32 / Q)
    double wq;       // w / Q
};

__device__ __forceinline__ uint32_t mul_shoup(uint32_t v, const Tw& t, uint32_t Q) {
    const uint32_t q = __umulhi(v, t.ws);
    return v * t.w - q * Q;
}
__device__ __forceinline__ uint32_t mul_fp64(uint32_t v, const Tw& t, uint32_t Q) {
    const double d = __hiloint2double(0x43300000, int(v)) - 4503599627370496.0;
    const double qd = fma(d, t.wq, 4503599627370496.0);
    const uint32_t q = uint32_t(__double2loint(qd));
    return v * t.w - q * Q + Q;  // the + Q folds into the IADD3 of the butterfly
}

template <int MIX>
__global__ void __launch_bounds__(256) bfly_kernel(uint32_t* out, uint32_t iters, uint32_t Q, uint32_t zero, uint32_t seed) {
    const uint32_t twoQ = 2 * Q;
    uint32_t x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = (threadIdx.x * 2654435761u + i * 40503u + seed) % Q;
    Tw w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t v = ((threadIdx.x + i + seed) * 2246822519u) % Q;
        w[i].w = v;
        w[i].ws = uint32_t((uint64_t(v) << 32) / Q);
        w[i].wq = double(v) / double(Q);
    }
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int st = 0; st < 4; ++st) {
            const int h = 8 >> st;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const int lo = (b / h) * 2 * h + (b % h);
                // spread the FP64 butterflies evenly over the 8
                const bool f = ((b + 1) * MIX) / 8 != (b * MIX) / 8;
                const uint32_t t = f ? mul_fp64(x[lo + h], w[st], Q) : mul_shoup(x[lo + h], w[st], Q);
                const uint32_t u = x[lo];
                x[lo] = u + t + zero;
                x[lo + h] = u + twoQ - t;
            }
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] = x[i] % Q * (zero + 1);  // canonical between rounds so that the variants can be compared
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) r = r * 31u + x[i] % Q;
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// same loop with the cheap bounding of the library's peak kernel (16 alu ops per 32 butterflies): this is the timed one
template <int MIX>
__global__ void __launch_bounds__(256) bfly_kernel_timed(uint32_t* out, uint32_t iters, uint32_t Q, uint32_t zero, uint32_t seed) {
    const uint32_t twoQ = 2 * Q;
    uint32_t x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = (threadIdx.x * 2654435761u + i * 40503u + seed) % Q;
    Tw w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t v = ((threadIdx.x + i + seed) * 2246822519u) % Q;
        w[i].w = v;
        w[i].ws = uint32_t((uint64_t(v) << 32) / Q);
        w[i].wq = double(v) / double(Q);
    }
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int st = 0; st < 4; ++st) {
            const int h = 8 >> st;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const int lo = (b / h) * 2 * h + (b % h);
                const bool f = ((b + 1) * MIX) / 8 != (b * MIX) / 8;
                const uint32_t t = f ? mul_fp64(x[lo + h], w[st], Q) : mul_shoup(x[lo + h], w[st], Q);
                const uint32_t u = x[lo];
                x[lo] = u + t + zero;
                x[lo + h] = u + twoQ - t;
            }
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] &= 0x07ffffffu;
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) r ^= x[i];
    if (r == 0x12345678u) out[0] = r;
}

template <int MIX>
static void run(uint32_t* d_out, uint32_t* h_ref, uint32_t* h_out, int sms, uint32_t Q) {
    const uint32_t blocks = sms * 4, threads = 256, iters = 4096;
    const size_t n = size_t(blocks) * threads;
    bfly_kernel<MIX><<<blocks, threads>>>(d_out, 16, Q, 0u, 3u);
    cudaMemcpy(h_out, d_out, n * 4, cudaMemcpyDeviceToHost);
    size_t bad = 0;
    if (MIX == 0) for (size_t i = 0; i < n; ++i) h_ref[i] = h_out[i];
    else for (size_t i = 0; i < n; ++i) bad += h_ref[i] != h_out[i];
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    bfly_kernel_timed<MIX><<<blocks, threads>>>(d_out, 64, Q, 0u, 1u);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        bfly_kernel_timed<MIX><<<blocks, threads>>>(d_out, iters, Q, 0u, 2u + rep);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const cudaError_t e = cudaGetLastError();
    printf("MIX %d/8 fp64: %.3f Tbfly/s  (%.3f ms)  mismatches vs Shoup: %zu  %s\n", MIX, double(n) * iters * 32 / (best * 1e-3) / 1e12,
           best, bad, e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const uint32_t Q = 134176769u;
    const size_t n = size_t(sms) * 4 * 256;
    uint32_t* d_out;
    cudaMalloc(&d_out, n * 4);
    uint32_t* h_ref = new uint32_t[n];
    uint32_t* h_out = new uint32_t[n];
    run<0>(d_out, h_ref, h_out, sms, Q);
    run<1>(d_out, h_ref, h_out, sms, Q);
    run<2>(d_out, h_ref, h_out, sms, Q);
    run<3>(d_out, h_ref, h_out, sms, Q);
    run<4>(d_out, h_ref, h_out, sms, Q);
    run<6>(d_out, h_ref, h_out, sms, Q);
    run<8>(d_out, h_ref, h_out, sms, Q);
    return 0;
}

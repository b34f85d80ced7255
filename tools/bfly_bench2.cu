// bfly_bench2.cu -- micro-benchmark, round 2: alternative instruction sequences for the lazy butterfly's t = y w mod Q in [0, 2Q).
// Round 1 (tools/bfly_bench.cu) measured the FP64 quotient as DADD + DFMA: flat at 4.2 Tbfly/s for every mix with Shoup's.  Here:
//
//   V0 Shoup     q = umulhi(y, w')                       t = y w - q Q          IMAD.HI + 2 IMAD                 8 fma-heavy cycles
//   V1 FP64 x2   d = (2^52 | y) - 2^52; q = lo(fma(d, w/Q, 2^52))               DADD + DFMA + 2 IMAD (+ MOV)
//   V2 FP64 x1   q = lo(fma_rz(2^52 | y, W, C)),  W = floor(w 2^52 / Q) / 2^52, C = 2^52 (1 - W): the exact sum is
//                2^52 + y W, rounded toward zero to an integer: q in {floor(y w / Q) - 1, floor(y w / Q)} as Shoup's
//                                                                              DFMA + 2 IMAD (+ MOV)            4 fma-heavy cycles
//   V3 Shoup, q Q by shifts: Q = 2^27 - 2^15 - 2^13 + 1 -> -q Q = ((5 q) << 13) - q - (q << 27)   IMAD.HI + IMAD + 3 alu
//   V4 V2 with q Q by shifts                                                    DFMA + IMAD + 3 alu (+ MOV)      2 fma-heavy cycles
//
// MIX = how many of every 8 butterflies use the variant (the rest use Shoup).  Checks residues mod Q against V0.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -o /tmp/bfly_bench2 tools/bfly_bench2.cu && /tmp/bfly_bench2
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

struct Tw {
    uint32_t w, ws;  // twiddle, floor(w 2^32 / Q)
    double wq;       // w / Q (V1)
    double W, C;     // V2
};

__device__ __forceinline__ uint32_t mul_shoup(uint32_t v, const Tw& t, uint32_t Q) {
    const uint32_t q = __umulhi(v, t.ws);
    return v * t.w - q * Q;
}
__device__ __forceinline__ uint32_t mul_fp64x2(uint32_t v, const Tw& t, uint32_t Q) {
    const double d = __hiloint2double(0x43300000, int(v)) - 4503599627370496.0;
    const double qd = fma(d, t.wq, 4503599627370496.0);
    const uint32_t q = uint32_t(__double2loint(qd));
    return v * t.w - q * Q + Q;
}
__device__ __forceinline__ uint32_t quot_fp64x1(uint32_t v, const Tw& t) {
    const double qd = __fma_rz(__hiloint2double(0x43300000, int(v)), t.W, t.C);
    return uint32_t(__double2loint(qd));
}
__device__ __forceinline__ uint32_t mul_fp64x1(uint32_t v, const Tw& t, uint32_t Q) { return v * t.w - quot_fp64x1(v, t) * Q; }
// -q Q for Q = 2^27 - 2^15 - 2^13 + 1
__device__ __forceinline__ uint32_t neg_qQ(uint32_t q) {
    const uint32_t q5 = q + (q << 2);
    return (q5 << 13) - q - (q << 27);
}
__device__ __forceinline__ uint32_t mul_shoup_sh(uint32_t v, const Tw& t, uint32_t) { return v * t.w + neg_qQ(__umulhi(v, t.ws)); }
__device__ __forceinline__ uint32_t mul_fp64x1_sh(uint32_t v, const Tw& t, uint32_t) { return v * t.w + neg_qQ(quot_fp64x1(v, t)); }

template <int V>
__device__ __forceinline__ uint32_t mul_var(uint32_t v, const Tw& t, uint32_t Q) {
    if (V == 1) return mul_fp64x2(v, t, Q);
    if (V == 2) return mul_fp64x1(v, t, Q);
    if (V == 3) return mul_shoup_sh(v, t, Q);
    if (V == 4) return mul_fp64x1_sh(v, t, Q);
    return mul_shoup(v, t, Q);
}

__device__ __forceinline__ void make_tw(Tw& t, uint32_t v, uint32_t Q) {
    t.w = v;
    t.ws = uint32_t((uint64_t(v) << 32) / Q);
    t.wq = double(v) / double(Q);
    // floor(v 2^52 / Q) by long division in two 26-bit steps (v < Q < 2^27)
    const uint64_t hi = (uint64_t(v) << 26) / Q, r = (uint64_t(v) << 26) % Q;
    const uint64_t Wi = (hi << 26) + ((r << 26) / Q);
    t.W = double(Wi) * (1.0 / 4503599627370496.0);
    t.C = double((1ull << 52) - Wi);
}

template <int V, int MIX, bool TIMED>
__global__ void __launch_bounds__(256) bfly_kernel(uint32_t* out, uint32_t iters, uint32_t Q, uint32_t zero, uint32_t seed) {
    const uint32_t twoQ = 2 * Q;
    uint32_t x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = (threadIdx.x * 2654435761u + i * 40503u + seed) % Q;
    Tw w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) make_tw(w[i], ((threadIdx.x + i + seed) * 2246822519u) % Q, Q);
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int st = 0; st < 4; ++st) {
            const int h = 8 >> st;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const int lo = (b / h) * 2 * h + (b % h);
                const bool f = ((b + 1) * MIX) / 8 != (b * MIX) / 8;
                const uint32_t t = f ? mul_var<V>(x[lo + h], w[st], Q) : mul_shoup(x[lo + h], w[st], Q);
                const uint32_t u = x[lo];
                x[lo] = u + t + zero;
                x[lo + h] = u + twoQ - t;
            }
        }
        if (TIMED) {
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] &= 0x07ffffffu;
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = x[i] % Q * (zero + 1);
        }
    }
    if (TIMED) {
        uint32_t r = 0;
#pragma unroll
        for (int i = 0; i < 16; ++i) r ^= x[i];
        if (r == 0x12345678u) out[0] = r;
    } else {
        uint32_t r = 0;
#pragma unroll
        for (int i = 0; i < 16; ++i) r = r * 31u + x[i] % Q;
        out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    }
}

static uint32_t* g_ref = nullptr;

template <int V, int MIX>
static void run(uint32_t* d_out, uint32_t* h_out, int sms, uint32_t Q, int ctas_per_sm) {
    const uint32_t blocks = sms * ctas_per_sm, threads = 256, iters = 4096;
    const size_t n = size_t(sms) * 4 * threads;
    bfly_kernel<V, MIX, false><<<sms * 4, threads>>>(d_out, 16, Q, 0u, 3u);
    cudaMemcpy(h_out, d_out, n * 4, cudaMemcpyDeviceToHost);
    size_t bad = 0;
    if (!g_ref) {
        g_ref = new uint32_t[n];
        for (size_t i = 0; i < n; ++i) g_ref[i] = h_out[i];
    } else {
        for (size_t i = 0; i < n; ++i) bad += g_ref[i] != h_out[i];
    }
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    bfly_kernel<V, MIX, true><<<blocks, threads>>>(d_out, 64, Q, 0u, 1u);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        bfly_kernel<V, MIX, true><<<blocks, threads>>>(d_out, iters, Q, 0u, 2u + rep);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const cudaError_t e = cudaGetLastError();
    const double bf = double(blocks) * threads * iters * 32;
    // SMSP-cycles per warp-butterfly at 1965 MHz: 148 SMs x 4 SMSPs x f / (bf / 32 per s)
    const double rate = bf / (best * 1e-3);
    printf("V%d MIX %d/8 @%d CTAs/SM: %.3f Tbfly/s  (%.3f ms)  %.2f SMSP-cycles per warp-butterfly  mismatches vs Shoup: %zu  %s\n", V, MIX,
           ctas_per_sm, rate / 1e12, best, double(sms) * 4 * 1.965e9 / (rate / 32), bad, e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const uint32_t Q = 134176769u;
    const size_t n = size_t(sms) * 4 * 256;
    uint32_t* d_out;
    cudaMalloc(&d_out, n * 4 * 2);
    uint32_t* h_out = new uint32_t[n];
    for (int c = 4; c <= 8; c += 4) {
        run<0, 0>(d_out, h_out, sms, Q, c);
        run<1, 4>(d_out, h_out, sms, Q, c);
        run<1, 8>(d_out, h_out, sms, Q, c);
        run<2, 2>(d_out, h_out, sms, Q, c);
        run<2, 4>(d_out, h_out, sms, Q, c);
        run<2, 6>(d_out, h_out, sms, Q, c);
        run<2, 8>(d_out, h_out, sms, Q, c);
        run<3, 4>(d_out, h_out, sms, Q, c);
        run<3, 8>(d_out, h_out, sms, Q, c);
        run<4, 4>(d_out, h_out, sms, Q, c);
        run<4, 8>(d_out, h_out, sms, Q, c);
    }
    return 0;
}

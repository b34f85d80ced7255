// exchange_microbench.cu -- the intra-warp exchange of the register-resident transforms, two ways (VERDICT r1, item 4b:
// north_star names "warp-shuffle butterflies"; the kernels exchange through padded shared memory + __syncwarp).
// The exchange between forward passes 2 and 3 (fast_kernels.cuh, patterns B -> C) is an 8 x 8 transpose among the 8 lanes
// that share lane & 3:   new x[r] of lane L  =  old x[L >> 2] of lane (r << 2 | (L & 3)).
//   smem : 8 STS + 8 LDS through a padded per-warp buffer (bank-conflict free), 2 __syncwarp        (what the kernels do)
//   shfl : 3 rounds of a butterfly transpose: 4 SHFL.BFLY + 12 SEL per round = 12 SHFL + 36 SEL     (the alternative)
// Both are timed alone and between two 3-stage butterfly passes (24 lazy Shoup butterflies per exchange: the real duty cycle),
// 32 warps per SM, results checked equal.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/exchange_microbench tools/exchange_microbench.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#define Q 134176769u

__device__ __forceinline__ void xchg_smem(uint32_t (&x)[8], uint32_t* E, uint32_t lane) {
    // warp-private buffer of 8 rows of 32 + 4 pad words: write pattern B (row r, column lane), read pattern C
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 8; ++r) E[36 * r + lane] = x[r];
    __syncwarp();
    const uint32_t base = 36 * (lane >> 2) + (lane & 3);
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = E[base + 4 * r];
}

__device__ __forceinline__ void xchg_shfl(uint32_t (&x)[8], uint32_t lane) {
#pragma unroll
    for (int b = 0; b < 3; ++b) {
        const bool up = (lane >> (2 + b)) & 1;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r & (1 << b)) continue;
            const int r2 = r | (1 << b);
            uint32_t v = up ? x[r] : x[r2];
            v = __shfl_xor_sync(0xffffffffu, v, 4 << b);
            if (up) x[r] = v; else x[r2] = v;
        }
    }
}

__device__ __forceinline__ void bfly(uint32_t& a, uint32_t& b, uint32_t w, uint32_t wp, uint32_t zero) {
    const uint32_t q = __umulhi(b, wp), t = b * w - q * Q, u = a;
    a = u + t + zero;
    b = u + 2 * Q - t;
}

template <int MODE, bool WORK>  // MODE 0 smem, 1 shfl
__global__ void __launch_bounds__(256) run(uint32_t* out, uint32_t iters, uint32_t w, uint32_t wp, uint32_t zero) {
    __shared__ uint32_t Es[8][8 * 36];
    const uint32_t lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    uint32_t x[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = (threadIdx.x * 2654435761u + r * 40503u + blockIdx.x) % Q;
    for (uint32_t it = 0; it < iters; ++it) {
        if (WORK) {
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int h = 4 >> s;
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    if (!(r & h)) bfly(x[r], x[r + h], w, wp, zero);
            }
#pragma unroll
            for (int r = 0; r < 8; ++r) x[r] &= 0x07ffffffu;
        }
        if (MODE == 0) xchg_smem(x, Es[wrp], lane);
        else xchg_shfl(x, lane);
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r = r * 31 + x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE, bool WORK>
static float timeit(uint32_t* out, int blocks, uint32_t iters) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    run<MODE, WORK><<<blocks, 256>>>(out, 16, 12345u, uint32_t((12345ull << 32) / Q), 0u);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(a);
        run<MODE, WORK><<<blocks, 256>>>(out, iters, 12345u, uint32_t((12345ull << 32) / Q), 0u);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int blocks = sms * 4;  // 32 warps per SM
    const uint32_t iters = 20000;
    uint32_t *o0, *o1;
    cudaMalloc(&o0, blocks * 256 * 4);
    cudaMalloc(&o1, blocks * 256 * 4);
    // same result
    run<0, true><<<blocks, 256>>>(o0, 7, 12345u, uint32_t((12345ull << 32) / Q), 0u);
    run<1, true><<<blocks, 256>>>(o1, 7, 12345u, uint32_t((12345ull << 32) / Q), 0u);
    static uint32_t h0[148 * 4 * 256 * 2], h1[148 * 4 * 256 * 2];
    cudaMemcpy(h0, o0, blocks * 256 * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(h1, o1, blocks * 256 * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int i = 0; i < blocks * 256; ++i) bad += h0[i] != h1[i];
    const double warps = double(blocks) * 8, clk = 1.965e9;
    auto rep = [&](const char* name, float ms, float base) {
        const double per = (ms - base) * 1e-3 * clk / iters / (warps / (sms * 4.0));  // SMSP cycles per warp-exchange
        printf("%-34s %8.3f ms   %6.1f SMSP-cycles per warp-exchange (above the baseline)\n", name, ms, per);
    };
    printf("{\"sms\": %d, \"mismatches\": %d}\n", sms, bad);
    const float s0 = timeit<0, false>(o0, blocks, iters), s1 = timeit<1, false>(o1, blocks, iters);
    rep("exchange alone, shared memory", s0, 0);
    rep("exchange alone, shuffles", s1, 0);
    const float w0 = timeit<0, true>(o0, blocks, iters), w1 = timeit<1, true>(o1, blocks, iters);
    printf("24 butterflies + exchange, shared memory : %8.3f ms\n", w0);
    printf("24 butterflies + exchange, shuffles      : %8.3f ms   (%+.1f%%)\n", w1, 100.0 * (w1 - w0) / w0);
    return bad != 0;
}

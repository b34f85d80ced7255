// pipe_microbench.cu -- measures the integer-pipe rates that bound the NTT phase on sm_100a (B200):
// IMAD.LO / IMAD.WIDE / IMAD.HI / IADD3+LOP3 / DFMA issue rates, and whole lazy butterflies (Montgomery-wide and Shoup-hi
// formulations) on register-resident data.  Output feeds DESIGN.md's choice of modular multiplication and the integer
// roofline denominator of bench.py.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/pipe_microbench tools/pipe_microbench.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

#define Q 134176769u
#define QINV_NEG 0u  // filled at run time through a kernel argument so nothing constant-folds

template <int MODE>
__global__ void __launch_bounds__(256) chain(uint32_t* sink, uint32_t iters, uint32_t m, uint32_t c) {
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 2654435761u + i * 40503u + c;
    unsigned long long w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = a[i];
    double d[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) d[i] = 1.0 + a[i] * 1e-9;
    const double dm = 1.0 + m * 1e-12, dc = c * 1e-7;
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 16; ++r) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (MODE == 0) a[i] = a[i] * m + c;                                       // IMAD (lo)
                if (MODE == 1) w[i] = (unsigned long long)(uint32_t)w[i] * m + w[i];      // IMAD.WIDE with 64-bit addend
                if (MODE == 2) a[i] = __umulhi(a[i], m) + c;                              // IMAD.HI
                if (MODE == 3) a[i] = (a[i] + c) ^ (a[i] >> 3);                            // alu pipe: IADD3 / SHF / LOP3
                if (MODE == 4) d[i] = fma(d[i], dm, dc);                                  // DFMA
                if (MODE == 5) w[i] = (unsigned long long)(uint32_t)w[i] * m;             // IMAD.WIDE, no addend
            }
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r ^= a[i] ^ uint32_t(w[i]) ^ uint32_t(w[i] >> 32) ^ uint32_t(__double2loint(d[i]));
    if (r == 0x12345678u) sink[0] = r;
}

// whole lazy butterflies on 16 register values: x' = x + t, y' = x + 2Q - t, t = y*w*2^-32 mod Q in [0, 2Q)
// MODE 0: Montgomery (IMAD.WIDE, IMAD.LO, IMAD.WIDE.hi)   MODE 1: Shoup (IMAD.HI, IMAD.LO, IMAD.LO)
template <int MODE>
__global__ void __launch_bounds__(128) bfly(uint32_t* sink, uint32_t iters, uint32_t qq, uint32_t qinv_neg, const uint32_t* tw) {
    uint32_t x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = (threadIdx.x * 2654435761u + i * 40503u) % qq;
    const uint32_t twoq = 2 * qq;
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int st = 0; st < 4; ++st) {
            const int h = 8 >> st;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const int lo = (b / h) * 2 * h + (b % h), hi = lo + h;
                const uint32_t w = tw[(it + st * 8 + b) & 63];
                uint32_t t;
                if (MODE == 0) {
                    const unsigned long long p = (unsigned long long)x[hi] * w;
                    const uint32_t k = uint32_t(p) * qinv_neg;
                    t = uint32_t((p + (unsigned long long)k * qq) >> 32);
                } else {
                    const uint32_t wp = tw[64 + ((it + st * 8 + b) & 63)];
                    const uint32_t qh = __umulhi(x[hi], wp);
                    t = x[hi] * w - qh * qq;
                }
                const uint32_t xl = x[lo];
                x[lo] = xl + t;
                x[hi] = xl + twoq - t;
            }
        }
        // keep values bounded: conditional-free fold (cheap, on the alu pipe)
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] = x[i] & 0x07ffffffu;
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) r ^= x[i];
    if (r == 0x12345678u) sink[0] = r;
}

template <class F>
static float time_ms(F f) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    f();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        f();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    int sms = 0, clk = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    uint32_t* sink;
    cudaMalloc(&sink, 4);
    uint32_t htw[128];
    for (int i = 0; i < 128; ++i) htw[i] = (i * 2654435761u) % Q;
    uint32_t* tw;
    cudaMalloc(&tw, sizeof htw);
    cudaMemcpy(tw, htw, sizeof htw, cudaMemcpyHostToDevice);
    uint32_t inv = 1;
    for (int i = 0; i < 5; ++i) inv *= 2u - Q * inv;
    const uint32_t iters = 2048;
    const char* names[] = {"IMAD.LO", "IMAD.WIDE+add64", "IMAD.HI", "ALU (IADD3+SHF+LOP3 = 3 ops)", "DFMA", "IMAD.WIDE"};
    printf("{\"sms\": %d, \"clock_khz\": %d", sms, clk);
    for (int occ = 0; occ < 2; ++occ) {
        const int blocks = sms * (occ ? 8 : 4), threads = 256;
        for (int mode = 0; mode < 6; ++mode) {
            float ms = 0;
            switch (mode) {
                case 0: ms = time_ms([&] { chain<0><<<blocks, threads>>>(sink, iters, 7u, 3u); }); break;
                case 1: ms = time_ms([&] { chain<1><<<blocks, threads>>>(sink, iters, 7u, 3u); }); break;
                case 2: ms = time_ms([&] { chain<2><<<blocks, threads>>>(sink, iters, 0xfffffff7u, 3u); }); break;
                case 3: ms = time_ms([&] { chain<3><<<blocks, threads>>>(sink, iters, 7u, 3u); }); break;
                case 4: ms = time_ms([&] { chain<4><<<blocks, threads>>>(sink, iters, 7u, 3u); }); break;
                case 5: ms = time_ms([&] { chain<5><<<blocks, threads>>>(sink, iters, 7u, 3u); }); break;
            }
            const double ops = double(blocks) * threads * iters * 16.0 * 8.0;
            printf(",\n \"%s @%d warps/SM\": {\"ms\": %.3f, \"Gop_s\": %.1f, \"per_clk_per_sm_at_1965\": %.1f}", names[mode],
                   (occ ? 64 : 32), ms, ops / ms / 1e6, ops / (ms * 1e-3) / sms / 1.965e9);
        }
    }
    for (int mode = 0; mode < 2; ++mode) {
        for (int bps = 2; bps <= 8; bps *= 2) {
            const int blocks = sms * bps;
            float ms = mode == 0 ? time_ms([&] { bfly<0><<<blocks, 128>>>(sink, iters, Q, 0u - inv, tw); })
                                 : time_ms([&] { bfly<1><<<blocks, 128>>>(sink, iters, Q, 0u - inv, tw); });
            const double bf = double(blocks) * 128 * iters * 32.0;
            printf(",\n \"butterfly %s @%d warps/SM\": {\"ms\": %.3f, \"Gbfly_s\": %.1f, \"clk_per_bfly_per_sm_x64\": %.2f}",
                   mode == 0 ? "montgomery-wide" : "shoup-hi", bps * 4, ms, bf / ms / 1e6,
                   64.0 / (bf / (ms * 1e-3) / sms / 1.965e9));
        }
    }
    printf("}\n");
    return 0;
}

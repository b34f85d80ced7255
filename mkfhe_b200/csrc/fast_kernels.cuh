// fast_kernels.cuh -- specialised kernels for the production shape N = 2048, digitsG - 1 = 2 (every STD100_* set and the
// STD128_*_LWE sets of binfhecontext.cpp:91-146), any party count whose accumulator fits shared memory.
//
// One CTA of 256 threads owns one gate.  A size-2048 transform is four register passes of 3 + 3 + 3 + 2 radix-2 stages
// with 8 coefficients per thread; between passes the coefficients are transposed through shared memory: one exchange
// crosses warps (block barrier), the other two stay inside a warp (warp w owns coefficients [256 w, 256 w + 256) on both
// sides, so __syncwarp suffices).  Padding (4 words per 32) makes every exchange access bank-conflict free.  Butterflies are
// lazy: no value is reduced below 2^32 until the end, which Q < 2^32 / 32 allows (bounds next to each routine).
// (A first version used 128 threads x 16 coefficients in three passes: 128 registers per thread, 4 warps per gate, a
// 100 KB instruction footprint -- ncu showed 16% of stalls waiting for instructions and only 3.5 warps per scheduler;
// profiles/r1_fast_v1_blind_rotate.txt.  This layout halves the unrolled code and doubles the warps.)
//
// Modular multiplication by a known constant (twiddles) is Shoup's: q = mulhi(y, w'), t = y*w - q*Q in [0, 2Q), with
// w' = floor(w * 2^32 / Q) -- one IMAD.HI (half rate on sm_100a, measured by tools/pipe_microbench.cu) and two IMAD.LO, one
// slot fewer on the fma pipe than Montgomery's two IMAD.WIDE.  Key products use Montgomery accumulation in 64 bits, one
// reduction per output (keys are stored in Montgomery form at upload).
//
// What is computed is the reference's HbProd recurrence (mk-acc-xzw.cpp:231-345, mk-acc-xzw_B.cpp:221-331) restated in the
// coefficient domain, exactly as generic_kernels.cuh does; only the schedule differs:
//   iteration it = 0..k-1 : party p_it (key owner last): t = (X^c - 1) a_p -> digits -> 2 forward NTTs -> X = <D, d>,
//                           V += <D, Pkey_p>;  then one inverse NTT: of X (non-owner; a_p += result) or of V (owner)
//   iteration it = k      : digits of INTT(V) -> 2 forward NTTs -> X_owner += <D, f>; inverse NTT; a_owner += result
// so every iteration is "2 forward + 1 inverse" and the code holds one instance of each.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>
#include <vector>

#include "generic_kernels.cuh"

namespace mkfhe {

constexpr uint32_t FN = 2048, FLOGN = 11, FT = 256, FPT = 8;  // ring size, threads per gate, coefficients per thread

struct FastTables {
    const uint2* tw2;     // [N]  forward twiddles, OpenFHE order: psi^i at bitrev(i), as Shoup pairs (w, floor(w 2^32 / Q))
    const uint2* itw2;    // [N]  cyclic inverse (DIT) twiddles: itw2[t + m] = omega^(-m N / 2t), m < t, t = 1, 2, .., N/2
    const uint2* scale2;  // [N]  psi^-i / N  (the twist and the 1/N of the inverse, folded into one Shoup multiply)
    const uint4* tw16;    // [7][128] the 14 twiddles of thread t in the last forward pass of the 16-coefficient transform:
                          // word 0: tw2[256+2t..], words 1-2: tw2[512+4t..], words 3-6: tw2[1024+8t..], two twiddles per word
    const uint4* tw4;     // [3][256] the six twiddles of thread t in the last forward pass, as three 128-bit words:
                          // (tw2[512+2t], tw2[513+2t]) | (tw2[1024+4t], tw2[1025+4t]) | (tw2[1026+4t], tw2[1027+4t]) -- one
                          // coalesced load each (16-byte lane stride) instead of six 8-byte loads at 16/32-byte stride
    uint2 twA[8];         // tw2[0..7]: the first forward pass uses the same twiddles in every thread -> kernel-parameter constants
    uint2 itwA[4];        // itw2[0..3]
    uint2 twB[64];        // tw2[0..63]: the second forward pass indexes them by warp number (uniform: constant-bank loads, off the LSU)
    uint2 itwB[32];       // itw2[0..31]: the second inverse pass (4 distinct addresses per warp)
    uint32_t zero;        // 0, opaque to the compiler: makes "u + t" a three-input add, which ptxas maps to IADD3 (alu pipe)
                          // instead of IMAD.IADD (fma pipe, the saturated one: profiles/r1_fast_v2_blind_rotate.txt)
    uint32_t prefetch;    // 1: every step asks L2 for the NEXT step's key tile (prefetch.global.L2, one 128-byte line per thread, no
                          // registers): with few gates in flight the tile comes from HBM and the step would wait for it
    uint32_t pair;        // 1: the 80- / 128-register builds take inverse transforms two at a time (fast_step PAIR); MKFHE_PAIR=0/1
    uint32_t stagger_ns;  // experiment (MKFHE_STAGGER_NS): CTA b of the gate kernel starts (b mod 4) * stagger_ns late, so that the
                          // resident gates of an SM run out of phase; 0 = off
};

// y * w mod Q in [0, 2Q) for ANY 32-bit y (Harvey's lazy Shoup multiplication)
__device__ __forceinline__ uint32_t shoup_lazy(uint32_t y, uint2 w, uint32_t Q) {
    const uint32_t q = __umulhi(y, w.y);
    return y * w.x - q * Q;
}
// [0, 2Q) -> [0, Q)
__device__ __forceinline__ uint32_t fold1(uint32_t x, uint32_t Q) { return min(x, x - Q); }

// Position of coefficient j (11 bits) in an exchange buffer: 4 words of padding after every 32 (pad(j) = j + 4 (j >> 5)), so
// a polynomial takes FNP = 2304 words.  Every access pattern of the transforms is then "per-thread base + compile-time
// constant" (one LDS/STS with an immediate offset, no address arithmetic in the unrolled code), 16-byte groups stay
// aligned, and each pattern is bank-conflict free:
//   A  j = tid + 256 r          -> (tid + 4 (tid >> 5)) + 288 r      consecutive lanes, consecutive words
//   B  j = 256 w + 32 r + lane  -> (288 w + lane) + 36 r             consecutive lanes, consecutive words
//   C  j = 32 h + 4 r + l2      -> (36 h + l2) + 4 r                 h = tid >> 2, l2 = tid & 3: bank = 4 (h mod 8) + l2
//   D  j = 8 tid + 4 c (128-bit)-> (8 tid + 4 h) + 4 c               a quarter warp covers 8 distinct 16-byte bank groups
// (The first version XOR-swizzled bits 2-4 with bits 5-7: same conflict freedom in 2048 words, but patterns B, C and D
// cost one LOP3 per access -- 64 per iteration of the step loop, 4% of its instructions; profiles/r1_s14_*.)
constexpr uint32_t FNP = FN + FN / 8;
struct ExIdx {
    uint32_t a, b, c, d;
    __device__ __forceinline__ explicit ExIdx(uint32_t tid)
        : a(tid + 4 * (tid >> 5)), b(288 * (tid >> 5) + (tid & 31)), c(36 * (tid >> 2) + (tid & 3)), d(8 * tid + 4 * (tid >> 2)) {}
    __device__ __forceinline__ uint32_t A(int r) const { return a + 288 * r; }
    __device__ __forceinline__ uint32_t B(int r) const { return b + 36 * r; }
    __device__ __forceinline__ uint32_t C(int r) const { return c + 4 * r; }
    __device__ __forceinline__ uint32_t D(int ch) const { return d + 4 * ch; }
};

// barrier over the threads that share one gate's transform: the whole CTA (bar = 0), or one 256-thread group of a
// party-parallel CTA (named barrier bar = 1 + group)
template <bool GRP>
__device__ __forceinline__ void gate_sync(uint32_t bar) {
    if (GRP) asm volatile("bar.sync %0, 256;" ::"r"(bar) : "memory");
    else __syncthreads();
}

#define MKFHE_BFLY(X0, X1, W)                          \
    {                                                  \
        const uint32_t t_ = shoup_lazy(X1, W, Q), u_ = X0; \
        X0 = u_ + t_ + T.zero;                         \
        X1 = u_ + twoQ - t_;                           \
    }

// ---- forward negacyclic NTT, NP polynomials at once (shared twiddles), Cooley-Tukey, natural -> bit-reversed --------
// in : x[p][r] = coefficient tid + 256 r          (values < 2Q)
// out: x[p][i] = slot 8 tid + i (OpenFHE order)    (values < 24Q: +2Q per stage, 11 stages)
// E: NP * FNP words of shared memory.  Contains one __syncthreads (after the pass-1 writes); the caller guarantees that
// every thread is done reading E from its previous use before any thread enters (a barrier lies in between).
// HALF: the 2048 words are TWO polynomials of 1024 coefficients ([0, 1024) and [1024, 2048)).  A size-2048 Cooley-Tukey
// transform is one stage (t = 1024) followed by two independent size-1024 transforms of the halves, so skipping that stage
// and indexing the twiddles per half (the table built by fast_init for N = 1024: entry m + b = twiddle of block b mod m/2
// of the size-1024 stage with m/2 blocks) transforms both polynomials with the same code, layouts and exchanges.
template <int NP, bool GRP = false, bool HALF = false>
__device__ __forceinline__ void fwd_ntt_regs(uint32_t (&x)[NP][8], uint32_t* E, const FastTables& T, uint32_t Q, uint32_t tid,
                                             uint32_t bar = 0) {
    const uint32_t twoQ = 2 * Q;
    const uint32_t wrp = tid >> 5, hi6 = tid >> 2;
    const ExIdx ex(tid);
    // pass 1: t = 1024, 512, 256  (m = 1, 2, 4 blocks; block of coefficient tid + 256 r at stage s is r >> (3 - s))
#pragma unroll
    for (int s = HALF ? 1 : 0; s < 3; ++s) {
        const int h = 4 >> s;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r & h) continue;
            const uint2 w = T.twA[(1 << s) + (r >> (3 - s))];
#pragma unroll
            for (int p = 0; p < NP; ++p) MKFHE_BFLY(x[p][r], x[p][r + h], w);
        }
    }
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < 8; ++r) E[p * FNP + ex.A(r)] = x[p][r];
    gate_sync<GRP>(bar);
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < 8; ++r) x[p][r] = E[p * FNP + ex.B(r)];
    // pass 2: t = 128, 64, 32  (m = 8, 16, 32; block = (warp << s) + (r >> (3 - s)): the same for all lanes of a warp)
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const int h = 4 >> s;
        const uint32_t twi = (8 << s) + (wrp << s);
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r & h) continue;
            const uint2 w = T.twB[twi + (r >> (3 - s))];
#pragma unroll
            for (int p = 0; p < NP; ++p) MKFHE_BFLY(x[p][r], x[p][r + h], w);
        }
    }
    // the remaining exchanges stay inside the warp: warp w holds coefficients [256 w, 256 w + 256) before and after
    __syncwarp();
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < 8; ++r) E[p * FNP + ex.B(r)] = x[p][r];
    __syncwarp();
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < 8; ++r) x[p][r] = E[p * FNP + ex.C(r)];
    // pass 3: t = 16, 8, 4  (m = 64, 128, 256; block = (hi6 << s) + (r >> (3 - s)))
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const int h = 4 >> s;
        const uint2* tws = T.tw2 + (64 << s) + (hi6 << s);
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r & h) continue;
            const uint2 w = __ldg(tws + (r >> (3 - s)));
#pragma unroll
            for (int p = 0; p < NP; ++p) MKFHE_BFLY(x[p][r], x[p][r + h], w);
        }
    }
    __syncwarp();
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < 8; ++r) E[p * FNP + ex.C(r)] = x[p][r];
    __syncwarp();
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const uint4 v = *reinterpret_cast<const uint4*>(E + p * FNP + ex.D(c));
            x[p][4 * c + 0] = v.x; x[p][4 * c + 1] = v.y; x[p][4 * c + 2] = v.z; x[p][4 * c + 3] = v.w;
        }
    // pass 4: t = 2, 1  (m = 512, 1024; block = (tid << (s + 1)) + (i >> (2 - s)))
    {
        const uint4 wa = __ldg(T.tw4 + tid), wb = __ldg(T.tw4 + 256 + tid), wc = __ldg(T.tw4 + 512 + tid);
        const uint2 w0[2] = {make_uint2(wa.x, wa.y), make_uint2(wa.z, wa.w)};
        const uint2 w1[4] = {make_uint2(wb.x, wb.y), make_uint2(wb.z, wb.w), make_uint2(wc.x, wc.y), make_uint2(wc.z, wc.w)};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (i & 2) continue;
#pragma unroll
            for (int p = 0; p < NP; ++p) MKFHE_BFLY(x[p][i], x[p][i + 2], w0[i >> 2]);
        }
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
#pragma unroll
            for (int p = 0; p < NP; ++p) MKFHE_BFLY(x[p][i], x[p][i + 1], w1[i >> 1]);
        }
    }
}

// ---- inverse negacyclic NTT of one polynomial, bit-reversed -> natural -------------------------------------------
// Written as the cyclic decimation-in-time transform with omega^-1 (Cooley-Tukey butterflies, so growth is additive),
// followed by one multiplication by psi^-i / N -- the same ring map as the reference's Gentleman-Sande inverse
// (transformnat-impl.h:491-552), hence the same canonical residues.
// in : x[i] = slot 8 tid + i, values < 6Q
// out: x[r] = coefficient tid + 256 r, values in [0, 2Q)
// bounds: first stage (twiddle 1, no multiplication) < 12Q, then +2Q per stage: < 32Q < 2^32.
// E: FNP words.  One __syncthreads inside (after the pass-3 writes); same entry condition as fwd_ntt_regs.
template <bool GRP = false, bool HALF = false>
__device__ __forceinline__ void inv_ntt_regs(uint32_t (&x)[8], uint32_t* E, const FastTables& T, uint32_t Q, uint32_t tid,
                                             uint32_t bar = 0) {
    const uint32_t twoQ = 2 * Q, sixQ = 6 * Q;
    const uint32_t lane = tid & 31, lo2 = tid & 3;
    const ExIdx ex(tid);
    // pass 1: t = 1 (twiddle 1), t = 2
#pragma unroll
    for (int i = 0; i < 8; i += 2) {
        const uint32_t u = x[i], v = x[i + 1];
        x[i] = u + v;
        x[i + 1] = u + sixQ - v;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (i & 2) continue;
        MKFHE_BFLY(x[i], x[i + 2], T.itwA[2 + (i & 1)]);
    }
    __syncwarp();
#pragma unroll
    for (int c = 0; c < 2; ++c)
        *reinterpret_cast<uint4*>(E + ex.D(c)) = make_uint4(x[4 * c], x[4 * c + 1], x[4 * c + 2], x[4 * c + 3]);
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = E[ex.C(r)];
    // pass 2: t = 4, 8, 16  (h = t / 4 in register units; m = 4 (r mod h) + lo2)
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const int h = 1 << s;
        const uint32_t twi = 4 * h + lo2;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r & h) continue;
            const uint2 w = T.itwB[twi + 4 * (r & (h - 1))];
            MKFHE_BFLY(x[r], x[r + h], w);
        }
    }
    // same positions written as read: no hazard inside the warp beyond the one with the next reads
#pragma unroll
    for (int r = 0; r < 8; ++r) E[ex.C(r)] = x[r];
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = E[ex.B(r)];
    // pass 3: t = 32, 64, 128  (h = t / 32; m = 32 (r mod h) + lane)
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const int h = 1 << s;
        const uint2* tws = T.itw2 + 32 * h + lane;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r & h) continue;
            const uint2 w = __ldg(tws + 32 * (r & (h - 1)));
            MKFHE_BFLY(x[r], x[r + h], w);
        }
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) E[ex.B(r)] = x[r];
    gate_sync<GRP>(bar);
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = E[ex.A(r)];
    // pass 4: t = 256, 512, 1024  (h = t / 256; m = 256 (r mod h) + tid); HALF: the halves stay separate (no t = 1024)
#pragma unroll
    for (int s = 0; s < (HALF ? 2 : 3); ++s) {
        const int h = 1 << s;
        const uint2* tws = T.itw2 + 256 * h + tid;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r & h) continue;
            const uint2 w = __ldg(tws + 256 * (r & (h - 1)));
            MKFHE_BFLY(x[r], x[r + h], w);
        }
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) x[r] = shoup_lazy(x[r], __ldg(T.scale2 + tid + 256 * r), Q);
}

// The same inverse transform on NP polynomials at once (shared twiddle loads: the stand-alone transform kernel is short of
// LSU wavefronts, and twiddles are a quarter of them).  E: NP * FNP words.  The gate kernels keep the one-polynomial form above.
template <int NP, bool HALF = false>
__device__ __forceinline__ void inv_ntt_regs_np(uint32_t (&x)[NP][8], uint32_t* E, const FastTables& T, uint32_t Q, uint32_t tid) {
    const uint32_t twoQ = 2 * Q, sixQ = 6 * Q;
    const uint32_t lane = tid & 31, lo2 = tid & 3;
    const ExIdx ex(tid);
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
            const uint32_t u = x[p][i], v = x[p][i + 1];
            x[p][i] = u + v;
            x[p][i + 1] = u + sixQ - v;
        }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (i & 2) continue;
#pragma unroll
        for (int p = 0; p < NP; ++p) MKFHE_BFLY(x[p][i], x[p][i + 2], T.itwA[2 + (i & 1)]);
    }
    __syncwarp();
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int c = 0; c < 2; ++c)
            *reinterpret_cast<uint4*>(E + p * FNP + ex.D(c)) = make_uint4(x[p][4 * c], x[p][4 * c + 1], x[p][4 * c + 2], x[p][4 * c + 3]);
    __syncwarp();
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < 8; ++r) x[p][r] = E[p * FNP + ex.C(r)];
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const int h = 1 << s;
        const uint32_t twi = 4 * h + lo2;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r & h) continue;
            const uint2 w = T.itwB[twi + 4 * (r & (h - 1))];
#pragma unroll
            for (int p = 0; p < NP; ++p) MKFHE_BFLY(x[p][r], x[p][r + h], w);
        }
    }
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < 8; ++r) E[p * FNP + ex.C(r)] = x[p][r];
    __syncwarp();
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < 8; ++r) x[p][r] = E[p * FNP + ex.B(r)];
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const int h = 1 << s;
        const uint2* tws = T.itw2 + 32 * h + lane;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r & h) continue;
            const uint2 w = __ldg(tws + 32 * (r & (h - 1)));
#pragma unroll
            for (int p = 0; p < NP; ++p) MKFHE_BFLY(x[p][r], x[p][r + h], w);
        }
    }
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < 8; ++r) E[p * FNP + ex.B(r)] = x[p][r];
    __syncthreads();
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < 8; ++r) x[p][r] = E[p * FNP + ex.A(r)];
#pragma unroll
    for (int s = 0; s < (HALF ? 2 : 3); ++s) {
        const int h = 1 << s;
        const uint2* tws = T.itw2 + 256 * h + tid;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r & h) continue;
            const uint2 w = __ldg(tws + 256 * (r & (h - 1)));
#pragma unroll
            for (int p = 0; p < NP; ++p) MKFHE_BFLY(x[p][r], x[p][r + h], w);
        }
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const uint2 sc = __ldg(T.scale2 + tid + 256 * r);
#pragma unroll
        for (int p = 0; p < NP; ++p) x[p][r] = shoup_lazy(x[p][r], sc, Q);
    }
}

// ---- helpers of the specialised step ---------------------------------------------------------------------------------

// SignedDigitDecompose of one canonical coefficient t (mk-acc.cpp:54-80), first digit dropped, two digits kept, each
// returned as r + Q in (0, 2Q) (same residue as the reference's "r < 0 ? r + Q : r"; the lazy transform takes < 2Q).
// sext_g(d) = ((d + B/2) mod B) - B/2 and (d - sext_g(d)) >> g = (d + B/2) >> g, so three rounding shifts suffice.
__device__ __forceinline__ void decompose2(uint32_t t, uint32_t Q, uint32_t g, uint32_t& o0, uint32_t& o1) {
    const int32_t half = 1 << (g - 1);
    const int32_t d = static_cast<int32_t>(t < (Q >> 1) ? t : t - Q);
    const int32_t d1 = (d + half) >> g;   // first digit ignored (approximate gadget decomposition)
    const int32_t d2 = (d1 + half) >> g;
    const int32_t d3 = (d2 + half) >> g;
    o0 = static_cast<uint32_t>(d1 - (d2 << g) + static_cast<int32_t>(Q));
    o1 = static_cast<uint32_t>(d2 - (d3 << g) + static_cast<int32_t>(Q));
}

// sum_j a_j * key_j * 2^-32 mod Q, in [0, 3Q): a_j < 24Q (lazy NTT output), key_j < Q in Montgomery form
__device__ __forceinline__ uint32_t mac2(uint32_t a0, uint32_t k0, uint32_t a1, uint32_t k1, const Mod m) {
    const uint64_t t = static_cast<uint64_t>(a0) * k0 + static_cast<uint64_t>(a1) * k1;
    const uint32_t q = static_cast<uint32_t>(t) * m.qinv_neg;
    return static_cast<uint32_t>((t + static_cast<uint64_t>(q) * m.Q) >> 32);
}

__device__ __forceinline__ uint4 ldg4(const uint32_t* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
// 128-bit key load that does not allocate in L1: a key tile is used once per gate-step (it is served by L2), and letting it
// through L1 evicts the 48 KiB of twiddle tables every transform reads (ncu: 33% L1 hit rate before, profiles/r1_s14_*)
__device__ __forceinline__ uint4 ldg4_stream(const uint32_t* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// L2 prefetch of the evaluation key(s) of step s (32 KiB each: 256 threads x one 128-byte line); s past the end: nothing
template <bool XZW>
__device__ __forceinline__ void prefetch_step_keys(const DevCtx& P, uint32_t s, uint32_t tid) {
    if (s >= P.k * P.n) return;
    const uint32_t u = s / P.n, i = s - u * P.n;
    const size_t per = size_t(2) * 2 * FN;
    const uint32_t* ev1 = P.bk + (size_t(u * P.S) * P.n + i) * per + 32 * tid;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(ev1));
    if (XZW) asm volatile("prefetch.global.L2 [%0];" ::"l"(ev1 + size_t(P.n) * per));
}

// 4 consecutive slots of one key polynomial as HbProd uses it in a step s > 0 (Montgomery form, canonical):
//   XZW_B: ev1                       (mk-acc-xzw_B.cpp:306-309)
//   XZW  : ev1 - ev2 * X^-c          (mk-acc-xzw.cpp:322-325: ev1 - ev2*(X^-c - 1) - ev2)
template <bool XZW>
__device__ __forceinline__ uint4 key4(const uint32_t* ev1, const uint32_t* ev2, const uint32_t (&xneg)[4], const Mod m) {
    uint4 a = ldg4_stream(ev1);
    if (XZW) {
        const uint4 b = ldg4_stream(ev2);
        a.x = submod(a.x, montmul(b.x, xneg[0], m), m.Q);
        a.y = submod(a.y, montmul(b.y, xneg[1], m), m.Q);
        a.z = submod(a.z, montmul(b.z, xneg[2], m), m.Q);
        a.w = submod(a.w, montmul(b.w, xneg[3], m), m.Q);
    }
    return a;
}

// Shared memory of one gate: a[k][N] | E1[2][FNP] | E2[FNP] (absent when ALIAS) | Ms[N] (XZW: Q - X^-c per slot) | crot[k*n] (u16).
// ALIAS: the inverse transform exchanges through E1[0] instead of its own buffer (one more block barrier per iteration);
// chosen by the host only when the separate buffer does not fit (see fast_use_alias).
__host__ __device__ inline size_t fast_smem_bytes(uint32_t k, uint32_t n, bool alias, bool xzw) {
    return (size_t(k) * FN + (alias ? 2 : 3) * FNP + (xzw ? FN : 0)) * 4 + ((size_t(k) * n * 2 + 15) & ~size_t(15));
}

// psi^e for the 4 slots 8 tid + 4 ch + q of X^x (x = c or -c mod 2N): slot j evaluates X at psi^(2 bitrev(j) + 1), and
// bitrev(8 tid + i) = bitrev3(i) << 8 | bitrev8(tid), so the exponent is e0 + 512 * (x * bitrev3(i) mod 8) with
// e0 = x * (2 bitrev8(tid) + 1).  Montgomery form (DevCtx::psipow).
__device__ __forceinline__ void monomial4(const uint32_t* psipow, uint32_t e0, uint32_t x, int ch, uint32_t (&m)[4]) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint32_t b3 = __brev(uint32_t(4 * ch + q)) >> 29;
        m[q] = __ldg(psipow + ((e0 + 512u * ((x * b3) & 7u)) & (2 * FN - 1)));
    }
}

// <D, key> at the 4 slots 8 tid + 4 ch + q of one key column (k1p / k2p / evsp point at digit 0 of that column of ev1 /
// ev2 / evk*, digit j at + 2 j N), result < 3Q.  D_j < 24Q (lazy transform output); keys canonical, Montgomery form.
template <bool XZW, bool FIRST>
__device__ __forceinline__ void key_mac4(const DevCtx& P, const uint32_t* k1p, const uint32_t* k2p, const uint32_t* evsp, uint32_t c,
                                         uint32_t cneg, uint32_t e0p, uint32_t e0n, int ch, const uint32_t (&D)[2][8],
                                         uint32_t (&x4)[4], const uint32_t* Ms = nullptr) {
    constexpr uint32_t N = FN, KCH = FN / 2;  // key polynomials are stored in register order: chunk ch of thread t at
                                              // words KCH ch + 4 t .. + 3 (fast_permute_keys), so a warp load is 512 contiguous bytes
    const uint32_t Q = P.mod.Q;
    if (FIRST) {
        uint32_t mp[4], mn[4];
        monomial4(P.psipow, e0p, c, ch, mp);
        if (XZW) monomial4(P.psipow, e0n, cneg, ch, mn);
        uint32_t kk[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const uint4 es = ldg4_stream(evsp + 2 * j * N + KCH * ch), e1 = ldg4_stream(k1p + 2 * j * N + KCH * ch);
            const uint32_t esv[4] = {es.x, es.y, es.z, es.w}, e1v[4] = {e1.x, e1.y, e1.z, e1.w};
            uint32_t e2v[4] = {0, 0, 0, 0};
            if (XZW) {
                const uint4 e2 = ldg4_stream(k2p + 2 * j * N + KCH * ch);
                e2v[0] = e2.x; e2v[1] = e2.y; e2v[2] = e2.z; e2v[3] = e2.w;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uint32_t r = addmod(esv[q], montmul(e1v[q], submod(mp[q], P.one_mont, Q), P.mod), Q);
                if (XZW) r = addmod(r, montmul(e2v[q], submod(mn[q], P.one_mont, Q), P.mod), Q);
                kk[j][q] = r;
            }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) x4[q] = mac2(D[0][4 * ch + q], kk[0][q], D[1][4 * ch + q], kk[1][q], P.mod);
    } else if (XZW) {
        // sum_j D_j (ev1_j - ev2_j X^-c) = sum_j D_j ev1_j + (Q - X^-c) * redc(sum_j D_j ev2_j): two 64-bit sums,
        // two reductions, five wide products per slot (mk-acc-xzw.cpp:322-325 assembles the key first: 2 more
        // reductions and the conditional subtractions)
        // Q - X^-c at the 4 slots: from the per-step staging in shared memory (Ms: this thread's own 8 values, written once
        // per step by fast_step), else gathered from the power table
        uint32_t mq[4];
        if (Ms != nullptr) {
            const uint4 v = *reinterpret_cast<const uint4*>(Ms + KCH * ch);
            mq[0] = v.x; mq[1] = v.y; mq[2] = v.z; mq[3] = v.w;
        } else {
            monomial4(P.psipow, e0n, cneg, ch, mq);
#pragma unroll
            for (int q = 0; q < 4; ++q) mq[q] = Q - mq[q];
        }
        const uint4 a0 = ldg4_stream(k1p + KCH * ch), a1 = ldg4_stream(k1p + 2 * N + KCH * ch);
        const uint4 b0 = ldg4_stream(k2p + KCH * ch), b1 = ldg4_stream(k2p + 2 * N + KCH * ch);
        const uint32_t a0v[4] = {a0.x, a0.y, a0.z, a0.w}, a1v[4] = {a1.x, a1.y, a1.z, a1.w};
        const uint32_t b0v[4] = {b0.x, b0.y, b0.z, b0.w}, b1v[4] = {b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t d0 = D[0][4 * ch + q], d1 = D[1][4 * ch + q];  // < 24Q
            const uint64_t s2 = static_cast<uint64_t>(d0) * b0v[q] + static_cast<uint64_t>(d1) * b1v[q];  // < 48 Q^2
            const uint32_t q2 = static_cast<uint32_t>(s2) * P.mod.qinv_neg;
            const uint32_t r2 = static_cast<uint32_t>((s2 + static_cast<uint64_t>(q2) * Q) >> 32);  // < 3Q
            const uint64_t s1 = static_cast<uint64_t>(d0) * a0v[q] + static_cast<uint64_t>(d1) * a1v[q] +
                                static_cast<uint64_t>(r2) * mq[q];  // < 51 Q^2 < 2^60
            const uint32_t q1 = static_cast<uint32_t>(s1) * P.mod.qinv_neg;
            x4[q] = static_cast<uint32_t>((s1 + static_cast<uint64_t>(q1) * Q) >> 32);  // < 3Q
        }
    } else {
        const uint4 a0 = ldg4_stream(k1p + KCH * ch), a1 = ldg4_stream(k1p + 2 * N + KCH * ch);
        const uint32_t a0v[4] = {a0.x, a0.y, a0.z, a0.w}, a1v[4] = {a1.x, a1.y, a1.z, a1.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) x4[q] = mac2(D[0][4 * ch + q], a0v[q], D[1][4 * ch + q], a1v[q], P.mod);
    }
}

// One step s of EvalAcc (mk-acc-xzw.cpp:106-129, mk-acc-xzw_B.cpp:118-132) with rotation c, on the register-resident
// transforms.  FIRST: the step (u, i) = (0, 0) -- AddToAccXZW0 (mk-acc-xzw.cpp:347-381, mk-acc-xzw_B.cpp:333-379): the KDM
// key evs + ev1 (X^c - 1) [+ ev2 (X^-c - 1)] is assembled per slot, the accumulator itself is decomposed (no monomial) and
// the result REPLACES it.  Runs once per gate, so it is a separate instantiation instead of selects in the hot loop.
// Entry: a[] visible to all threads, nobody still reading E1 / E2.  Exit: as entry (ends with a block barrier).
// PAIR (the builds with 80 or 128 registers): inverse transforms are taken two at a time.  A non-owner party's product X_p is
// parked in registers until the next product is ready -- the next non-owner's, or the owner's V -- and both go back to
// coefficients in ONE two-polynomial inverse (inv_ntt_regs_np<2>): the pair shares every twiddle load (22 of the inverse's 66
// LSU instructions) and doubles the independent butterflies per stage, which is what the low-occupancy builds are short of
// (stall `wait`, profiles/r2_s1_k8_296.txt).  Same exact sums, in a different order of evaluation.  The pair exchanges through
// E1 (both halves): no barrier is needed against the forward transforms around it, because a transform's last exchange and the
// next transform's first exchange use the same per-thread-private positions (forward ends on pattern D and the inverse begins
// on D; the inverse ends on A and the next forward begins on A).
template <bool XZW, bool FIRST, bool ALIAS, bool PREFETCH = false, bool PAIR = false>  // PREFETCH: the builds that serve small batches
__device__ __forceinline__ void fast_step(const DevCtx& P, const FastTables& T, uint32_t* a, uint32_t* E1, uint32_t* E2, uint32_t* Ms,
                                          uint32_t s, uint32_t c, uint32_t tid) {
    constexpr uint32_t N = FN, M = 2 * FN;
    const uint32_t Q = P.mod.Q, k = P.k, n = P.n, g = P.gbits;
    const uint32_t u = FIRST ? 0u : s / n, i = FIRST ? 0u : s - u * n;
    const uint32_t cneg = (M - c) & (M - 1);
    const size_t per = size_t(2) * 2 * N;  // words of one UniEncEvalKey: [j][d|f][N]
    const uint32_t* ev1 = P.bk + (size_t(u * P.S) * n + i) * per + 4 * tid;
    const uint32_t* ev2 = XZW ? ev1 + size_t(n) * per : ev1;
    const uint32_t* evs = P.bk + size_t(k) * P.S * n * per + 4 * tid;  // evk* = ek[0][0][n] (mk-acc-xzw.cpp:51,72-74)
    const uint32_t brv8 = __brev(tid) >> 24;
    const uint32_t e0n = cneg * (2 * brv8 + 1), e0p = c * (2 * brv8 + 1);
    if (PREFETCH && T.prefetch) prefetch_step_keys<XZW>(P, s + 1, tid);
    uint32_t keep[8];  // V while parties are processed, then X of the key owner
    uint32_t y[8];
    uint32_t hold[8];  // PAIR: a non-owner's product waiting for a partner
    uint32_t hp = 0;
    bool have = false;
    // XZW: Q - X^-c at this thread's 8 slots is the same in all k + 1 iterations of the step.  Gathering it from the power
    // table costs 8 scattered 4-byte loads per iteration (up to 32 L1 wavefronts each: as much LSU traffic as everything
    // else in the iteration); it is gathered once per step instead and parked in shared memory at this thread's own words
    // (register order, so it comes back as two 128-bit loads; no barrier: nobody else reads them).
    uint32_t* Mt = Ms + 4 * tid;
    if (XZW && !FIRST) {
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            uint32_t mq[4];
            monomial4(P.psipow, e0n, cneg, ch, mq);
            *reinterpret_cast<uint4*>(Mt + (N / 2) * ch) = make_uint4(Q - mq[0], Q - mq[1], Q - mq[2], Q - mq[3]);
        }
    }
#pragma unroll 1
    for (uint32_t it = 0; it <= k; ++it) {
        uint32_t D[2][8];
        uint32_t p = u;
        if (it < k) {
            p = (it + 1 < k) ? (it < u ? it : it + 1) : u;  // the key owner last: its f-term needs the complete V
            const uint32_t* ap = a + p * N;
#pragma unroll
            for (int r = 0; r < 8; ++r)
                decompose2(FIRST ? ap[tid + 256 * r] : rot_sub(ap, c, tid + 256 * r, N, Q), Q, g, D[0][r], D[1][r]);
        } else {
#pragma unroll
            for (int r = 0; r < 8; ++r) decompose2(fold1(y[r], Q), Q, g, D[0][r], D[1][r]);
        }
        fwd_ntt_regs<2>(D, E1, T, Q, tid);
        const int col = it < k ? 0 : 1;  // "d" column for the parties, "f" column for the V term
        const uint32_t* pk = P.pkey + size_t(p) * 2 * N + 4 * tid;
        const uint32_t* k1p = ev1 + col * N;  // digit j at + 2 j N
        const uint32_t* k2p = ev2 + col * N;
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            uint32_t x4[4];  // <D, key> at slots 8 tid + 4 ch + q, < 3Q
            key_mac4<XZW, FIRST>(P, k1p, k2p, evs + col * N, c, cneg, e0p, e0n, ch, D, x4, (XZW && !FIRST) ? Mt : nullptr);
            if (it < k) {
                const uint4 q0 = ldg4(pk + (N / 2) * ch), q1 = ldg4(pk + N + (N / 2) * ch);
                const uint32_t pp0[4] = {q0.x, q0.y, q0.z, q0.w}, pp1[4] = {q1.x, q1.y, q1.z, q1.w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int sl = 4 * ch + q;
                    const uint32_t v = mac2(D[0][sl], pp0[q], D[1][sl], pp1[q], P.mod);
                    uint32_t vs = (it == 0 ? 0u : keep[sl]) + v;  // < 7Q
                    vs = min(vs, vs - 4 * Q);                     // < 4Q
                    if (it + 1 < k) {
                        keep[sl] = vs;
                        y[sl] = x4[q];
                    } else {  // owner: transform V next, hold X
                        y[sl] = vs;
                        keep[sl] = x4[q];
                    }
                }
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q) y[4 * ch + q] = keep[4 * ch + q] + x4[q];  // < 6Q
            }
        }
        if (PAIR && !FIRST && it < k) {
            if (it + 1 < k && !have) {  // a non-owner with nothing parked: park it, no inverse yet
#pragma unroll
                for (int r = 0; r < 8; ++r) hold[r] = y[r];
                hp = p;
                have = true;
                continue;
            }
            if (have) {  // parked product + this one (a non-owner's X_p, or the owner's V) in one two-polynomial inverse
                uint32_t z[2][8];
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    z[0][r] = hold[r];
                    z[1][r] = y[r];
                }
                inv_ntt_regs_np<2>(z, E1, T, Q, tid);
                uint32_t* ah = a + hp * N;
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    uint32_t v = ah[tid + 256 * r] + z[0][r];  // < 3Q
                    v = min(v, v - 2 * Q);
                    ah[tid + 256 * r] = min(v, v - Q);
                    y[r] = z[1][r];
                }
                have = false;
                if (it + 1 < k) {
                    uint32_t* ap = a + p * N;
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        uint32_t v = ap[tid + 256 * r] + y[r];
                        v = min(v, v - 2 * Q);
                        ap[tid + 256 * r] = min(v, v - Q);
                    }
                }
                continue;  // (the owner's V stays in y for the next iteration)
            }
        }
        inv_ntt_regs(y, ALIAS ? E1 : E2, T, Q, tid);
        if (it + 1 != k) {  // a_p (+)= result (the owner's V goes on to the next iteration in registers instead)
            uint32_t* ap = a + p * N;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                if (FIRST) {
                    ap[tid + 256 * r] = fold1(y[r], Q);
                } else {
                    uint32_t v = ap[tid + 256 * r] + y[r];  // < 3Q
                    v = min(v, v - 2 * Q);
                    ap[tid + 256 * r] = min(v, v - Q);
                }
            }
        }
        // (a_p itself needs no barrier here: it was read -- at shifted positions, by other threads -- before the two
        // barriers inside this iteration's transforms, and is read again only after the barrier that ends the step.  Nor does
        // ALIAS: the inverse's last exchange reads, and the next forward's first exchange writes, pattern A -- positions
        // private to each thread; round 1 had a block barrier here that bought nothing.)
    }
    __syncthreads();  // a[] complete before the next step's shifted reads
}

// MINB: gates the launch bounds promise per SM = the register budget (4: 64 registers, 3: 80, 2: 128).  The host picks the
// largest MINB shared memory allows anyway, so party counts that fit fewer gates per SM get the registers the absent gates
// would have used (no spills, longer load-to-use distances).
// PAIR: inverse transforms two at a time (fast_step), a build of its own: as a run-time switch inside one kernel the paired step's
// 240 B of spills in the 80-register XZW build slowed the UNPAIRED path of the same kernel too (boolean-mkntru, 4096 gates:
// 242.4 -> 249.8 ms with either setting, profiles/r2_s14_pair_alias.txt, r2_s27_pair_noprefetch.txt).
template <bool GATE, bool XZW, bool ALIAS, int MINB, bool PAIR = false>
__global__ void __launch_bounds__(FT, MINB) blind_rotate_fast(const DevCtx P, const FastTables T, const GateIO io) {
    extern __shared__ __align__(16) uint32_t sm[];
    const uint32_t k = P.k, n = P.n, tid = threadIdx.x;
    constexpr uint32_t N = FN;
    uint32_t* a = sm;
    uint32_t* E1 = a + k * N;
    uint32_t* E2 = E1 + 2 * FNP;
    uint32_t* Ms = ALIAS ? E2 : E2 + FNP;
    uint16_t* crot = reinterpret_cast<uint16_t*>(XZW ? Ms + N : Ms);

    if (T.stagger_ns) __nanosleep((blockIdx.x & 3u) * T.stagger_ns);
    rotate_prologue<GATE>(P, io, blockIdx.x, a, crot);
    // (no key prefetch in this kernel: batches of at most one gate per SM, the only ones it pays for, run on the cluster and
    // party-parallel kernels; carrying the code here cost the 80-register XZW build 240 B of spills and 3% at 4096 gates)
    // (inlined on purpose: as a __noinline__ function the first step takes the kernel parameters by address, which moves them to
    // local memory for the whole kernel: 178.6 -> 189.6 ms at 4096 gates, profiles/r2_s29_first_noinline.txt)
    // (nor does its presence cost the step loop anything: a build without the first step runs the loop at the same speed,
    // 178.80 vs 178.75 ms, profiles/r2_s34_nofirst.txt)
    fast_step<XZW, true, ALIAS, false, false>(P, T, a, E1, E2, Ms, 0, crot[0], tid);
    for (uint32_t s = 1; s < k * n; ++s) {
        const uint32_t c = crot[s];
        if (c == 0) continue;  // X^0 - 1 = 0: the step leaves acc unchanged
        fast_step<XZW, false, ALIAS, false, PAIR>(P, T, a, E1, E2, Ms, s, c, tid);
    }
    rotate_epilogue<GATE>(P, io, blockIdx.x, a, E1);
}

// ---- accumulators in global memory (L2) for large party counts ---------------------------------------------------------------
// From k = 8 on the accumulator a[k][N] (8 KiB per party) decides how many gates share an SM: 2 at k = 8, 1 at k = 16, and an SM
// with 8-16 warps is latency-bound (stall `wait`, fma-heavy pipe 57-61% busy against 71% at k = 2 with 32 warps;
// profiles/r2_s1_k8_296.txt, r2_s11_k16_148.txt).  Here the accumulator of a gate lives in a per-CTA slice of a global workspace
// instead: only the resident CTAs' slices are live (148 x MINB x k x 8 KiB <= 76 MiB at k = 16, inside the 126 MB L2), shared memory
// holds the exchange buffers and rotation amounts alone, and MINB = 4 gates (32 warps) share an SM at every party count.  The step
// function is the same code: `a` is a generic pointer, the accesses (coalesced 4-byte rows, one wavefront each as in shared memory)
// become LDG / STG served by L2, and the block barriers that ordered them in shared memory order them in global memory as well
// (all accesses to a slice come from the one CTA that owns it).  The CTA is persistent: gate g = blockIdx.x, + gridDim.x, ...
// Shared memory: E1[2][FNP] | E2[FNP] | Ms[N] (XZW) | crot[k*n] (u16).
// MEASURED (see fast_ag_wanted): bit-exact, and not faster except at k = 16 boolean-mklwe (+2%) -- kept as a selectable experiment.
__host__ __device__ inline size_t fast_ag_smem_bytes(uint32_t k, uint32_t n, bool xzw) {
    return (size_t(3) * FNP + (xzw ? FN : 0)) * 4 + ((size_t(k) * n * 2 + 15) & ~size_t(15));
}
template <bool GATE, bool XZW, int MINB>
__global__ void __launch_bounds__(FT, MINB) blind_rotate_fast_ag(const DevCtx P, const FastTables T, const GateIO io, uint32_t* accws,
                                                                 uint32_t batch) {
    extern __shared__ __align__(16) uint32_t sm[];
    const uint32_t k = P.k, n = P.n, tid = threadIdx.x;
    constexpr uint32_t N = FN;
    uint32_t* E1 = sm;
    uint32_t* E2 = E1 + 2 * FNP;
    uint32_t* Ms = E2 + FNP;
    uint16_t* crot = reinterpret_cast<uint16_t*>(XZW ? Ms + N : Ms);
    uint32_t* a = accws + size_t(blockIdx.x) * k * N;
    for (uint32_t g = blockIdx.x; g < batch; g += gridDim.x) {
        rotate_prologue<GATE>(P, io, g, a, crot);
        fast_step<XZW, true, false, false, false>(P, T, a, E1, E2, Ms, 0, crot[0], tid);
        for (uint32_t s = 1; s < k * n; ++s) {
            const uint32_t c = crot[s];
            if (c == 0) continue;
            fast_step<XZW, false, false, false, false>(P, T, a, E1, E2, Ms, s, c, tid);
        }
        rotate_epilogue<GATE>(P, io, g, a, E1);
        __syncthreads();  // the next gate's prologue rewrites crot and a
    }
}

// ---- party-parallel variant for large party counts ----------------------------------------------------------------------
// At k >= 8 one gate's accumulator takes 64-128 KiB of shared memory, so only 2 (k = 8) or 1 (k = 16) gates fit an SM and a
// 256-thread CTA per gate leaves the SM with 16 or 8 warps.  Here one CTA of G x 256 threads owns one gate and the G groups
// process different parties of the same step concurrently: the parties' terms of HbProd (mk-acc-xzw.cpp:240-271) are
// independent except for sumV, which the groups accumulate in shared memory with atomic adds (sums of canonical residues:
// k Q < 2^32).  Then the owner's group runs the sumV term (:273-289) while the others finish their last party: a step takes
// ceil(k / G) + 1 transform rounds instead of k + 1, and the same 3k + 3 transforms as the sequential schedule.  Group barriers are named barriers (1 + group); block barriers separate the phases.
// Shared memory: a[k][N] | per group E1[2][FNP] | E2[FNP] (absent when ALIAS) | Vs[N] | Ms[N] (XZW: Q - X^-c per slot, shared by
// the groups).  The rotation amounts are not kept on chip (k = 16 needs the space): rotation_amount() re-derives the step's
// amount from the gate's inputs.
__host__ __device__ inline size_t grouped_smem_bytes(uint32_t k, uint32_t n, uint32_t G, bool alias) {
    (void)n;
    return (size_t(k) * FN + size_t(G) * (alias ? 2 : 3) * FNP + 2 * FN) * 4;
}

template <bool XZW, bool FIRST, bool ALIAS>
__device__ __forceinline__ void grouped_step(const DevCtx& P, const FastTables& T, uint32_t* a, uint32_t* E1, uint32_t* E2, uint32_t* Vs,
                                             uint32_t* Ms, uint32_t s, uint32_t c, uint32_t tid, uint32_t grp, uint32_t G) {
    constexpr uint32_t N = FN, M = 2 * FN;
    const uint32_t Q = P.mod.Q, k = P.k, n = P.n, g = P.gbits, bar = 1 + grp;
    const uint32_t u = FIRST ? 0u : s / n, i = FIRST ? 0u : s - u * n;
    const uint32_t cneg = (M - c) & (M - 1);
    const size_t per = size_t(2) * 2 * N;
    const uint32_t* ev1 = P.bk + (size_t(u * P.S) * n + i) * per + 4 * tid;
    const uint32_t* ev2 = XZW ? ev1 + size_t(n) * per : ev1;
    const uint32_t* evs = P.bk + size_t(k) * P.S * n * per + 4 * tid;
    const uint32_t brv8 = __brev(tid) >> 24;
    const uint32_t e0n = cneg * (2 * brv8 + 1), e0p = c * (2 * brv8 + 1);
    uint32_t y[8];
    if (T.prefetch && grp == 0) prefetch_step_keys<XZW>(P, s + 1, tid);
    // XZW: Q - X^-c per slot, once per step (see fast_step).  Every group writes the same values to the same words (slot =
    // f(tid) only), each thread reads back the words it wrote itself; steps are separated by block barriers.
    uint32_t* Mt = Ms + 4 * tid;
    if (XZW && !FIRST) {
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            uint32_t mq[4];
            monomial4(P.psipow, e0n, cneg, ch, mq);
            *reinterpret_cast<uint4*>(Mt + (N / 2) * ch) = make_uint4(Q - mq[0], Q - mq[1], Q - mq[2], Q - mq[3]);
        }
    }
    // ---- phase 1: the parties, round-robin over the groups (order index k - 1 = the key owner, always the LAST party of its
    // group).  Every party's <D_p, d> is transformed back and added to a_p right away -- except a group's last party: its
    // product stays in registers (y) across the barrier that completes sumV, so that the critical path of a step (the owner
    // group: party terms, then the sumV term) loses one inverse transform: the owner's <D_u, d> is added to <g^-1(sumV), f> in
    // the evaluation domain and transformed back once (as the sequential kernel does), and the other groups transform their
    // held product back WHILE the owner group works on the sumV term, instead of idling through it.
#pragma unroll 1
    for (uint32_t it = grp; it < k; it += G) {
        uint32_t D[2][8];
        const uint32_t p = (it + 1 < k) ? (it < u ? it : it + 1) : u;
        uint32_t* ap = a + p * N;
#pragma unroll
        for (int r = 0; r < 8; ++r) decompose2(FIRST ? ap[tid + 256 * r] : rot_sub(ap, c, tid + 256 * r, N, Q), Q, g, D[0][r], D[1][r]);
        fwd_ntt_regs<2, true>(D, E1, T, Q, tid, bar);
        const uint32_t* pk = P.pkey + size_t(p) * 2 * N + 4 * tid;
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            uint32_t x4[4];
            key_mac4<XZW, FIRST>(P, ev1, ev2, evs, c, cneg, e0p, e0n, ch, D, x4, (XZW && !FIRST) ? Mt : nullptr);
            const uint4 q0 = ldg4(pk + (N / 2) * ch), q1 = ldg4(pk + N + (N / 2) * ch);
            const uint32_t pp0[4] = {q0.x, q0.y, q0.z, q0.w}, pp1[4] = {q1.x, q1.y, q1.z, q1.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int sl = 4 * ch + q;
                uint32_t v = mac2(D[0][sl], pp0[q], D[1][sl], pp1[q], P.mod);  // < 3Q
                v = min(v, v - 2 * Q);
                atomicAdd(Vs + 256 * sl + tid, min(v, v - Q));  // canonical: the sum over k <= 16 parties stays below 2^31
                y[sl] = x4[q];
            }
        }
        if (it + G < k) {  // not this group's last party: back to coefficients now
            inv_ntt_regs<true>(y, ALIAS ? E1 : E2, T, Q, tid, bar);
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                if (FIRST) {
                    ap[tid + 256 * r] = fold1(y[r], Q);
                } else {
                    uint32_t v = ap[tid + 256 * r] + y[r];  // < 3Q
                    v = min(v, v - 2 * Q);
                    ap[tid + 256 * r] = min(v, v - Q);
                }
            }
            if (ALIAS) gate_sync<true>(bar);
        }
    }
    // the party whose product this group still holds in y (evaluation domain, < 3Q): order index of its last iteration
    const uint32_t itl = grp + ((k - 1 - grp) / G) * G;
    const uint32_t pl = (itl + 1 < k) ? (itl < u ? itl : itl + 1) : u;
    __syncthreads();  // sumV complete in Vs
    if (grp == (k - 1) % G) {
        // ---- phase 2 (owner group): the sumV term <g^-1(sumV), f>, added to the held <D_u, d>, one inverse, a_u ----
        uint32_t D[2][8], v8[8];
#pragma unroll
        for (int sl = 0; sl < 8; ++sl) {  // Vs[256 sl + tid] = slot 8 tid + sl: conflict-free for the atomics and here
            uint32_t v = Vs[256 * sl + tid];  // < 16Q
            Vs[256 * sl + tid] = 0;           // ready for the next step
            v = min(v, v - 8 * Q);
            v8[sl] = min(v, v - 4 * Q);       // < 4Q
        }
        inv_ntt_regs<true>(v8, ALIAS ? E1 : E2, T, Q, tid, bar);
        if (ALIAS) gate_sync<true>(bar);
#pragma unroll
        for (int r = 0; r < 8; ++r) decompose2(fold1(v8[r], Q), Q, g, D[0][r], D[1][r]);
        fwd_ntt_regs<2, true>(D, E1, T, Q, tid, bar);
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            uint32_t x4[4];
            key_mac4<XZW, FIRST>(P, ev1 + N, ev2 + N, evs + N, c, cneg, e0p, e0n, ch, D, x4, (XZW && !FIRST) ? Mt : nullptr);  // "f" column
#pragma unroll
            for (int q = 0; q < 4; ++q) y[4 * ch + q] += x4[q];  // < 6Q
        }
    }
    // every group: its held product (the owner group's now includes the sumV term) back to coefficients, into a_pl
    {
        inv_ntt_regs<true>(y, ALIAS ? E1 : E2, T, Q, tid, bar);
        uint32_t* ap = a + pl * N;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (FIRST) {
                ap[tid + 256 * r] = fold1(y[r], Q);
            } else {
                uint32_t v = ap[tid + 256 * r] + y[r];
                v = min(v, v - 2 * Q);
                ap[tid + 256 * r] = min(v, v - Q);
            }
        }
    }
    __syncthreads();  // a[] complete (and every group past its last exchange) before the next step
}

template <bool GATE, bool XZW, bool ALIAS, int MAXT>  // MAXT = 256 G: 2 groups get 128 registers per thread, 3 groups 80, 4 groups 64
__global__ void __launch_bounds__(MAXT, 1) blind_rotate_grouped(const DevCtx P, const FastTables T, const GateIO io) {
    extern __shared__ __align__(16) uint32_t sm[];
    const uint32_t k = P.k, n = P.n, G = blockDim.x >> 8, grp = threadIdx.x >> 8, tid = threadIdx.x & 255;
    constexpr uint32_t N = FN;
    constexpr uint32_t EW = (ALIAS ? 2 : 3) * FNP;  // exchange words per group
    uint32_t* a = sm;
    uint32_t* E1 = a + k * N + grp * EW;
    uint32_t* E2 = E1 + 2 * FNP;
    uint32_t* Vs = a + k * N + G * EW;
    uint32_t* Ms = Vs + N;

    for (uint32_t e = threadIdx.x; e < N; e += blockDim.x) Vs[e] = 0;
    rotate_prologue<GATE>(P, io, blockIdx.x, a, nullptr);  // ends with a block barrier
    grouped_step<XZW, true, ALIAS>(P, T, a, E1, E2, Vs, Ms, 0, rotation_amount<GATE>(P, io, blockIdx.x, 0), tid, grp, G);
    for (uint32_t s = 1; s < k * n; ++s) {
        const uint32_t c = rotation_amount<GATE>(P, io, blockIdx.x, s);
        if (c == 0) continue;
        grouped_step<XZW, false, ALIAS>(P, T, a, E1, E2, Vs, Ms, s, c, tid, grp, G);
    }
    rotate_epilogue<GATE>(P, io, blockIdx.x, a, a + k * N);  // scratch: group 0's exchange buffer
}

// ---- standalone batched transform (config 5, mkfhe_ntt): one polynomial per CTA-iteration, persistent grid -----------
// HBM-bound at large batch (16 KiB moved per 11 264 butterflies), so the next polynomial's coefficients are fetched into
// registers before the current one is transformed: loads of iteration i+1 overlap the arithmetic of iteration i.
template <bool HALF, bool INV>  // HALF: N = 1024, one CTA-iteration transforms the polynomials 2b and 2b + 1 (batch counts pairs);
                                // INV: one direction per kernel, as in ntt_batch_tma2 (the A/B twins are built the way the default is)
__global__ void __launch_bounds__(FT, 4) ntt_batch_fast(const FastTables T, const uint32_t Q, const uint2 one2, uint32_t* polys,
                                                        const uint64_t batch) {
    constexpr bool inverse = INV;
    __shared__ __align__(16) uint32_t E[FNP];
    const uint32_t tid = threadIdx.x;
    uint32_t nx[8];
    auto fetch = [&](uint64_t b) {
        const uint32_t* p = polys + b * FN;
        if (!inverse) {
#pragma unroll
            for (int r = 0; r < 8; ++r) nx[r] = __ldcs(p + tid + 256 * r);
        } else {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const uint4 v = __ldcs(reinterpret_cast<const uint4*>(p + 8 * tid + 4 * c));
                nx[4 * c] = v.x; nx[4 * c + 1] = v.y; nx[4 * c + 2] = v.z; nx[4 * c + 3] = v.w;
            }
        }
    };
    if (blockIdx.x < batch) fetch(blockIdx.x);
    for (uint64_t b = blockIdx.x; b < batch; b += gridDim.x) {
        uint32_t* p = polys + b * FN;
        if (!inverse) {
            uint32_t x[1][8];
#pragma unroll
            for (int r = 0; r < 8; ++r) x[0][r] = nx[r];
            if (b + gridDim.x < batch) fetch(b + gridDim.x);
            fwd_ntt_regs<1, false, HALF>(x, E, T, Q, tid);
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint4 v;
                v.x = fold1(shoup_lazy(x[0][4 * c + 0], one2, Q), Q);
                v.y = fold1(shoup_lazy(x[0][4 * c + 1], one2, Q), Q);
                v.z = fold1(shoup_lazy(x[0][4 * c + 2], one2, Q), Q);
                v.w = fold1(shoup_lazy(x[0][4 * c + 3], one2, Q), Q);
                __stcs(reinterpret_cast<uint4*>(p + 8 * tid + 4 * c), v);
            }
        } else {
            uint32_t x[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) x[r] = nx[r];
            if (b + gridDim.x < batch) fetch(b + gridDim.x);
            inv_ntt_regs<false, HALF>(x, E, T, Q, tid);
#pragma unroll
            for (int r = 0; r < 8; ++r) __stcs(p + tid + 256 * r, fold1(x[r], Q));
        }
        __syncthreads();  // E is reused by the next polynomial
    }
}

// ---- the same transform fed by the bulk-copy engine (TMA): cp.async.bulk + mbarrier ring ------------------------------------
// One elected thread streams whole polynomials (8 KiB, contiguous in HBM) global -> shared with `cp.async.bulk` (1-D bulk copy,
// SASS UBLKCP), completion signalled on an mbarrier per ring slot (`complete_tx::bytes`); the CTA's 256 threads pick their
// coefficients out of the slot, transform in registers exactly as ntt_batch_fast does, put the result back INTO THE SAME SLOT
// (every thread has passed a block barrier since it read the slot), and the elected thread sends the slot back with
// `cp.async.bulk.global.shared::cta` (bulk store).  NTT_STAGES slots per CTA: while polynomial i is transformed, the loads of
// i+1 .. i+S-1 and the stores of i-1, i-2 are in flight -- no thread holds a register for data in flight, no LDG/STG issue
// slots, no long-scoreboard stalls in the compute warps.  Slot reuse: the store of iteration i-1 must have finished READING
// its slot before the load of polynomial i+S-1 may overwrite it (`cp.async.bulk.wait_group.read 1`: all bulk groups but the
// newest).  Generic-proxy writes (STS of the results) are made visible to the async proxy by `fence.proxy.async` before the
// barrier that precedes the store.
// Resident CTAs per SM (CTAS) and ring depth: the transform needs 48 registers with one polynomial in flight per thread, so 5
// CTAs (40 warps) fit the register file with a 4-deep ring (42 KiB each) and 6 CTAs (40 registers) with a 3-deep ring.
__host__ __device__ constexpr int ntt_stages(int ctas) { return ctas >= 6 ? 3 : 4; }
__host__ __device__ inline size_t ntt_tma_smem_bytes(int ctas) { return (size_t(ntt_stages(ctas)) * FN + FNP) * 4 + ntt_stages(ctas) * 8; }

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    uint32_t done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(a), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gmem_dst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// x in [0, 2^32) -> [0, 2Q) without a multiplication by a reciprocal: with sh = bits of Q and delta = 2^sh - Q,
// x - (x >> sh) Q = (x mod 2^sh) + (x >> sh) delta < 2^sh + 32 delta <= 2Q whenever 32 delta <= Q - delta (true for the
// reference's Q = 2^27 - 40959 and for every "previous prime below 2^27"); one SHF + one IMAD instead of IMAD.HI + IMAD.
__device__ __forceinline__ uint32_t fold_shift27(uint32_t x, uint32_t Q) { return x - (x >> 27) * Q; }
__host__ __device__ inline bool fold_shift27_ok(uint32_t Q) { return Q < (1u << 27) && Q > (1u << 26) && 33ull * ((1u << 27) - Q) <= Q; }

template <bool HALF, bool SHIFTFOLD, int CTAS, bool INV>
__global__ void __launch_bounds__(FT, CTAS) ntt_batch_tma(const FastTables T, const uint32_t Q, const uint2 one2, uint32_t* polys,
                                                       const uint64_t batch) {
    constexpr bool inverse = INV;
    extern __shared__ __align__(128) uint32_t sm[];
    constexpr int NTT_STAGES = ntt_stages(CTAS);
    uint32_t* ring = sm;
    uint32_t* E = ring + NTT_STAGES * FN;
    uint64_t* full = reinterpret_cast<uint64_t*>(E + FNP);
    const uint32_t tid = threadIdx.x;
    constexpr uint32_t BYTES = FN * 4;
    const uint64_t mine = blockIdx.x < batch ? (batch - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;  // polynomials of this CTA
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < NTT_STAGES; ++s) mbar_init(full + s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (uint64_t i = 0; i < mine && i < NTT_STAGES - 1; ++i) {
            mbar_expect_tx(full + i, BYTES);
            bulk_load(ring + i * FN, polys + (blockIdx.x + i * gridDim.x) * FN, BYTES, full + i);
        }
    }
    __syncthreads();
    for (uint64_t i = 0; i < mine; ++i) {
        const uint32_t s = uint32_t(i % NTT_STAGES), parity = uint32_t(i / NTT_STAGES) & 1u;
        uint32_t* buf = ring + s * FN;
        mbar_wait(full + s, parity);
        if (!inverse) {
            uint32_t x[1][8];
#pragma unroll
            for (int r = 0; r < 8; ++r) x[0][r] = buf[tid + 256 * r];
            fwd_ntt_regs<1, false, HALF>(x, E, T, Q, tid);
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint4 v;
                if (SHIFTFOLD) {
                    v.x = fold1(fold_shift27(x[0][4 * c + 0], Q), Q); v.y = fold1(fold_shift27(x[0][4 * c + 1], Q), Q);
                    v.z = fold1(fold_shift27(x[0][4 * c + 2], Q), Q); v.w = fold1(fold_shift27(x[0][4 * c + 3], Q), Q);
                } else {
                    v.x = fold1(shoup_lazy(x[0][4 * c + 0], one2, Q), Q); v.y = fold1(shoup_lazy(x[0][4 * c + 1], one2, Q), Q);
                    v.z = fold1(shoup_lazy(x[0][4 * c + 2], one2, Q), Q); v.w = fold1(shoup_lazy(x[0][4 * c + 3], one2, Q), Q);
                }
                *reinterpret_cast<uint4*>(buf + 8 * tid + 4 * c) = v;
            }
        } else {
            uint32_t x[8];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const uint4 v = *reinterpret_cast<const uint4*>(buf + 8 * tid + 4 * c);
                x[4 * c] = v.x; x[4 * c + 1] = v.y; x[4 * c + 2] = v.z; x[4 * c + 3] = v.w;
            }
            inv_ntt_regs<false, HALF>(x, E, T, Q, tid);
#pragma unroll
            for (int r = 0; r < 8; ++r) buf[tid + 256 * r] = fold1(x[r], Q);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // this thread's STS -> visible to the bulk-copy engine
        __syncthreads();  // results complete in the slot; E free for the next polynomial
        if (tid == 0) {
            bulk_store(polys + (blockIdx.x + i * gridDim.x) * FN, buf, BYTES);
            const uint64_t nx = i + NTT_STAGES - 1;  // goes to slot (i - 1) % S, whose store was issued one iteration ago
            if (nx < mine) {
                if (i > 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                const uint32_t ns = uint32_t(nx % NTT_STAGES);
                mbar_expect_tx(full + ns, BYTES);
                bulk_load(ring + ns * FN, polys + (blockIdx.x + nx * gridDim.x) * FN, BYTES, full + ns);
            }
        }
    }
    if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // shared memory stays valid until the stores have read it
}

// Two polynomials per ring slot and CTA iteration (16 KiB bulk copies): the transforms of a pair share every twiddle load, which
// takes 13 (forward) / 22 (inverse) of the ~90 / ~108 LSU wavefronts per warp and polynomial off the kernel's busiest unit.
// 3 slots of 16 KiB + two exchange buffers = 66 KiB per CTA, 3 CTAs per SM at up to 80 registers.  `pairs` counts slots; an odd
// last polynomial is the caller's business (it goes through ntt_batch_tma).
constexpr int NTT2_STAGES = 3;
__host__ __device__ inline size_t ntt_tma2_smem_bytes() { return (size_t(NTT2_STAGES) * 2 * FN + 2 * FNP) * 4 + NTT2_STAGES * 8; }

// INV is a template parameter, not a kernel argument: one direction per kernel (a run-time switch between two inlined transform
// bodies is what slowed the gate kernel's paired step, see blind_rotate_fast)
template <bool HALF, bool SHIFTFOLD, bool INV>
__global__ void __launch_bounds__(FT, 3) ntt_batch_tma2(const FastTables T, const uint32_t Q, const uint2 one2, uint32_t* polys,
                                                        const uint64_t pairs) {
    constexpr bool inverse = INV;
    extern __shared__ __align__(128) uint32_t sm[];
    uint32_t* ring = sm;
    uint32_t* E = ring + NTT2_STAGES * 2 * FN;
    uint64_t* full = reinterpret_cast<uint64_t*>(E + 2 * FNP);
    const uint32_t tid = threadIdx.x;
    constexpr uint32_t BYTES = 2 * FN * 4;
    const uint64_t mine = blockIdx.x < pairs ? (pairs - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < NTT2_STAGES; ++s) mbar_init(full + s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (uint64_t i = 0; i < mine && i < NTT2_STAGES - 1; ++i) {
            mbar_expect_tx(full + i, BYTES);
            bulk_load(ring + i * 2 * FN, polys + (blockIdx.x + i * gridDim.x) * 2 * FN, BYTES, full + i);
        }
    }
    __syncthreads();
    for (uint64_t i = 0; i < mine; ++i) {
        const uint32_t s = uint32_t(i % NTT2_STAGES), parity = uint32_t(i / NTT2_STAGES) & 1u;
        uint32_t* buf = ring + s * 2 * FN;
        mbar_wait(full + s, parity);
        uint32_t x[2][8];
        if (!inverse) {
#pragma unroll
            for (int p = 0; p < 2; ++p)
#pragma unroll
                for (int r = 0; r < 8; ++r) x[p][r] = buf[p * FN + tid + 256 * r];
            fwd_ntt_regs<2, false, HALF>(x, E, T, Q, tid);
#pragma unroll
            for (int p = 0; p < 2; ++p)
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint4 v;
                    if (SHIFTFOLD) {
                        v.x = fold1(fold_shift27(x[p][4 * c + 0], Q), Q); v.y = fold1(fold_shift27(x[p][4 * c + 1], Q), Q);
                        v.z = fold1(fold_shift27(x[p][4 * c + 2], Q), Q); v.w = fold1(fold_shift27(x[p][4 * c + 3], Q), Q);
                    } else {
                        v.x = fold1(shoup_lazy(x[p][4 * c + 0], one2, Q), Q); v.y = fold1(shoup_lazy(x[p][4 * c + 1], one2, Q), Q);
                        v.z = fold1(shoup_lazy(x[p][4 * c + 2], one2, Q), Q); v.w = fold1(shoup_lazy(x[p][4 * c + 3], one2, Q), Q);
                    }
                    *reinterpret_cast<uint4*>(buf + p * FN + 8 * tid + 4 * c) = v;
                }
        } else {
#pragma unroll
            for (int p = 0; p < 2; ++p)
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const uint4 v = *reinterpret_cast<const uint4*>(buf + p * FN + 8 * tid + 4 * c);
                    x[p][4 * c] = v.x; x[p][4 * c + 1] = v.y; x[p][4 * c + 2] = v.z; x[p][4 * c + 3] = v.w;
                }
            inv_ntt_regs_np<2, HALF>(x, E, T, Q, tid);
#pragma unroll
            for (int p = 0; p < 2; ++p)
#pragma unroll
                for (int r = 0; r < 8; ++r) buf[p * FN + tid + 256 * r] = fold1(x[p][r], Q);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            bulk_store(polys + (blockIdx.x + i * gridDim.x) * 2 * FN, buf, BYTES);
            const uint64_t nx = i + NTT2_STAGES - 1;
            if (nx < mine) {
                if (i > 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                const uint32_t ns = uint32_t(nx % NTT2_STAGES);
                mbar_expect_tx(full + ns, BYTES);
                bulk_load(ring + ns * 2 * FN, polys + (blockIdx.x + nx * gridDim.x) * 2 * FN, BYTES, full + ns);
            }
        }
    }
    if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// ---- butterfly micro-benchmark: the measured peak the gate kernel's roofline is quoted against -----------------------------
// Register-resident lazy Shoup butterflies (the exact instruction sequence of the transforms: IMAD.HI + 2 IMAD + 2 IADD3), 16
// values per thread, 4 radix-2 stages per iteration, twiddles in registers, no memory traffic: what the integer pipes of
// this GPU sustain when they do nothing but modular butterflies.
__global__ void __launch_bounds__(256) butterfly_peak_kernel(uint32_t* sink, uint32_t iters, uint32_t Q, uint32_t zero, uint32_t seed) {
    struct { uint32_t zero; } T{zero};
    const uint32_t twoQ = 2 * Q;
    uint32_t x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = (threadIdx.x * 2654435761u + i * 40503u + seed) % Q;
    uint2 w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t v = ((threadIdx.x + i + seed) * 2246822519u) % Q;
        w[i] = make_uint2(v, uint32_t((uint64_t(v) << 32) / Q));
    }
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int st = 0; st < 4; ++st) {
            const int h = 8 >> st;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const int lo = (b / h) * 2 * h + (b % h);
                MKFHE_BFLY(x[lo], x[lo + h], w[st]);
            }
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] &= 0x07ffffffu;  // keep the lazy values bounded: 16 alu ops per 32 butterflies
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) r ^= x[i];
    if (r == 0x12345678u) sink[0] = r;
}

// ---- 16 coefficients per thread: the batched transform with TWO exchanges instead of three ------------------------------
// 128 threads per polynomial, passes of 4 + 4 + 3 radix-2 stages (forward; 3 + 4 + 4 inverse).  One exchange crosses warps,
// the other stays inside a warp (warp w owns [512 w, 512 w + 512) on both sides).  Exchange index: pad16(j) = j + 4 (j >> 5)
// + 8 (j >> 8), which makes the three access patterns "base + immediate" and bank-conflict free:
//   A  j = tid + 128 r           -> (tid + 4 (tid >> 5)) + 152 r + 8 (r >> 1) ... consecutive lanes, consecutive words
//   B  j = 128 h + 8 r + l3      -> h = tid >> 3, l3 = tid & 7: banks 16 h + 8 (h >> 1) + l3 (mod 32): {0,16,8,24} + l3
//   C  j = 16 tid + r (128-bit)  -> a quarter warp covers 8 distinct 16-byte bank groups
// Used by the NTT micro-benchmark entry (mkfhe_ntt, N = 2048); the gate kernel keeps the 8-coefficient transforms this round.
constexpr uint32_t FT16 = 128, FNP16 = 2368;
__device__ __forceinline__ constexpr uint32_t pad16(uint32_t j) { return j + 4 * (j >> 5) + 8 * (j >> 8); }
struct ExIdx16 {
    uint32_t a, b, c;
    __device__ __forceinline__ explicit ExIdx16(uint32_t tid)
        : a(tid + 4 * (tid >> 5)), b(pad16(128 * (tid >> 3)) + (tid & 7)), c(pad16(16 * tid)) {}
    // A: j = tid + 128 r: j >> 5 = (tid >> 5) + 4 r, j >> 8 = r >> 1
    __device__ __forceinline__ uint32_t A(int r) const { return a + 128 * r + 16 * r + 8 * (r >> 1); }
    // B: j = 128 h + 8 r + l3: j >> 5 = 4 h + (r >> 2), j >> 8 = h >> 1 (in the base)
    __device__ __forceinline__ uint32_t B(int r) const { return b + 8 * r + 4 * (r >> 2); }
    // C: j = 16 tid + r: j >> 5 = tid >> 1, j >> 8 = tid >> 4 (both in the base)
    __device__ __forceinline__ uint32_t C(int r) const { return c + r; }
};

// forward: in x[r] = coefficient tid + 128 r (< 2Q), out x[r] = slot 16 tid + r (< 24Q).  E: FNP16 words.
__device__ __forceinline__ void fwd_ntt16(uint32_t (&x)[16], uint32_t* E, const FastTables& T, uint32_t Q, uint32_t tid) {
    const uint32_t twoQ = 2 * Q;
    const ExIdx16 ex(tid);
    const uint32_t h4 = tid >> 3;
    // pass 1: t = 1024, 512, 256, 128; block of coefficient tid + 128 r at stage s is r >> (4 - s): the same in every thread
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int h = 8 >> s;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            if (r & h) continue;
            MKFHE_BFLY(x[r], x[r + h], T.twB[(1 << s) + (r >> (4 - s))]);
        }
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) E[ex.A(r)] = x[r];
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 16; ++r) x[r] = E[ex.B(r)];
    // pass 2: t = 64, 32, 16, 8 (m = 16 << s); block = (h4 << s) + (r >> (4 - s))
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int h = 8 >> s;
        const uint2* tws = T.tw2 + (16 << s) + (h4 << s);
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            if (r & h) continue;
            const uint2 w = __ldg(tws + (r >> (4 - s)));
            MKFHE_BFLY(x[r], x[r + h], w);
        }
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 16; ++r) E[ex.B(r)] = x[r];
    __syncwarp();
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const uint4 v = *reinterpret_cast<const uint4*>(E + ex.C(4 * c));
        x[4 * c] = v.x; x[4 * c + 1] = v.y; x[4 * c + 2] = v.z; x[4 * c + 3] = v.w;
    }
    // pass 3: t = 4, 2, 1 (m = 256 << s); block = (tid << (1 + s)) + (r >> (3 - s)): 2 + 4 + 8 twiddles, contiguous per thread,
    // read as seven coalesced 128-bit words from the packed table
    uint2 w[14];
#pragma unroll
    for (int i = 0; i < 7; ++i) {
        const uint4 v = __ldg(T.tw16 + 128 * i + tid);
        w[2 * i] = make_uint2(v.x, v.y);
        w[2 * i + 1] = make_uint2(v.z, v.w);
    }
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const int h = 4 >> s;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            if (r & h) continue;
            MKFHE_BFLY(x[r], x[r + h], w[(2 << s) - 2 + (r >> (3 - s))]);
        }
    }
}

// inverse: in x[r] = slot 16 tid + r (< 6Q), out x[r] = coefficient tid + 128 r in [0, 2Q).  Same ring map as inv_ntt_regs.
__device__ __forceinline__ void inv_ntt16(uint32_t (&x)[16], uint32_t* E, const FastTables& T, uint32_t Q, uint32_t tid) {
    const uint32_t twoQ = 2 * Q, sixQ = 6 * Q;
    const ExIdx16 ex(tid);
    const uint32_t l3 = tid & 7;
    // pass 1: t = 1 (twiddle 1), 2, 4: twiddle itw[t + (r mod t)], the same in every thread
#pragma unroll
    for (int r = 0; r < 16; r += 2) {
        const uint32_t u = x[r], v = x[r + 1];
        x[r] = u + v;
        x[r + 1] = u + sixQ - v;
    }
#pragma unroll
    for (int s = 1; s < 3; ++s) {
        const int h = 1 << s;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            if (r & h) continue;
            MKFHE_BFLY(x[r], x[r + h], T.itwB[h + (r & (h - 1))]);
        }
    }
    __syncwarp();
#pragma unroll
    for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(E + ex.C(4 * c)) = make_uint4(x[4 * c], x[4 * c + 1], x[4 * c + 2], x[4 * c + 3]);
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 16; ++r) x[r] = E[ex.B(r)];
    // pass 2: t = 8, 16, 32, 64 (h = t / 8 in register units); m = (j mod t) = 8 (r mod h) + l3
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int h = 1 << s;
        const uint2* tws = T.itw2 + 8 * h + l3;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            if (r & h) continue;
            const uint2 w = __ldg(tws + 8 * (r & (h - 1)));
            MKFHE_BFLY(x[r], x[r + h], w);
        }
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) E[ex.B(r)] = x[r];
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 16; ++r) x[r] = E[ex.A(r)];
    // pass 3: t = 128, 256, 512, 1024 (h = t / 128); m = 128 (r mod h) + tid
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int h = 1 << s;
        const uint2* tws = T.itw2 + 128 * h + tid;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            if (r & h) continue;
            const uint2 w = __ldg(tws + 128 * (r & (h - 1)));
            MKFHE_BFLY(x[r], x[r + h], w);
        }
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) x[r] = shoup_lazy(x[r], __ldg(T.scale2 + tid + 128 * r), Q);
}

__global__ void __launch_bounds__(FT16, 4) ntt_batch_fast16(const FastTables T, const uint32_t Q, const uint2 one2, uint32_t* polys,
                                                           const uint64_t batch, const int inverse) {
    __shared__ __align__(16) uint32_t E[FNP16];
    const uint32_t tid = threadIdx.x;
    uint32_t nx[16];  // the next polynomial, fetched while the current one is transformed
    auto fetch = [&](uint64_t b) {
        const uint32_t* p = polys + b * FN;
        if (!inverse) {
#pragma unroll
            for (int r = 0; r < 16; ++r) nx[r] = __ldcs(p + tid + 128 * r);
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const uint4 v = __ldcs(reinterpret_cast<const uint4*>(p + 16 * tid + 4 * c));
                nx[4 * c] = v.x; nx[4 * c + 1] = v.y; nx[4 * c + 2] = v.z; nx[4 * c + 3] = v.w;
            }
        }
    };
    if (blockIdx.x < batch) fetch(blockIdx.x);
    for (uint64_t b = blockIdx.x; b < batch; b += gridDim.x) {
        uint32_t* p = polys + b * FN;
        uint32_t x[16];
#pragma unroll
        for (int r = 0; r < 16; ++r) x[r] = nx[r];
        if (b + gridDim.x < batch) fetch(b + gridDim.x);
        if (!inverse) {
            fwd_ntt16(x, E, T, Q, tid);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint4 v;
                v.x = fold1(shoup_lazy(x[4 * c + 0], one2, Q), Q);
                v.y = fold1(shoup_lazy(x[4 * c + 1], one2, Q), Q);
                v.z = fold1(shoup_lazy(x[4 * c + 2], one2, Q), Q);
                v.w = fold1(shoup_lazy(x[4 * c + 3], one2, Q), Q);
                __stcs(reinterpret_cast<uint4*>(p + 16 * tid + 4 * c), v);
            }
        } else {
            inv_ntt16(x, E, T, Q, tid);
#pragma unroll
            for (int r = 0; r < 16; ++r) __stcs(p + tid + 128 * r, fold1(x[r], Q));
        }
        __syncthreads();  // E is reused by the next polynomial
    }
}

// ---- bootstrapping-key generation on the register-resident transforms (N = 2048; semantics of keygen_bk_kernel) ----------
// One CTA per evk: NTT(r), then per digit NTT(e0_j), NTT(e1_j) as one two-polynomial transform, canonicalised, and the
// pointwise assembly on the 8 slots the thread holds; the evk leaves as 128-bit stores.  R[j] (slot j of NTT(r), the
// reference's scalar "skrPoly[i]") is published through shared memory by thread 0, which holds slots 0..7.
__global__ void __launch_bounds__(FT, 2) keygen_bk_fast(const DevCtx P, const FastTables T, const uint2 one2, const GadgetPow G,
                                                        const uint32_t* __restrict__ sinv_m, const uint32_t* __restrict__ crs_m,
                                                        const uint32_t* __restrict__ r, const uint32_t* __restrict__ e,
                                                        const uint8_t* __restrict__ m, uint32_t* __restrict__ out) {
    __shared__ __align__(16) uint32_t E[2 * FNP];
    __shared__ uint32_t Rlow[8];
    constexpr uint32_t N = FN;
    const uint32_t Q = P.mod.Q, ell = P.ell, n = P.n, tid = threadIdx.x;
    const uint32_t ev = blockIdx.x, last = gridDim.x - 1;
    uint32_t u = 0;
    bool kdm = true;
    if (ev != last) {
        u = ev / (P.S * n);
        kdm = (u == 0 && (ev - u * P.S * n) % n == 0);
    }
    const bool msg = m[ev] != 0;
    uint32_t R[1][8];
#pragma unroll
    for (int q = 0; q < 8; ++q) R[0][q] = r[size_t(ev) * N + tid + 256 * q];
    fwd_ntt_regs<1>(R, E, T, Q, tid);
#pragma unroll
    for (int i = 0; i < 8; ++i) R[0][i] = fold1(shoup_lazy(R[0][i], one2, Q), Q);
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) Rlow[i] = R[0][i];
    }
    uint32_t sv[8];
    {
        const uint4 a = ldg4(sinv_m + size_t(u) * N + 8 * tid), b = ldg4(sinv_m + size_t(u) * N + 8 * tid + 4);
        sv[0] = a.x; sv[1] = a.y; sv[2] = a.z; sv[3] = a.w; sv[4] = b.x; sv[5] = b.y; sv[6] = b.z; sv[7] = b.w;
    }
    __syncthreads();  // Rlow visible; E free again
    for (uint32_t j = 0; j < ell; ++j) {
        uint32_t X[2][8];
        const uint32_t* ee = e + (size_t(ev) * ell + j) * 2 * N;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            X[0][q] = ee[tid + 256 * q];
            X[1][q] = ee[N + tid + 256 * q];
        }
        fwd_ntt_regs<2>(X, E, T, Q, tid);
        const uint32_t Rj = Rlow[j];
        const uint32_t* cj = crs_m + size_t(j) * N + 8 * tid;
        const uint4 c0 = ldg4(cj), c1 = ldg4(cj + 4);
        const uint32_t cv[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        uint32_t d[8], f[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t e0 = fold1(shoup_lazy(X[0][i], one2, Q), Q), e1 = fold1(shoup_lazy(X[1][i], one2, Q), Q);
            f[i] = montmul(addmod(e1, montmul(R[0][i], G.mont[j], P.mod), Q), sv[i], P.mod);
            uint32_t dd = e0;
            if (msg) dd = addmod(dd, kdm ? montmul(G.plain[j], sv[i], P.mod) : G.plain[j], Q);
            d[i] = addmod(dd, montmul(Rj, cv[i], P.mod), Q);
        }
        uint32_t* o = out + (size_t(ev) * ell + j) * 2 * N + 8 * tid;
        *reinterpret_cast<uint4*>(o) = make_uint4(d[0], d[1], d[2], d[3]);
        *reinterpret_cast<uint4*>(o + 4) = make_uint4(d[4], d[5], d[6], d[7]);
        *reinterpret_cast<uint4*>(o + N) = make_uint4(f[0], f[1], f[2], f[3]);
        *reinterpret_cast<uint4*>(o + N + 4) = make_uint4(f[4], f[5], f[6], f[7]);
        __syncthreads();  // E is reused by the next digit's transform
    }
}

// ---- host side ------------------------------------------------------------------------------------------------------
struct FastState {
    FastTables T{};
    uint2 *d_tw2{nullptr}, *d_itw2{nullptr}, *d_scale2{nullptr};
    uint4* d_tw4{nullptr};
    uint4* d_tw16{nullptr};
    uint2 one2{};
    int sms{148};
    int max_clusters[2][3] = {{-1, -1, -1}, {-1, -1, -1}};  // cluster_kernels.cuh: co-resident clusters of 2 / 4 / 8 CTAs, [gate][size]
    uint32_t* accws{nullptr};  // blind_rotate_fast_ag: [sms * 4][k][N] accumulator slices, allocated by fast_init when MKFHE_AG=1
    bool half{false};  // tables for N = 1024 (two polynomials per 2048-word pass): only the batched transform uses them
};

// accumulators in global memory (blind_rotate_fast_ag): an experiment, selected with MKFHE_AG=1 at context creation (the
// workspace is allocated then) -- measured on B200 (profiles/r2_s18_ag.txt, gates/s, shared-memory kernels vs this one):
//   k = 16  boolean-mklwe 436 vs 446 (+2%)   boolean-mkntru 343 vs 329 (4 per SM, 64 registers) / 303 (3 per SM, 80)
//   k = 8   boolean-mklwe 1740 vs 1695       boolean-mkntru 1368 vs 1248
//   k = 4   boolean-mklwe 6685 vs 6120       boolean-mkntru 4979 vs 4286         k = 2 boolean-mklwe 22.09 k vs 20.46 k
// Four times the warps at k = 16 buy back what the L2 round trips of the accumulator rows (and the 64-bit addresses: 112-212 B of
// spills in the step loop) cost, no more; at equal occupancy (k = 2) the global accumulator costs 8%.  Off by default.
inline bool fast_ag_wanted(const DevCtx&) {
    const char* e = getenv("MKFHE_AG");
    return e != nullptr && e[0] == '1';
}

inline bool fast_supported(const DevCtx& d) {
    return d.N == FN && d.ell == 2 && d.mod.Q < (1u << 27) && d.gbits >= 2 && d.gbits <= 16 &&
           fast_smem_bytes(d.k, d.n, true, d.variant == 0) <= 227 * 1024;
}

// N = 1024 with a modulus the lazy butterflies can take: the batched transform (mkfhe_ntt, config 5) runs on the HALF path
inline bool fast_ntt_half_supported(const DevCtx& d) { return d.N == FN / 2 && d.mod.Q < (1u << 27); }

inline cudaError_t fast_init(FastState& F, const DevCtx& d, uint32_t psi, cudaStream_t st, uint64_t* dev_bytes) {
    const uint32_t Q = d.mod.Q, N = FN;
    const bool half = d.N == FN / 2;
    const uint32_t NT = half ? FN / 2 : FN, LOGT = half ? FLOGN - 1 : FLOGN;  // size of the transform the tables describe
    auto mul = [&](uint32_t a, uint32_t b) { return uint32_t((uint64_t(a) * b) % Q); };
    auto pw = [&](uint32_t a, uint64_t e) {
        uint64_t r = 1, b = a % Q;
        for (; e; e >>= 1, b = b * b % Q)
            if (e & 1) r = r * b % Q;
        return uint32_t(r);
    };
    auto pair = [&](uint32_t w) { return make_uint2(w, uint32_t((uint64_t(w) << 32) / Q)); };
    auto brv = [&](uint32_t x) {
        uint32_t r = 0;
        for (uint32_t i = 0; i < LOGT; ++i) r |= ((x >> i) & 1u) << (LOGT - 1 - i);
        return r;
    };
    const uint32_t psi_inv = pw(psi, uint64_t(Q) - 2), omega_inv = mul(psi_inv, psi_inv), n_inv = pw(NT, uint64_t(Q) - 2);
    std::vector<uint2> tw(N), itw(N), sc(N);
    uint32_t x = 1, xi = n_inv;
    for (uint32_t i = 0; i < NT; ++i) {
        tw[brv(i)] = pair(x);
        sc[i] = pair(xi);
        x = mul(x, psi);
        xi = mul(xi, psi_inv);
    }
    if (half) {
        // 2048-word view of two size-1024 transforms: stage with m blocks (m = 2 .. 1024) = the size-1024 stage with m/2
        // blocks in each half, block b of the view = block b mod m/2 of its half; scale by coefficient index mod 1024
        std::vector<uint2> t1(tw.begin(), tw.begin() + NT);
        for (uint32_t m = 2; m < N; m <<= 1)
            for (uint32_t b = 0; b < m; ++b) tw[m + b] = t1[m / 2 + (b & (m / 2 - 1))];
        tw[0] = tw[1] = pair(1);
        for (uint32_t i = 0; i < NT; ++i) sc[NT + i] = sc[i];
    }
    itw[0] = pair(1);
    for (uint32_t t = 1; t < N; t <<= 1) {
        if (t >= NT) {  // HALF: no stage joins the halves; entries unused
            for (uint32_t m = 0; m < t; ++m) itw[t + m] = pair(1);
            continue;
        }
        const uint32_t step = pw(omega_inv, NT / (2 * t));  // omega^-(NT/2t): a primitive 2t-th root
        uint32_t v = 1;
        for (uint32_t m = 0; m < t; ++m) {
            itw[t + m] = pair(v);
            v = mul(v, step);
        }
    }
    cudaError_t e;
    for (auto pr : {std::make_pair(&F.d_tw2, &tw), std::make_pair(&F.d_itw2, &itw), std::make_pair(&F.d_scale2, &sc)}) {
        if ((e = cudaMalloc(reinterpret_cast<void**>(pr.first), N * sizeof(uint2))) != cudaSuccess) return e;
        *dev_bytes += N * sizeof(uint2);
        if ((e = cudaMemcpyAsync(*pr.first, pr.second->data(), N * sizeof(uint2), cudaMemcpyHostToDevice, st)) != cudaSuccess) return e;
    }
    std::vector<uint4> t4(3 * 256);
    for (uint32_t t = 0; t < 256; ++t) {
        auto q4 = [&](uint32_t i) { return make_uint4(tw[i].x, tw[i].y, tw[i + 1].x, tw[i + 1].y); };
        t4[t] = q4(512 + 2 * t);
        t4[256 + t] = q4(1024 + 4 * t);
        t4[512 + t] = q4(1026 + 4 * t);
    }
    if ((e = cudaMalloc(reinterpret_cast<void**>(&F.d_tw4), t4.size() * sizeof(uint4))) != cudaSuccess) return e;
    *dev_bytes += t4.size() * sizeof(uint4);
    if ((e = cudaMemcpyAsync(F.d_tw4, t4.data(), t4.size() * sizeof(uint4), cudaMemcpyHostToDevice, st)) != cudaSuccess) return e;
    std::vector<uint4> t16(7 * 128);
    for (uint32_t t = 0; t < 128; ++t) {
        auto q4 = [&](uint32_t i) { return make_uint4(tw[i].x, tw[i].y, tw[i + 1].x, tw[i + 1].y); };
        t16[t] = q4(256 + 2 * t);
        for (uint32_t i = 0; i < 2; ++i) t16[128 * (1 + i) + t] = q4(512 + 4 * t + 2 * i);
        for (uint32_t i = 0; i < 4; ++i) t16[128 * (3 + i) + t] = q4(1024 + 8 * t + 2 * i);
    }
    if ((e = cudaMalloc(reinterpret_cast<void**>(&F.d_tw16), t16.size() * sizeof(uint4))) != cudaSuccess) return e;
    *dev_bytes += t16.size() * sizeof(uint4);
    if ((e = cudaMemcpyAsync(F.d_tw16, t16.data(), t16.size() * sizeof(uint4), cudaMemcpyHostToDevice, st)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
    F.T.tw16 = F.d_tw16;
    F.T.tw2 = F.d_tw2; F.T.itw2 = F.d_itw2; F.T.scale2 = F.d_scale2; F.T.tw4 = F.d_tw4;
    for (int i = 0; i < 8; ++i) F.T.twA[i] = tw[i];
    for (int i = 0; i < 4; ++i) F.T.itwA[i] = itw[i];
    for (int i = 0; i < 64; ++i) F.T.twB[i] = tw[i];
    for (int i = 0; i < 32; ++i) F.T.itwB[i] = itw[i];
    F.T.zero = 0;
    F.one2 = pair(1);
    F.half = half;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&F.sms, cudaDevAttrMultiProcessorCount, dev);
    if (!half && fast_ag_wanted(d)) {  // at set-up, so that the gate calls never allocate
        const size_t bytes = size_t(F.sms) * 4 * d.k * FN * sizeof(uint32_t);
        if ((e = cudaMalloc(reinterpret_cast<void**>(&F.accws), bytes)) != cudaSuccess) return e;
        *dev_bytes += bytes;
    }
    return cudaSuccess;
}
inline void fast_destroy(FastState& F) {
    if (F.accws) cudaFree(F.accws);
    for (uint2* p : {F.d_tw2, F.d_itw2, F.d_scale2})
        if (p) cudaFree(p);
    if (F.d_tw4) cudaFree(F.d_tw4);
    if (F.d_tw16) cudaFree(F.d_tw16);
    F = FastState{};
}
// Key polynomials -> register order, in place, one CTA per polynomial: thread t of a gate holds the slots 8t .. 8t+7 after
// a forward transform and multiplies them by the key in two chunks of 4; chunk ch of thread t moves to words 1024 ch + 4 t,
// so each 128-bit key load of a warp reads 512 contiguous bytes (slot order would read half of every 32-byte sector twice).
__global__ void __launch_bounds__(FT) fast_permute_keys(uint32_t* polys) {
    uint32_t* p = polys + size_t(blockIdx.x) * FN;
    const uint32_t t = threadIdx.x;
    const uint4 c0 = *reinterpret_cast<const uint4*>(p + 8 * t), c1 = *reinterpret_cast<const uint4*>(p + 8 * t + 4);
    __syncthreads();
    *reinterpret_cast<uint4*>(p + 4 * t) = c0;
    *reinterpret_cast<uint4*>(p + FN / 2 + 4 * t) = c1;
}
inline cudaError_t fast_permute(uint32_t* polys, uint64_t words, cudaStream_t st, uint64_t* launches) {
    fast_permute_keys<<<unsigned(words / FN), FT, 0, st>>>(polys);
    ++*launches;
    return cudaGetLastError();
}
inline cudaError_t fast_load_bk(FastState&, const DevCtx& d, uint64_t words, cudaStream_t st, uint64_t*, uint64_t* launches) {
    return fast_permute(const_cast<uint32_t*>(d.bk), words, st, launches);
}
inline cudaError_t fast_load_pkey(FastState&, const DevCtx& d, uint64_t words, cudaStream_t st, uint64_t*, uint64_t* launches) {
    return fast_permute(const_cast<uint32_t*>(d.pkey), words, st, launches);
}

// gates resident per SM: 64 registers x 256 threads allow 4; shared memory (1 KiB reserved per CTA) may allow fewer
inline int fast_gates_per_sm(size_t smem) { return int(std::min<size_t>(4, (227 * 1024) / (smem + 1024))); }
// ALIAS (the inverse transform exchanges through E1[0]; no buffer of its own: 9 KiB less per gate).  Round 1 paid a block barrier
// per iteration for it and measured it slower (k = 4: 3783 vs 4246 gates/s); the barrier was never needed (fast_step), and
// without it the layout is free.  It is chosen when the separate buffer does not fit at all, and -- for the XZW_B flavour,
// whose 64-register build does not spill -- when it lets one more gate share an SM and the batch has gates for it:
// boolean-mklwe k = 4, 888 gates: 4 per SM aliased 6436 gates/s, 3 per SM with the separate buffer 6034 (6198 with paired
// inverses); profiles/r2_s14_pair_alias.txt.  The XZW flavour at 64 registers spills and stays at 3 per SM.
inline bool fast_use_alias(const DevCtx& d, uint32_t batch, int sms) {
    if (const char* e = getenv("MKFHE_FAST_ALIAS")) return e[0] == '1';
    const bool xzw = d.variant == 0;
    const size_t plain = fast_smem_bytes(d.k, d.n, false, xzw), aliased = fast_smem_bytes(d.k, d.n, true, xzw);
    if (plain > 227 * 1024) return true;
    const int g0 = fast_gates_per_sm(plain), g1 = fast_gates_per_sm(aliased);
    return !xzw && g1 > g0 && uint64_t(batch) > uint64_t(g0) * uint64_t(sms);
}

// party-parallel groups per gate, overridable for experiments with MKFHE_GROUPS.  Measured on B200, gates/s
// (profiles/r1_s34_k16_groups.txt, r1_s17_grouped_quick.txt):   k = 16: 1 group 281, 2 groups 350, 4 groups 306 (1 gate per SM
// in all three: 8 / 16 / 32 warps, 128 / 128 / 64 registers);   k = 8: 1398 vs 1252 (2 groups);   k = 4: 4246 vs 3013 -- where
// several gates fit an SM anyway, the idle groups of the sumV phase cost more than the shorter step brings.
// So: two groups from k = 16 on.
// Small batches are a different regime: with at most one gate per SM the SM is latency-bound, and the shorter step wins
// (profiles/r1_s49_latency_groups.txt, ms per launch at 148 gates, 1 group vs 2: boolean-mklwe 11.29 vs 9.31, boolean-mkntru
// 13.80 vs 12.71, k = 4 43.2 vs 37.0; at 296 gates one group is ahead again, 16.1 vs 18.6).  So: two groups whenever the batch
// does not exceed the SM count -- the single gate of the drop-in accumulator and the narrow levels of a circuit.
inline uint32_t fast_groups(const DevCtx& d, uint32_t batch, int sms) {
    uint32_t G = (d.k >= 16 || (d.k >= 2 && batch <= uint32_t(sms))) ? 2 : 1;
    if (const char* e = getenv("MKFHE_GROUPS")) G = uint32_t(atoi(e));
    if (G < 2 || G > 4) return 1;
    if (d.k > 16 || d.k < G) return 1;
    return grouped_smem_bytes(d.k, d.n, G, true) <= 227 * 1024 ? G : 1;
}

inline cudaError_t fast_blind_rotate(FastState& F, const DevCtx& d, bool gate, const GateIO& io, uint32_t batch, cudaStream_t st,
                                     uint64_t* launches) {
    const bool xzw = d.variant == 0;
    // large party counts, more gates than SMs: the accumulators move to L2 and four (XZW: three, as at k = 2) gates share an SM
    if (F.accws != nullptr && fast_ag_wanted(d) && batch > uint32_t(F.sms)) {
        F.T.prefetch = 0u;
        F.T.pair = 0u;
        int minb = xzw ? 3 : 4;
        if (const char* ev = getenv("MKFHE_MINB")) minb = std::min(4, std::max(3, atoi(ev)));
        const size_t smem = fast_ag_smem_bytes(d.k, d.n, xzw);
        const uint32_t grid = std::min<uint32_t>(batch, uint32_t(F.sms) * uint32_t(minb));
        auto launch = [&](auto kern) -> cudaError_t {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            if (e != cudaSuccess) return e;
            kern<<<grid, FT, smem, st>>>(d, F.T, io, F.accws, batch);
            return cudaSuccess;
        };
        cudaError_t e;
        if (minb >= 4)
            e = gate ? (xzw ? launch(blind_rotate_fast_ag<true, true, 4>) : launch(blind_rotate_fast_ag<true, false, 4>))
                     : (xzw ? launch(blind_rotate_fast_ag<false, true, 4>) : launch(blind_rotate_fast_ag<false, false, 4>));
        else
            e = gate ? (xzw ? launch(blind_rotate_fast_ag<true, true, 3>) : launch(blind_rotate_fast_ag<true, false, 3>))
                     : (xzw ? launch(blind_rotate_fast_ag<false, true, 3>) : launch(blind_rotate_fast_ag<false, false, 3>));
        if (e != cudaSuccess) return e;
        ++*launches;
        return cudaGetLastError();
    }
    const uint32_t G = fast_groups(d, batch, F.sms);
    F.T.stagger_ns = getenv("MKFHE_STAGGER_NS") ? uint32_t(atoi(getenv("MKFHE_STAGGER_NS"))) : 0u;
    // key prefetch: on in latency mode (at most one gate per SM, k <= 8), where the tile comes from HBM and nothing hides the
    // wait: 1, 16 gates boolean-mklwe 8.88 -> 8.75, 9.02 -> 8.74 ms; boolean-mkntru 12.22 -> 11.93 ms; k = 4 36.9 -> 36.2 ms;
    // no gain from 148 gates on and -0.7% at k = 16 (profiles/r2_s6_latency.txt).  MKFHE_PREFETCH=0/1 forces it for A/B runs
    F.T.prefetch = (batch <= uint32_t(F.sms) && d.k <= 8) ? 1u : 0u;
    if (const char* pf = getenv("MKFHE_PREFETCH")) F.T.prefetch = pf[0] == '1' ? 1u : 0u;
    F.T.pair = 1u;  // (resolved below, once the register budget is known)
    const char* pair_env = getenv("MKFHE_PAIR");
    const bool alias = G > 1 ? grouped_smem_bytes(d.k, d.n, G, false) > 227 * 1024 : fast_use_alias(d, batch, F.sms);
    const size_t smem = G > 1 ? grouped_smem_bytes(d.k, d.n, G, alias) : fast_smem_bytes(d.k, d.n, alias, xzw);
    auto launch = [&](auto kern) -> cudaError_t {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        if (e != cudaSuccess) return e;
        kern<<<batch, FT * G, smem, st>>>(d, F.T, io);
        return cudaSuccess;
    };
    cudaError_t e;
    const int sel = (gate ? 4 : 0) | (xzw ? 2 : 0) | (alias ? 1 : 0);
    int minb = G > 1 ? 1 : std::max(2, fast_gates_per_sm(smem));
    // XZW at 64 registers spills in the step loop; 3 gates per SM with 80 registers measured 3% faster than 4 with 64
    // (16.66k vs 16.17k gates/s, boolean-mkntru k = 2, profiles/r1_s28_xzw_quick.txt); XZW_B measured equal, stays at 4
    // A thin last wave does not change this: 4096 gates are 10 waves of 444 at 3 per SM (the last 22% full) against 7 waves of
    // 592 at 4, yet 3 per SM measured 242.4 ms and 4 per SM 247.6 ms (profiles/r2_s2_bench.json, r2_s11_mkntru4096.txt) -- the
    // gates of a thin wave have their SMs to themselves and finish in half a wave time.
    if (xzw && minb > 3) minb = 3;
    // a batch that puts fewer gates on an SM than the budget allows gets the registers of the absent ones: ms per launch at
    // 296 gates, 64 / 80 / 128 registers: 16.14 / 15.57 / 14.91 (boolean-mklwe); at 444: 22.31 / 21.56 / 22.72
    // (profiles/r1_s57_minb_midrange.txt)
    minb = std::min(minb, std::max(2, int((batch + uint32_t(F.sms) - 1) / uint32_t(F.sms))));
    if (const char* ev = getenv("MKFHE_MINB")) minb = std::min(4, std::max(2, atoi(ev)));
    if (alias) minb = 4;  // the fallback layout is built for the 64-register budget only
    if (G > 1) {
#define MKFHE_LAUNCH_G(G_, X_, A_)                                                                                         \
    (G == 2 ? launch(blind_rotate_grouped<G_, X_, A_, 512>)                                                                 \
            : G == 3 ? launch(blind_rotate_grouped<G_, X_, A_, 768>) : launch(blind_rotate_grouped<G_, X_, A_, 1024>))
        switch (sel) {
            case 0: e = MKFHE_LAUNCH_G(false, false, false); break;
            case 1: e = MKFHE_LAUNCH_G(false, false, true); break;
            case 2: e = MKFHE_LAUNCH_G(false, true, false); break;
            case 3: e = MKFHE_LAUNCH_G(false, true, true); break;
            case 4: e = MKFHE_LAUNCH_G(true, false, false); break;
            case 5: e = MKFHE_LAUNCH_G(true, false, true); break;
            case 6: e = MKFHE_LAUNCH_G(true, true, false); break;
            default: e = MKFHE_LAUNCH_G(true, true, true); break;
        }
#undef MKFHE_LAUNCH_G
    } else if (alias) {
        switch (sel >> 1) {
            case 0: e = launch(blind_rotate_fast<false, false, true, 4>); break;
            case 1: e = launch(blind_rotate_fast<false, true, true, 4>); break;
            case 2: e = launch(blind_rotate_fast<true, false, true, 4>); break;
            default: e = launch(blind_rotate_fast<true, true, true, 4>); break;
        }
    } else {
        // paired inverses: as kernels of their own they pay at k >= 8 only (128 registers, two gates per SM: boolean-mkntru k = 8
        // 1357 -> 1396 gates/s); unpaired is ahead everywhere else (ms per launch, unpaired / paired: boolean-mkntru 4096 gates
        // 241.8 / 248.5, k = 4 888 gates 176.7 / 179.5; boolean-mklwe 296 gates 14.97 / 15.33, 444 gates 21.71 / 22.13, k = 4 444
        // gates 71.9 / 72.8; profiles/r2_s28_pair_split.txt)
        bool pair = minb == 2 && d.k >= 8;
        if (pair_env) pair = pair_env[0] == '1' && minb < 4;
        F.T.pair = pair ? 1u : 0u;
#define MKFHE_LAUNCH_MINB(G_, X_)                                                                                                   \
    (minb >= 4 ? launch(blind_rotate_fast<G_, X_, false, 4>)                                                                        \
               : minb == 3 ? (pair ? launch(blind_rotate_fast<G_, X_, false, 3, true>) : launch(blind_rotate_fast<G_, X_, false, 3>)) \
                           : (pair ? launch(blind_rotate_fast<G_, X_, false, 2, true>) : launch(blind_rotate_fast<G_, X_, false, 2>)))
        switch (sel >> 1) {
            case 0: e = MKFHE_LAUNCH_MINB(false, false); break;
            case 1: e = MKFHE_LAUNCH_MINB(false, true); break;
            case 2: e = MKFHE_LAUNCH_MINB(true, false); break;
            default: e = MKFHE_LAUNCH_MINB(true, true); break;
        }
#undef MKFHE_LAUNCH_MINB
    }
    if (e != cudaSuccess) return e;
    ++*launches;
    return cudaGetLastError();
}

inline cudaError_t fast_keygen_bk(FastState& F, const DevCtx& d, const GadgetPow& G, const uint32_t* sinv_m, const uint32_t* crs_m,
                                  const uint32_t* r, const uint32_t* e, const uint8_t* m, uint32_t* out, uint32_t evks, cudaStream_t st,
                                  uint64_t* launches) {
    keygen_bk_fast<<<evks, FT, 0, st>>>(d, F.T, F.one2, G, sinv_m, crs_m, r, e, m, out);
    ++*launches;
    return cudaGetLastError();
}

inline cudaError_t fast_ntt(FastState& F, const DevCtx& d, uint32_t* polys, uint64_t batch, int inverse, cudaStream_t st,
                            uint64_t* launches) {
    // HALF: batch counts PAIRS of size-1024 polynomials (the caller sends an odd last one through the generic kernel)
    const unsigned grid = unsigned(std::min<uint64_t>(batch, uint64_t(F.sms) * 4));
    const char* var = getenv("MKFHE_NTT_VARIANT");
    // default: the bulk-copy ring with TWO polynomials per slot (ntt_batch_tma2: +2.4% / +3.6% at N = 2048 / 1024 over one per
    // slot, profiles/r2_s13_ntt*.json); MKFHE_NTT_VARIANT = 1: one polynomial per slot (ntt_batch_tma), b: the same with the
    // Barrett-style output fold, ldg: the register-prefetch kernel of round 1
    if ((!var || var[0] == '2' || var[0] == 't') && !getenv("MKFHE_NTT16") && batch >= 2) {  // (HALF: a slot holds two pairs)
        const uint64_t pairs = batch / 2;
        const size_t smem = ntt_tma2_smem_bytes();
        const unsigned g = unsigned(std::min<uint64_t>(pairs, uint64_t(F.sms) * 3));
        const bool sf = fold_shift27_ok(d.mod.Q);
        auto launch = [&](auto kern) -> cudaError_t {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            if (e != cudaSuccess) return e;
            kern<<<g, FT, smem, st>>>(F.T, d.mod.Q, F.one2, polys, pairs);
            return cudaSuccess;
        };
#define MKFHE_TMA2(H_, S_) (inverse ? launch(ntt_batch_tma2<H_, S_, true>) : launch(ntt_batch_tma2<H_, S_, false>))
        cudaError_t e = F.half ? (sf ? MKFHE_TMA2(true, true) : MKFHE_TMA2(true, false)) : (sf ? MKFHE_TMA2(false, true) : MKFHE_TMA2(false, false));
#undef MKFHE_TMA2
        if (e != cudaSuccess) return e;
        ++*launches;
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (batch & 1) {  // the odd last one: the one-polynomial kernel
            polys += (batch - 1) * FN;
            batch = 1;
        } else {
            return cudaSuccess;
        }
    }
    if (!getenv("MKFHE_NTT16") && !(var && var[0] == 'l')) {
        int ctas = 4;  // 4, 5 and 6 resident CTAs measured equal on B200 (profiles/r2_s3_experiments.txt): the kernel is not latency-bound
        if (const char* c = getenv("MKFHE_NTT_CTAS")) ctas = std::min(6, std::max(4, atoi(c)));
        const size_t smem = ntt_tma_smem_bytes(ctas);
        const unsigned g = unsigned(std::min<uint64_t>(batch, uint64_t(F.sms) * ctas));
        const bool sf = fold_shift27_ok(d.mod.Q) && !(var && var[0] == 'b');  // 'b': Barrett-style final fold, for A/B runs
        auto launch = [&](auto kern) -> cudaError_t {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            if (e != cudaSuccess) return e;
            kern<<<g, FT, smem, st>>>(F.T, d.mod.Q, F.one2, polys, batch);
            return cudaSuccess;
        };
#define MKFHE_NTT_TMA_D(H_, S_, I_) (ctas == 4 ? launch(ntt_batch_tma<H_, S_, 4, I_>) : ctas == 5 ? launch(ntt_batch_tma<H_, S_, 5, I_>) : launch(ntt_batch_tma<H_, S_, 6, I_>))
#define MKFHE_NTT_TMA(H_, S_) (inverse ? MKFHE_NTT_TMA_D(H_, S_, true) : MKFHE_NTT_TMA_D(H_, S_, false))
        cudaError_t e = F.half ? (sf ? MKFHE_NTT_TMA(true, true) : MKFHE_NTT_TMA(true, false))
                               : (sf ? MKFHE_NTT_TMA(false, true) : MKFHE_NTT_TMA(false, false));
#undef MKFHE_NTT_TMA
#undef MKFHE_NTT_TMA_D
        if (e != cudaSuccess) return e;
        ++*launches;
        return cudaGetLastError();
    }
    if (!F.half && getenv("MKFHE_NTT16")) {  // experimental: 16 coefficients per thread, two exchanges
        const unsigned g16 = unsigned(std::min<uint64_t>(batch, uint64_t(F.sms) * 8));
        ntt_batch_fast16<<<g16, FT16, 0, st>>>(F.T, d.mod.Q, F.one2, polys, batch, inverse);
    } else if (F.half) {
        if (inverse) ntt_batch_fast<true, true><<<grid, FT, 0, st>>>(F.T, d.mod.Q, F.one2, polys, batch);
        else ntt_batch_fast<true, false><<<grid, FT, 0, st>>>(F.T, d.mod.Q, F.one2, polys, batch);
    } else {
        if (inverse) ntt_batch_fast<false, true><<<grid, FT, 0, st>>>(F.T, d.mod.Q, F.one2, polys, batch);
        else ntt_batch_fast<false, false><<<grid, FT, 0, st>>>(F.T, d.mod.Q, F.one2, polys, batch);
    }
    ++*launches;
    return cudaGetLastError();
}

}  // namespace mkfhe

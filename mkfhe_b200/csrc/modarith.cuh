// modarith.cuh -- 32-bit modular arithmetic for sm_100a (Q prime, Q < 2^30; the parameter sets use 27-bit Q).
//
// Everything on the path is exact arithmetic in Z_Q, so any correct multiplication algorithm yields the reference's
// canonical residues (the reference uses Barrett `ModMulFastEq` ubintnat.h:1397-1410 and Shoup `ModMulFastConstEq`).
// Here: Montgomery multiplication with R = 2^32.  Constants (twiddles, keys) are stored in Montgomery form x*R mod Q,
// data stays in plain form, so montmul(data, constant) returns the plain product.
#pragma once
#include <cstdint>

namespace mkfhe {

struct Mod {
    uint32_t Q;        // modulus
    uint32_t qinv_neg; // -Q^{-1} mod 2^32
};

// a * b * 2^-32 mod Q, canonical.  Requires a * b < Q * 2^32 (one factor canonical, the other any 32-bit word).
__device__ __forceinline__ uint32_t montmul(uint32_t a, uint32_t b, const Mod m) {
    const uint64_t t = static_cast<uint64_t>(a) * b;
    const uint32_t k = static_cast<uint32_t>(t) * m.qinv_neg;
    const uint32_t r = static_cast<uint32_t>((t + static_cast<uint64_t>(k) * m.Q) >> 32);
    return r >= m.Q ? r - m.Q : r;
}
// same without the final conditional subtraction: result in [0, 2Q)
__device__ __forceinline__ uint32_t montmul_lazy(uint32_t a, uint32_t b, const Mod m) {
    const uint64_t t = static_cast<uint64_t>(a) * b;
    const uint32_t k = static_cast<uint32_t>(t) * m.qinv_neg;
    return static_cast<uint32_t>((t + static_cast<uint64_t>(k) * m.Q) >> 32);
}
// Montgomery reduction of a 64-bit accumulator T < Q * 2^32: T * 2^-32 mod Q, canonical
__device__ __forceinline__ uint32_t montredc(uint64_t t, const Mod m) {
    const uint32_t k = static_cast<uint32_t>(t) * m.qinv_neg;
    const uint32_t r = static_cast<uint32_t>((t + static_cast<uint64_t>(k) * m.Q) >> 32);
    return r >= m.Q ? r - m.Q : r;
}
__device__ __forceinline__ uint32_t addmod(uint32_t a, uint32_t b, uint32_t Q) {
    const uint32_t s = a + b;
    return s >= Q ? s - Q : s;
}
__device__ __forceinline__ uint32_t submod(uint32_t a, uint32_t b, uint32_t Q) {
    const uint32_t d = a - b;
    return a >= b ? d : d + Q;
}

// floor(0.5 + v*q/Q) mod q in IEEE double with the reference's evaluation order (mntru-pke.cpp:12-17, mklwe-pke.cpp:12-17).
// v*q < 2^53 is exact; the division is the only rounding; explicit _rn intrinsics forbid FMA contraction.
__device__ __forceinline__ uint32_t round_qQ(uint32_t v, uint32_t q, uint32_t Q) {
    const double prod = __dmul_rn(static_cast<double>(v), static_cast<double>(q));
    const double quo  = __ddiv_rn(prod, static_cast<double>(Q));
    const double r    = floor(__dadd_rn(0.5, quo));
    return static_cast<uint32_t>(static_cast<unsigned long long>(r) % q);
}

// One coefficient of UniEncAccumulator::SignedDigitDecompose (mk-acc.cpp:54-80) at NATIVEINT = 32:
// centre t to d in [-(Q+1)/2, Q/2), drop the lowest base-2^gbits digit, then emit signed digits r in [-B/2, B/2).
// The reference's (d << (32-g)) >> (32-g) is "sign-extend the low g bits"; (d - r) >> g is then an exact division.
struct Decomposer {
    int32_t d;
    int gbits;
    __device__ __forceinline__ Decomposer(uint32_t t, uint32_t Q, int g) : gbits(g) {
        d = static_cast<int32_t>(t < (Q >> 1) ? t : t - Q);
        (void)next();  // approximate gadget decomposition: the first digit is ignored (mk-acc.cpp:63-64)
    }
    __device__ __forceinline__ int32_t next() {
        const int sh    = 32 - gbits;
        const int32_t r = static_cast<int32_t>(static_cast<uint32_t>(d) << sh) >> sh;
        d               = (d - r) >> gbits;
        return r;
    }
};

}  // namespace mkfhe

// cluster_kernels.cuh -- one gate on a thread-block CLUSTER of CL = 2, 4 or 8 SMs (sm_100a), for the batches that leave most of the
// GPU idle: the single accumulator of the drop-in UniEncAccumulator::EvalAcc (mk-acc.h:67-71), one EvalBinGate, the narrow levels
// of a circuit.
//
// Why: a gate's steps are sequential, and within a step the chain  party terms -> sumV -> sumV term  (mk-acc-xzw.cpp:240-289) is
// what the clock sees.  On ONE SM (blind_rotate_grouped, two groups of 256 threads) a single boolean-mklwe gate keeps the fma-heavy
// pipe about 55% busy -- both groups' transforms queue on the same four sub-partitions -- and takes 8.7 ms.  Here CTA r of the
// cluster owns the parties p = r (mod CL): their accumulators live in ITS shared memory, its 8 warps have an SM's pipes to
// themselves, and the only traffic between the SMs is the partial sum  sum_p <D_p, Pkey_p>  of each CTA that does not own the
// step's key: 8 KiB per CTA and step, written straight into the key owner's shared memory (distributed shared memory,
// st.shared::cluster on a mapa address), followed by one cluster barrier (barrier.cluster.arrive / wait.acquire).  A CTA that does
// not own the key never waits for the owner's sumV term: after the barrier it transforms its last product back, updates its
// accumulator and starts the next step's party terms, so the key owner's chain (its parties, then the sumV term) is the only
// critical path.  Measured: 6.0 ms (k = 2, 2 SMs), 16.0 ms (k = 4, 4 SMs; one SM: 36.0), 92.8 ms (k = 16, 8 SMs; one SM: 428).
//
// Per executed step e (s = (u, i), rotation != 0), CTA r, owner CTA o = u mod CL:
//   for own parties p != u, then p = u (if r == o):  digits of (X^c - 1) a_p -> 2 forward NTTs -> X_p = <D_p, d>, V_r += <D_p, Pkey_p>
//   r != o:  V_r -> slot (r - o - 1 mod CL) of Vs[e & 1] in CTA o (remote);  arrive.release;  a_p += INTT(X_p) for the last own party;  wait
//   r == o:  arrive.relaxed;  wait;  V = V_o + sum of the slots of Vs[e & 1];  W = digits of INTT(V);  X_u += <NTT(W), f>;  a_u += INTT(X_u)
// Vs is double-buffered by the parity of the count of EXECUTED steps (steps with rotation 0 are skipped, identically in every CTA):
// CTA r's write of step e + 2 cannot pass the owner's read of step e, because r must first pass the barrier of step e + 1, at which
// the owner of step e arrives only after step e.
// Same exact sums as the sequential kernel (fast_step), evaluated in a different order: results are bit-identical (the tests hold
// this kernel to the oracle and to the compiled reference, for every cluster size and for party counts that do not divide it).
#pragma once
#include "fast_kernels.cuh"

namespace mkfhe {

// CL (template parameter below): CTAs per cluster = SMs per gate: 2, 4 or 8 (the portable maximum), at most the party count

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
// arrival of a CTA that has written nothing the others will read (the key owner): no release fence (ncu showed the fence of
// arrive.release as 5% of the owner's samples, profiles/r2_s21_cluster74.txt).  What the owner's arrival of step e + 1 orders is a
// write-after-read: the other CTAs may overwrite Vs[e & 1] only after the owner has read it in step e.  Those reads are shared-memory
// loads whose values were consumed by the sumV term of step e, thousands of instructions and at least three block barriers (bar.sync
// orders the CTA's prior shared-memory accesses) before this arrival, so no fence is needed for them.
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// address of `p` (a shared-memory address of this CTA) in the shared memory of CTA `rank` of the cluster
__device__ __forceinline__ uint32_t cluster_map(const void* p, uint32_t rank) {
    uint32_t local = static_cast<uint32_t>(__cvta_generic_to_shared(p)), remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(rank));
    return remote;
}
__device__ __forceinline__ void cluster_st4(uint32_t addr, uint4 v) {
    asm volatile("st.shared::cluster.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void cluster_st1(uint32_t addr, uint32_t v) {
    asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// parties of CTA r: p = r, r + CL, ...; party p lives at local row p / CL
__host__ __device__ inline uint32_t cluster_rows(uint32_t k, uint32_t CL) { return (k + CL - 1) / CL; }
// Shared memory of one CTA: a[rows][N] | E1[2][FNP] | E2[FNP] | Vs[2][CL - 1][N] | Ms[N] (XZW) | xchg[8] | crot[k*n] (u16)
__host__ __device__ inline size_t cluster_smem_bytes(uint32_t k, uint32_t n, bool xzw, uint32_t CL) {
    return (size_t(cluster_rows(k, CL)) * FN + 3 * FNP + 2 * (CL - 1) * FN + (xzw ? FN : 0) + 8) * 4 + ((size_t(k) * n * 2 + 15) & ~size_t(15));
}

// ---- key words in registers ahead of their use -----------------------------------------------------------------------------------
// With two warps per sub-partition nothing hides a global load between a transform and the multiply-accumulate that follows it
// (phase timing, profiles/r2_s23_cluster_timing.txt: 1.5 k cycles per <D, key> against 2.4 k for the two forward transforms before
// it; boolean-mkntru: 2.5 k more at the top of every step for the 8 scattered words of Q - X^-c).  This kernel has the register
// file of an SM for 256 threads, so the loads are issued BEFORE the forward transforms and land while they run.
template <bool XZW>
struct KeyRegs {
    uint4 a[2][2];                 // ev1, [digit][chunk of 4 slots]
    uint4 b[XZW ? 2 : 1][2];       // ev2 (XZW)
};
template <bool XZW>
__device__ __forceinline__ void key_load(KeyRegs<XZW>& K, const uint32_t* k1p, const uint32_t* k2p) {
    constexpr uint32_t N = FN, KCH = FN / 2;
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            K.a[j][ch] = ldg4_stream(k1p + 2 * j * N + KCH * ch);
            if (XZW) K.b[j][ch] = ldg4_stream(k2p + 2 * j * N + KCH * ch);
        }
}
// key_mac4's two steady-state branches (XZW_B: ev1; XZW: ev1 - ev2 X^-c as two 64-bit sums) on preloaded words; mq = Q - X^-c at this
// thread's 8 slots (XZW)
template <bool XZW>
__device__ __forceinline__ void key_mac_regs(const DevCtx& P, const KeyRegs<XZW>& K, const uint32_t (&mq)[8], int ch, const uint32_t (&D)[2][8],
                                             uint32_t (&x4)[4]) {
    const uint32_t Q = P.mod.Q;
    const uint32_t a0v[4] = {K.a[0][ch].x, K.a[0][ch].y, K.a[0][ch].z, K.a[0][ch].w};
    const uint32_t a1v[4] = {K.a[1][ch].x, K.a[1][ch].y, K.a[1][ch].z, K.a[1][ch].w};
    if (XZW) {
        const uint32_t b0v[4] = {K.b[0][ch].x, K.b[0][ch].y, K.b[0][ch].z, K.b[0][ch].w};
        const uint32_t b1v[4] = {K.b[XZW ? 1 : 0][ch].x, K.b[XZW ? 1 : 0][ch].y, K.b[XZW ? 1 : 0][ch].z, K.b[XZW ? 1 : 0][ch].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t d0 = D[0][4 * ch + q], d1 = D[1][4 * ch + q];  // < 24Q
            const uint64_t s2 = static_cast<uint64_t>(d0) * b0v[q] + static_cast<uint64_t>(d1) * b1v[q];
            const uint32_t q2 = static_cast<uint32_t>(s2) * P.mod.qinv_neg;
            const uint32_t r2 = static_cast<uint32_t>((s2 + static_cast<uint64_t>(q2) * Q) >> 32);  // < 3Q
            const uint64_t s1 = static_cast<uint64_t>(d0) * a0v[q] + static_cast<uint64_t>(d1) * a1v[q] +
                                static_cast<uint64_t>(r2) * mq[4 * ch + q];  // < 51 Q^2 < 2^60
            const uint32_t q1 = static_cast<uint32_t>(s1) * P.mod.qinv_neg;
            x4[q] = static_cast<uint32_t>((s1 + static_cast<uint64_t>(q1) * Q) >> 32);  // < 3Q
        }
    } else {
#pragma unroll
        for (int q = 0; q < 4; ++q) x4[q] = mac2(D[0][4 * ch + q], a0v[q], D[1][4 * ch + q], a1v[q], P.mod);
    }
}

// Phase timing of the key owner's chain (development aid: nvcc -DMKFHE_CLUSTER_TIMING; thread 0 of every CTA prints its totals)
#ifdef MKFHE_CLUSTER_TIMING
#define MKFHE_TICK(i_) { const uint32_t now_ = uint32_t(clock64()); tk[i_] += now_ - tlast; tlast = now_; }
#else
#define MKFHE_TICK(i_)
#endif

// One step of EvalAcc on the cluster (see the header comment).  `a`: this CTA's rows; `par`: parity of the count of executed steps
// (the same in both CTAs: they skip the same zero rotations).  Ends with a block barrier; every thread executes exactly one
// cluster arrive and one cluster wait.  Exchange-buffer discipline as in fast_step: forward transforms through E1, inverse ones
// through E2, and between two uses of either buffer lies the block barrier inside a transform on the other.
template <bool XZW, bool FIRST, int CL>
__device__ __forceinline__ void cluster_step(const DevCtx& P, const FastTables& T, uint32_t* a, uint32_t* E1, uint32_t* E2, uint32_t* Vs,
                                             uint32_t* Ms, uint32_t s, uint32_t c, uint32_t par, uint32_t tid, uint32_t rank
#ifdef MKFHE_CLUSTER_TIMING
                                             , uint32_t (&tk)[16]
#endif
                                             ) {
#ifdef MKFHE_CLUSTER_TIMING
    uint32_t tlast = uint32_t(clock64());
#endif
    constexpr uint32_t N = FN, M = 2 * FN;
    const uint32_t Q = P.mod.Q, k = P.k, n = P.n, g = P.gbits;
    const uint32_t u = FIRST ? 0u : s / n, i = FIRST ? 0u : s - u * n;
    const bool owner = (u % CL) == rank;
    const uint32_t ju = u / CL;                          // local row of the key owner's accumulator (in the owner CTA)
    const uint32_t rows = (k - rank + CL - 1) / CL;      // own parties
    const uint32_t cneg = (M - c) & (M - 1);
    const size_t per = size_t(2) * 2 * N;
    const uint32_t* ev1 = P.bk + (size_t(u * P.S) * n + i) * per + 4 * tid;
    const uint32_t* ev2 = XZW ? ev1 + size_t(n) * per : ev1;
    const uint32_t* evs = P.bk + size_t(k) * P.S * n * per + 4 * tid;
    const uint32_t brv8 = __brev(tid) >> 24;
    const uint32_t e0n = cneg * (2 * brv8 + 1), e0p = c * (2 * brv8 + 1);
    if (T.prefetch && rank == 0) prefetch_step_keys<XZW>(P, s + 1, tid);
    (void)Ms;
    uint32_t mq[8] = {};  // XZW: Q - X^-c at this thread's 8 slots, the same for every multiply-accumulate of the step: loaded once,
                          // into registers (the one-SM kernels park it in shared memory), first used after the forward transforms
    if (XZW && !FIRST) {
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            uint32_t m4[4];
            monomial4(P.psipow, e0n, cneg, ch, m4);
#pragma unroll
            for (int q = 0; q < 4; ++q) mq[4 * ch + q] = Q - m4[q];
        }
    }
    uint32_t* Vp = Vs + par * (CL - 1) * N + 8 * tid;  // this thread's 8 slots of the step's partial-sum buffers (one per other CTA)
    uint32_t keep[8];                       // V_r = sum over own parties of <D_p, Pkey_p>, < 4Q
    uint32_t y[8];                          // the last own party's <D_p, d>, evaluation domain, < 3Q
    uint32_t* alast = a;
#pragma unroll 1
    for (uint32_t j = 0; j < rows; ++j) {
        // own parties in increasing order, the key owner's (if it is ours) last: its f-term needs the complete V
        const uint32_t row = owner ? (j + 1 < rows ? (j < ju ? j : j + 1) : ju) : j;
        const uint32_t p = rank + CL * row;
        uint32_t* ap = a + row * N;
        uint32_t D[2][8];
#pragma unroll
        for (int r = 0; r < 8; ++r) decompose2(FIRST ? ap[tid + 256 * r] : rot_sub(ap, c, tid + 256 * r, N, Q), Q, g, D[0][r], D[1][r]);
        MKFHE_TICK(0)
        const uint32_t* pk = P.pkey + size_t(p) * 2 * N + 4 * tid;
        KeyRegs<XZW> K;
        uint4 pq[2][2];  // Pkey_p, [digit][chunk]
        if (!FIRST) key_load<XZW>(K, ev1, ev2);  // "d" column: in flight during the transforms
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            pq[0][ch] = ldg4(pk + (N / 2) * ch);
            pq[1][ch] = ldg4(pk + N + (N / 2) * ch);
        }
        fwd_ntt_regs<2>(D, E1, T, Q, tid);
        MKFHE_TICK(1)
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            uint32_t x4[4];
            if (FIRST) key_mac4<XZW, FIRST>(P, ev1, ev2, evs, c, cneg, e0p, e0n, ch, D, x4, nullptr);  // "d" column (KDM key, once per gate)
            else key_mac_regs<XZW>(P, K, mq, ch, D, x4);
            const uint4 q0 = pq[0][ch], q1 = pq[1][ch];
            const uint32_t pp0[4] = {q0.x, q0.y, q0.z, q0.w}, pp1[4] = {q1.x, q1.y, q1.z, q1.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int sl = 4 * ch + q;
                const uint32_t v = mac2(D[0][sl], pp0[q], D[1][sl], pp1[q], P.mod);  // < 3Q
                uint32_t vs = (j == 0 ? 0u : keep[sl]) + v;                          // < 7Q
                keep[sl] = min(vs, vs - 4 * Q);                                      // < 4Q
                y[sl] = x4[q];
            }
        }
        alast = ap;
        MKFHE_TICK(2)
        if (j + 1 < rows) {  // not the last own party: back to coefficients now
            inv_ntt_regs(y, E2, T, Q, tid);
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
    }
    if (!owner) {
        // V_r into the key owner's shared memory (the thread with the same tid there picks these 8 slots up), then the barrier
        const uint32_t slot = (rank + CL - 1 - u % CL) % CL;  // 0 .. CL - 2: this CTA's buffer in the owner's shared memory
        const uint32_t remote = cluster_map(Vp + slot * N, u % CL);
        cluster_st4(remote, make_uint4(keep[0], keep[1], keep[2], keep[3]));
        cluster_st4(remote + 16, make_uint4(keep[4], keep[5], keep[6], keep[7]));
        cluster_arrive();
        MKFHE_TICK(3)
    } else {
        cluster_arrive_relaxed();
        cluster_wait();
        MKFHE_TICK(4)
        // ---- the sumV term (mk-acc-xzw.cpp:273-289): V = V_o + V_r, W = digits of INTT(V), X_u += <NTT(W), f> ----
        uint32_t xu[8], D[2][8];
#pragma unroll
        for (int sl = 0; sl < 8; ++sl) {
            xu[sl] = y[sl];
            y[sl] = keep[sl];
        }
#pragma unroll
        for (int o = 0; o < CL - 1; ++o) {
            const uint4 r0 = *reinterpret_cast<const uint4*>(Vp + o * N), r1 = *reinterpret_cast<const uint4*>(Vp + o * N + 4);
            const uint32_t rv[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
            for (int sl = 0; sl < 8; ++sl) {
                const uint32_t v = y[sl] + rv[sl];  // < 8Q
                y[sl] = min(v, v - 4 * Q);          // < 4Q
            }
        }
        inv_ntt_regs(y, E2, T, Q, tid);
        MKFHE_TICK(5)
#pragma unroll
        for (int r = 0; r < 8; ++r) decompose2(fold1(y[r], Q), Q, g, D[0][r], D[1][r]);
        MKFHE_TICK(6)
        KeyRegs<XZW> K;
        if (!FIRST) key_load<XZW>(K, ev1 + N, ev2 + N);  // "f" column
        fwd_ntt_regs<2>(D, E1, T, Q, tid);
        MKFHE_TICK(7)
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            uint32_t x4[4];
            if (FIRST) key_mac4<XZW, FIRST>(P, ev1 + N, ev2 + N, evs + N, c, cneg, e0p, e0n, ch, D, x4, nullptr);  // "f" column
            else key_mac_regs<XZW>(P, K, mq, ch, D, x4);
#pragma unroll
            for (int q = 0; q < 4; ++q) y[4 * ch + q] = xu[4 * ch + q] + x4[q];  // < 6Q
        }
    }
    // the last own party's product (the owner's now includes the sumV term) back to coefficients
    MKFHE_TICK(8)
    inv_ntt_regs(y, E2, T, Q, tid);
    MKFHE_TICK(9)
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        if (FIRST) {
            alast[tid + 256 * r] = fold1(y[r], Q);
        } else {
            uint32_t v = alast[tid + 256 * r] + y[r];  // < 3Q
            v = min(v, v - 2 * Q);
            alast[tid + 256 * r] = min(v, v - Q);
        }
    }
    MKFHE_TICK(10)
    if (!owner) cluster_wait();  // (free by now: the owner arrived before its sumV term)
    MKFHE_TICK(11)
    __syncthreads();             // a[] complete before the next step's shifted reads
    MKFHE_TICK(12)
}

// Key switch of this CTA's parties (keyswitch_epilogue of generic_kernels.cuh restricted to the rows p = rank (mod CL)); the two
// CTAs switch their parties side by side.  MKLWE: b needs the sum of the B column over ALL parties (mklwe-pke.cpp:277-280): CTA r > 0
// leaves its share in CTA 0's xchg[r] before the closing cluster barrier of the kernel, CTA 0 finishes b after it.
__device__ __noinline__ uint32_t cluster_keyswitch(const DevCtx& P, const GateIO& io, uint32_t g, const uint32_t* a, uint32_t* scratch,
                                                   uint32_t rank, uint32_t CL) {
    const uint32_t N = P.N, Q = P.mod.Q, k = P.k, n = P.n;
    const uint32_t tid = threadIdx.x, T = blockDim.x;
    uint32_t* ms = scratch;
    uint32_t* sums = scratch + N;
    uint32_t bsum = 0;  // thread 0: this CTA's share of the B column
    for (uint32_t u = rank, row = 0; u < k; u += CL, ++row) {
        const uint32_t* au = a + row * N;
        for (uint32_t e = tid; e < N; e += T) {  // Transpose + INTT = [a0, -a_{N-1}, ..., -a_1]; RoundqQ (mntru-pke.cpp:359-374)
            const uint32_t v = e == 0 ? au[0] : au[N - e];
            const uint32_t ext = (e == 0 || v == 0) ? v : Q - v;
            ms[e] = round_qQ(ext, P.qKS, Q);
        }
        for (uint32_t x = tid; x < P.npad; x += T) sums[x] = 0;
        __syncthreads();
        switch (P.dKS) {
            case 3: keyswitch_party<3>(P, u, ms, sums); break;
            case 4: keyswitch_party<4>(P, u, ms, sums); break;
            default: keyswitch_party<0>(P, u, ms, sums); break;
        }
        __syncthreads();
        uint32_t* o = io.out + (size_t(g) * k + u) * n;
        for (uint32_t x = tid; x < n; x += T) {
            const uint32_t sv = sums[x] % P.qKS;
            o[x] = P.scheme == 1 ? (sv == 0 ? 0u : P.qKS - sv) : sv;
        }
        if (P.scheme == 1 && tid == 0) bsum = (bsum + sums[n] % P.qKS) % P.qKS;
        __syncthreads();
    }
    return bsum;
}

template <bool GATE, bool XZW, int CL>  // launched with the cluster-dimension attribute (CL, 1, 1)
__global__ void __launch_bounds__(FT, 1) blind_rotate_cluster(const DevCtx P, const FastTables T, const GateIO io) {
    extern __shared__ __align__(16) uint32_t sm[];
    const uint32_t k = P.k, n = P.n, tid = threadIdx.x, rank = cluster_rank(), g = blockIdx.x / CL;
    constexpr uint32_t N = FN, M = 2 * FN;
    const uint32_t rows = (k - rank + CL - 1) / CL;
    uint32_t* a = sm;
    uint32_t* E1 = a + cluster_rows(k, CL) * N;
    uint32_t* E2 = E1 + 2 * FNP;
    uint32_t* Vs = E2 + FNP;
    uint32_t* Ms = Vs + 2 * (CL - 1) * N;
    uint32_t* xchg = XZW ? Ms + N : Ms;
    uint16_t* crot = reinterpret_cast<uint16_t*>(xchg + 8);
    const uint32_t Q = P.mod.Q;

    // ---- prologue: rotation amounts (both CTAs keep all of them) and this CTA's rows of the initial accumulator ----
    for (uint32_t s = tid; s < k * n; s += FT) crot[s] = static_cast<uint16_t>(rotation_amount<GATE>(P, io, g, s));
    if (GATE) {
        // initial accumulator in COEFFICIENT form: party 0 only (binfhe-base-scheme.cpp:1018-1055 MKLWE, :1086-1110 MNTRU), as
        // rotate_prologue builds it
        uint32_t b_hat = 0;
        if (P.scheme == 1) {
            const uint32_t bt = submod((5u * P.q / 8u) % P.q, addmod(io.b1[g], io.b2[g], P.q), P.q);
            b_hat = round_qQ(bt, M, P.q);
        }
        const uint32_t Q2p = Q / 8 + 1, Q2pNeg = Q - Q2p;
        for (uint32_t e = tid; e < rows * N; e += FT) {
            uint32_t v = 0;
            if (rank == 0 && e < N) {
                if (P.scheme == 0) {
                    v = e < N / 2 ? Q2pNeg : Q2p;
                } else {
                    const uint32_t j = (e + M - b_hat) & (N - 1);
                    const uint32_t Rx = j < N / 2 ? Q2p : Q2pNeg;
                    const uint32_t index = b_hat + j;
                    v = (index >= N && index < M) ? Q - Rx : Rx;
                }
            }
            a[e] = v;
        }
        __syncthreads();
    } else {
        for (uint32_t row = 0; row < rows; ++row) {
            const uint32_t* ai = io.acc_in + (size_t(g) * k + rank + CL * row) * N;
            for (uint32_t e = tid; e < N; e += FT) a[row * N + e] = ai[e];
        }
        __syncthreads();
        ntt_inv_smem(a, rows, P);
        for (uint32_t e = tid; e < rows * N; e += FT) a[e] = montmul(a[e], P.ninv_mont, P.mod);
        __syncthreads();
    }
    // both CTAs are running (their shared memory may be written from now on)
    cluster_arrive();
    cluster_wait();

    uint32_t par = 0;
#ifdef MKFHE_CLUSTER_TIMING
    uint32_t tk[16] = {};
    const uint32_t tstart = uint32_t(clock64());
#define MKFHE_TK , tk
#else
#define MKFHE_TK
#endif
    cluster_step<XZW, true, CL>(P, T, a, E1, E2, Vs, Ms, 0, crot[0], par, tid, rank MKFHE_TK);
    for (uint32_t s = 1; s < k * n; ++s) {
        const uint32_t c = crot[s];
        if (c == 0) continue;  // X^0 - 1 = 0: the step leaves acc unchanged (the same decision in both CTAs)
        par ^= 1u;
        cluster_step<XZW, false, CL>(P, T, a, E1, E2, Vs, Ms, s, c, par, tid, rank MKFHE_TK);
    }
#undef MKFHE_TK
#ifdef MKFHE_CLUSTER_TIMING
    if (tid == 0 && g == 0)
        printf("cluster timing rank %u: total %u | decomp %u fwd2 %u mac %u | nonowner: stV+arrive %u | owner: arrive+wait %u invV %u decomp %u fwd2 %u "
               "macf %u | inv %u update %u wait %u sync %u\n", rank, uint32_t(clock64()) - tstart, tk[0], tk[1], tk[2], tk[3], tk[4], tk[5], tk[6],
               tk[7], tk[8], tk[9], tk[10], tk[11], tk[12]);
#endif

    // ---- epilogue: this CTA's rows -> outputs ----
    uint32_t bsum = 0;
    if (GATE && P.ksk16 != nullptr) {
        bsum = cluster_keyswitch(P, io, g, a, E1, rank, CL);
        if (P.scheme == 1 && rank != 0 && tid == 0) cluster_st1(cluster_map(xchg + rank, 0), bsum);
    } else if (GATE) {
        for (uint32_t row = 0; row < rows; ++row) {
            uint32_t* out = io.ms_out + (size_t(g) * k + rank + CL * row) * N;
            const uint32_t* au = a + row * N;
            for (uint32_t e = tid; e < N; e += FT) {
                const uint32_t v = e == 0 ? au[0] : au[N - e];
                const uint32_t ext = (e == 0 || v == 0) ? v : Q - v;
                out[e] = round_qQ(ext, P.qKS, Q);
            }
        }
    } else {
        ntt_fwd_smem(a, rows, P);
        for (uint32_t row = 0; row < rows; ++row) {
            uint32_t* out = io.acc_out + (size_t(g) * k + rank + CL * row) * N;
            for (uint32_t e = tid; e < N; e += FT) out[e] = a[row * N + e];
        }
    }
    // closing barrier: CTA 1's share of b has landed, and no CTA exits while its shared memory may still be written
    cluster_arrive();
    cluster_wait();
    if (GATE && P.ksk16 != nullptr && P.scheme == 1 && rank == 0 && tid == 0) {
        uint32_t all = bsum;  // b = RoundqQ(Q/8 + 1) - sum B (binfhe-base-scheme.cpp:450-458)
        for (uint32_t r = 1; r < CL; ++r) all = (all + xchg[r]) % P.qKS;
        io.bout[g] = submod(round_qQ((Q >> 3) + 1, P.qKS, Q), all, P.qKS);
    }
}

// ---- host side ----
inline bool cluster_supported(const DevCtx& d, uint32_t CL) {
    return fast_supported(d) && d.k >= CL && cluster_smem_bytes(d.k, d.n, d.variant == 0, CL) <= 227 * 1024;
}
template <class K>
inline cudaError_t cluster_launch(K kern, uint32_t CL, uint32_t batch, size_t smem, cudaStream_t st, const DevCtx& d, const FastTables& T,
                                  const GateIO& io, int* max_clusters) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(batch * CL, 1, 1);
    cfg.blockDim = dim3(FT, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (max_clusters != nullptr) return cudaOccupancyMaxActiveClusters(max_clusters, kern, &cfg);
    return cudaLaunchKernelEx(&cfg, kern, d, T, io);
}
// kernel of (gate, xzw, CL); max_clusters != null: only ask how many clusters of this shape the GPU holds at once
inline cudaError_t cluster_dispatch(FastState& F, const DevCtx& d, bool gate, uint32_t CL, const GateIO& io, uint32_t batch, cudaStream_t st,
                                    int* max_clusters) {
    const bool xzw = d.variant == 0;
    const size_t smem = cluster_smem_bytes(d.k, d.n, xzw, CL);
#define MKFHE_CL(G_, X_)                                                                                                  \
    (CL == 2 ? cluster_launch(blind_rotate_cluster<G_, X_, 2>, 2, batch, smem, st, d, F.T, io, max_clusters)               \
             : CL == 4 ? cluster_launch(blind_rotate_cluster<G_, X_, 4>, 4, batch, smem, st, d, F.T, io, max_clusters)     \
                       : cluster_launch(blind_rotate_cluster<G_, X_, 8>, 8, batch, smem, st, d, F.T, io, max_clusters))
    return gate ? (xzw ? MKFHE_CL(true, true) : MKFHE_CL(true, false)) : (xzw ? MKFHE_CL(false, true) : MKFHE_CL(false, false));
#undef MKFHE_CL
}
// SMs per gate for a batch: the largest cluster (8, 4, 2 CTAs, at most one per party) of which the GPU holds `batch` at once -- a
// second wave would double the latency the cluster is there to cut; 0 = no cluster (the one-SM kernels).
// MKFHE_CLUSTER = 0: never; 2 / 4 / 8: that size whenever the parameter set supports it (A/B runs, tests); 1: this policy.
inline uint32_t cluster_size_for(FastState& F, const DevCtx& d, bool gate, uint32_t batch) {
    uint32_t forced = 0;
    if (const char* e = getenv("MKFHE_CLUSTER")) {
        forced = uint32_t(atoi(e));
        if (forced == 0) return 0;
        if (forced == 2 || forced == 4 || forced == 8) return cluster_supported(d, forced) ? forced : 0;
        forced = 1;  // any other value: the default policy below, also next to hand-picked one-SM variants
    }
    // a run that selects a one-SM kernel variant by hand (experiments, the variant tests) keeps it
    if (!forced && (getenv("MKFHE_GROUPS") || getenv("MKFHE_MINB") || getenv("MKFHE_FAST_ALIAS") || getenv("MKFHE_AG"))) return 0;
    for (uint32_t CL = 8; CL >= 2; CL >>= 1) {
        if (!cluster_supported(d, CL) || uint64_t(batch) * CL > uint64_t(F.sms)) continue;
        int& cached = F.max_clusters[gate ? 1 : 0][CL == 8 ? 2 : CL == 4 ? 1 : 0];
        if (cached < 0) {
            int m = 0;
            GateIO none{};
            if (cluster_dispatch(F, d, gate, CL, none, 1, nullptr, &m) != cudaSuccess) {
                cudaGetLastError();
                m = 0;
            }
            cached = m;
        }
        if (int64_t(batch) <= int64_t(cached)) return CL;
    }
    return 0;
}
inline cudaError_t cluster_blind_rotate(FastState& F, const DevCtx& d, bool gate, uint32_t CL, const GateIO& io, uint32_t batch,
                                        cudaStream_t st, uint64_t* launches) {
    F.T.prefetch = d.k <= 8 ? 1u : 0u;
    if (const char* pf = getenv("MKFHE_PREFETCH")) F.T.prefetch = pf[0] == '1' ? 1u : 0u;
    const cudaError_t e = cluster_dispatch(F, d, gate, CL, io, batch, st, nullptr);
    if (e != cudaSuccess) return e;
    ++*launches;
    return cudaGetLastError();
}

}  // namespace mkfhe

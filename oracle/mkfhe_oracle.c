/*
 * mkfhe_oracle.c -- plain-C restatement of the reference's multi-key bootstrapping path (see mkfhe_oracle.h).
 * TEST INFRASTRUCTURE ONLY.  Written for clarity: every modular operation is a 64-bit `%`; canonical residues throughout.
 * file:line citations are into /root/reference.
 */
#include "mkfhe_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

struct orc_ctx {
    orc_params p;
    uint32_t ell;      /* digitsG - 1 digits are used (mk-acc.cpp:63-64) */
    uint32_t logN;
    uint32_t* tw;      /* Table[bitrev(i)] = psi^i            (transformnat-impl.h:714-727) */
    uint32_t* twi;     /* TableI[bitrev(i)] = psi^-i                                         */
    uint32_t Ninv;     /* N^-1 mod Q                                                         */
};

/* ---------- modular helpers ---------- */
static inline uint32_t mulmod(uint32_t a, uint32_t b, uint32_t m) { return (uint32_t)(((uint64_t)a * b) % m); }
static inline uint32_t addmod(uint32_t a, uint32_t b, uint32_t m) { uint64_t s = (uint64_t)a + b; return (uint32_t)(s >= m ? s - m : s); }
static inline uint32_t submod(uint32_t a, uint32_t b, uint32_t m) { return a >= b ? a - b : a + m - b; }
static uint32_t powmod(uint32_t a, uint64_t e, uint32_t m) {
    uint64_t r = 1, b = a % m;
    while (e) {
        if (e & 1) r = (r * b) % m;
        b = (b * b) % m;
        e >>= 1;
    }
    return (uint32_t)r;
}
static uint32_t invmod(uint32_t a, uint32_t p) { return powmod(a, (uint64_t)p - 2, p); } /* p prime */
static uint32_t bitrev(uint32_t x, uint32_t bits) {
    uint32_t r = 0;
    for (uint32_t i = 0; i < bits; ++i) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}

static int is_prime(uint64_t n) { /* deterministic Miller-Rabin for n < 3.2e9 (bases 2,3,5,7) */
    if (n < 2) return 0;
    static const uint32_t small[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (unsigned i = 0; i < sizeof small / sizeof *small; ++i) {
        if (n == small[i]) return 1;
        if (n % small[i] == 0) return 0;
    }
    uint64_t d = n - 1;
    int s = 0;
    while ((d & 1) == 0) { d >>= 1; ++s; }
    static const uint32_t bases[] = {2, 3, 5, 7};
    for (unsigned i = 0; i < 4; ++i) {
        uint64_t x = powmod(bases[i], d, (uint32_t)n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int r = 1; r < s; ++r) {
            x = (x * x) % n;
            if (x == n - 1) { comp = 0; break; }
        }
        if (comp) return 0;
    }
    return 1;
}

/* nbtheory-impl.h:334-358: first prime >= 2^nbits + 1 that is 1 mod m */
uint32_t orc_first_prime(uint32_t nbits, uint32_t m) {
    uint64_t q = 1ull << nbits, r = q % m, q2 = q + 1;
    if (r > 0) q2 += m - r;
    while (!is_prime(q2)) q2 += m;
    return (uint32_t)q2;
}
/* nbtheory-impl.h:371-379 */
uint32_t orc_previous_prime(uint32_t q, uint32_t m) {
    uint64_t c = (uint64_t)q - m;
    while (!is_prime(c)) c -= m;
    return (uint32_t)c;
}
/* nbtheory-impl.h:183-231: the MINIMUM primitive m-th root of unity (m a power of two) */
uint32_t orc_root_of_unity(uint32_t m, uint32_t Q) {
    uint32_t g = 0;
    for (uint32_t x = 2; x < Q; ++x) {
        uint32_t c = powmod(x, ((uint64_t)Q - 1) / m, Q);
        if (powmod(c, m / 2, Q) == Q - 1) { g = c; break; } /* order exactly m */
    }
    uint32_t best = g, x = g, g2 = mulmod(g, g, Q);
    for (uint32_t e = 1; e < m; e += 2) { /* all odd powers = all primitive roots */
        if (x < best) best = x;
        x = mulmod(x, g2, Q);
    }
    return best;
}

/* ---------- context ---------- */
orc_ctx* orc_create(const orc_params* p) {
    orc_ctx* c = (orc_ctx*)calloc(1, sizeof *c);
    c->p = *p;
    c->ell = p->digitsG - 1;
    c->logN = 0;
    while ((1u << c->logN) < p->N) ++c->logN;
    c->tw = (uint32_t*)malloc(sizeof(uint32_t) * p->N);
    c->twi = (uint32_t*)malloc(sizeof(uint32_t) * p->N);
    uint32_t psi_inv = invmod(p->psi, p->Q), x = 1, xi = 1;
    for (uint32_t i = 0; i < p->N; ++i) { /* transformnat-impl.h:721-727 */
        uint32_t r = bitrev(i, c->logN);
        c->tw[r] = x;
        c->twi[r] = xi;
        x = mulmod(x, p->psi, p->Q);
        xi = mulmod(xi, psi_inv, p->Q);
    }
    c->Ninv = invmod(p->N % p->Q, p->Q);
    return c;
}
void orc_destroy(orc_ctx* c) {
    if (!c) return;
    free(c->tw);
    free(c->twi);
    free(c);
}

/* ---------- NTT ---------- */
void orc_ntt(const orc_ctx* c, uint32_t* a, int inverse) {
    const uint32_t N = c->p.N, Q = c->p.Q;
    if (!inverse) {
        /* ForwardTransformToBitReverseInPlace, transformnat-impl.h:299-354 (the trailing t=1 loop is the same butterfly) */
        for (uint32_t m = 1, t = N >> 1; m < N; m <<= 1, t >>= 1) {
            for (uint32_t i = 0; i < m; ++i) {
                uint32_t w = c->tw[i + m];
                for (uint32_t j = 2 * i * t; j < 2 * i * t + t; ++j) {
                    uint32_t of = mulmod(a[j + t], w, Q), lo = a[j];
                    a[j] = addmod(lo, of, Q);
                    a[j + t] = submod(lo, of, Q);
                }
            }
        }
    } else {
        /* InverseTransformFromBitReverseInPlace, transformnat-impl.h:491-552: first stage carries N^-1 */
        for (uint32_t i = 0; i < N; i += 2) {
            uint32_t w = c->twi[(i + N) >> 1], hi = a[i + 1], lo = a[i];
            uint32_t of = submod(lo, hi, Q);
            lo = addmod(lo, hi, Q);
            a[i] = mulmod(lo, c->Ninv, Q);
            a[i + 1] = mulmod(mulmod(of, w, Q), c->Ninv, Q);
        }
        for (uint32_t m = N >> 2, t = 2, logt = 2; m >= 1; m >>= 1, t <<= 1, ++logt) {
            for (uint32_t i = 0; i < m; ++i) {
                uint32_t w = c->twi[i + m];
                for (uint32_t j = i << logt; j < (i << logt) + t; ++j) {
                    uint32_t hi = a[j + t], lo = a[j];
                    a[j] = addmod(lo, hi, Q);
                    a[j + t] = mulmod(submod(lo, hi, Q), w, Q);
                }
            }
        }
    }
}

/* ---------- SignedDigitDecompose, mk-acc.cpp:54-80 (NATIVEINT = 32: SignedNativeInt = int32_t, MaxBits() = 32) ---------- */
void orc_decompose(const orc_ctx* c, const uint32_t* in, uint32_t* out) {
    const uint32_t N = c->p.N;
    const uint32_t QHalf = c->p.Q >> 1;
    const int32_t Q_int = (int32_t)c->p.Q;
    const int32_t gBits = (int32_t)c->p.baseG_log2;
    const int32_t gBitsMaxBits = 32 - gBits;
    for (uint32_t k = 0; k < N; ++k) {
        uint32_t t0 = in[k];
        int32_t d0 = (int32_t)(t0 < QHalf ? t0 : t0 - (uint32_t)Q_int);
        /* (d0 << s) >> s : sign-extend the low gBits bits (arithmetic right shift, as gcc implements it for int32) */
        int32_t r0 = (int32_t)((uint32_t)d0 << gBitsMaxBits) >> gBitsMaxBits;
        d0 = (d0 - r0) >> gBits; /* first digit is dropped */
        for (uint32_t d = 0; d < c->ell; ++d) {
            r0 = (int32_t)((uint32_t)d0 << gBitsMaxBits) >> gBitsMaxBits;
            d0 = (d0 - r0) >> gBits;
            if (r0 < 0) r0 += Q_int;
            out[(size_t)d * N + k] = (uint32_t)r0; /* reference does += into a zero vector */
        }
    }
}

/* ---------- monomials, mk-cryptoparameters.cpp:53-70 ---------- */
void orc_monomial(const orc_ctx* c, uint32_t m, uint32_t* out) {
    const uint32_t N = c->p.N, Q = c->p.Q;
    memset(out, 0, sizeof(uint32_t) * N);
    out[0] = submod(out[0], 1, Q);
    if (m < N)
        out[m] = addmod(out[m], 1, Q);
    else
        out[m - N] = submod(out[m - N], 1, Q);
    orc_ntt(c, out, 0);
}

/* ---------- HbProd, mk-acc-xzw.cpp:231-290 (identical in mk-acc-xzw_B.cpp:221-279) ---------- */
static void hbprod(const orc_ctx* c, const uint32_t* d /*[ell][N]*/, const uint32_t* f /*[ell][N]*/, uint32_t index,
                   const uint32_t* pkey /*[k][ell][N]*/, uint32_t* acc /*[k][N] EVALUATION, in/out*/, uint32_t* work) {
    const uint32_t N = c->p.N, Q = c->p.Q, k = c->p.k, ell = c->ell;
    uint32_t* ct = work;           /* [N]      */
    uint32_t* dct = work + N;      /* [ell][N] */
    uint32_t* sumV = dct + ell * N; /* [N]     */
    memset(sumV, 0, sizeof(uint32_t) * N);
    for (uint32_t u = 0; u < k; ++u) {
        memcpy(ct, acc + (size_t)u * N, sizeof(uint32_t) * N);
        orc_ntt(c, ct, 1);                                       /* ct[u].SetFormat(COEFFICIENT) :240-243 */
        orc_decompose(c, ct, dct);                               /* :255 */
        for (uint32_t i = 0; i < ell; ++i) orc_ntt(c, dct + (size_t)i * N, 0); /* :259-261 */
        for (uint32_t x = 0; x < N; ++x) {                       /* :263-269 */
            uint32_t uj = 0, v = 0;
            for (uint32_t i = 0; i < ell; ++i) {
                uj = addmod(uj, mulmod(dct[(size_t)i * N + x], d[(size_t)i * N + x], Q), Q);
                v = addmod(v, mulmod(dct[(size_t)i * N + x], pkey[((size_t)u * ell + i) * N + x], Q), Q);
            }
            sumV[x] = addmod(sumV[x], v, Q);
            acc[(size_t)u * N + x] = uj;
        }
    }
    orc_ntt(c, sumV, 1);                                         /* :273 */
    orc_decompose(c, sumV, dct);                                 /* :275 */
    for (uint32_t i = 0; i < ell; ++i) orc_ntt(c, dct + (size_t)i * N, 0);
    for (uint32_t x = 0; x < N; ++x) {                           /* :281-288 */
        uint32_t w = 0;
        for (uint32_t i = 0; i < ell; ++i) w = addmod(w, mulmod(dct[(size_t)i * N + x], f[(size_t)i * N + x], Q), Q);
        acc[(size_t)index * N + x] = addmod(acc[(size_t)index * N + x], w, Q);
    }
}

/* evk(u, s, i) -> pointer to [ell][2][N]; i == n with (u,s) == (0,0) is evk* */
static const uint32_t* evk_at(const orc_ctx* c, const uint32_t* bk, uint32_t u, uint32_t s, uint32_t i) {
    const uint32_t S = c->p.variant == 0 ? 2 : 1;
    const size_t per = (size_t)c->ell * 2 * c->p.N;
    if (i == c->p.n) return bk + (size_t)c->p.k * S * c->p.n * per;
    return bk + (((size_t)u * S + s) * c->p.n + i) * per;
}

/* EvalAcc: mk-acc-xzw.cpp:89-130 (+AddToAccXZW :292-345, AddToAccXZW0 :347-381),
 *          mk-acc-xzw_B.cpp:103-133 (+AddToAccXZW :281-331, AddToAccXZW0 :333-379) */
void orc_eval_acc(const orc_ctx* c, const uint32_t* bk, const uint32_t* pkey, uint32_t* acc, const uint32_t* rot,
                  uint32_t rot_mod) {
    const uint32_t N = c->p.N, Q = c->p.Q, k = c->p.k, n = c->p.n, ell = c->ell, M = 2 * N;
    uint32_t* mono = (uint32_t*)malloc(sizeof(uint32_t) * N * 2);
    uint32_t* monoNeg = mono + N;
    uint32_t* d = (uint32_t*)malloc(sizeof(uint32_t) * N * ell * 2);
    uint32_t* f = d + (size_t)ell * N;
    uint32_t* tmp = (uint32_t*)malloc(sizeof(uint32_t) * N * k);
    uint32_t* work = (uint32_t*)malloc(sizeof(uint32_t) * N * (ell + 2));
    for (uint32_t u = 0; u < k; ++u) {
        for (uint32_t i = 0; i < n; ++i) {
            uint32_t r = rot[(size_t)u * n + i];
            /* XZW: ct[u][i]*2*N/mod in NativeInteger<uint32> arithmetic (mk-acc-xzw.cpp:110,125); XZW_B: ct[u][i] itself */
            uint32_t cpos = c->p.variant == 0 ? (uint32_t)((r * 2u * N) / rot_mod) : r;
            uint32_t cneg = (M - cpos) % M; /* NativeInteger(0).ModSubFast(c, M) */
            orc_monomial(c, cpos == M ? 0 : cpos, mono);
            const uint32_t* ev1 = evk_at(c, bk, u, 0, i);
            if (c->p.variant == 0) {
                const uint32_t* ev2 = evk_at(c, bk, u, 1, i);
                orc_monomial(c, cneg == M ? 0 : cneg, monoNeg);
                if (u == 0 && i == 0) { /* AddToAccXZW0 :370-377 */
                    const uint32_t* evs = evk_at(c, bk, 0, 0, n);
                    for (uint32_t j = 0; j < ell; ++j)
                        for (uint32_t x = 0; x < N; ++x) {
                            size_t o0 = ((size_t)j * 2 + 0) * N + x, o1 = ((size_t)j * 2 + 1) * N + x;
                            d[(size_t)j * N + x] = addmod(addmod(evs[o0], mulmod(ev1[o0], mono[x], Q), Q), mulmod(ev2[o0], monoNeg[x], Q), Q);
                            f[(size_t)j * N + x] = addmod(addmod(evs[o1], mulmod(ev1[o1], mono[x], Q), Q), mulmod(ev2[o1], monoNeg[x], Q), Q);
                        }
                } else { /* AddToAccXZW :322-325 */
                    for (uint32_t j = 0; j < ell; ++j)
                        for (uint32_t x = 0; x < N; ++x) {
                            size_t o0 = ((size_t)j * 2 + 0) * N + x, o1 = ((size_t)j * 2 + 1) * N + x;
                            d[(size_t)j * N + x] = submod(submod(ev1[o0], mulmod(ev2[o0], monoNeg[x], Q), Q), ev2[o0], Q);
                            f[(size_t)j * N + x] = submod(submod(ev1[o1], mulmod(ev2[o1], monoNeg[x], Q), Q), ev2[o1], Q);
                        }
                }
            } else {
                if (u == 0 && i == 0) { /* mk-acc-xzw_B.cpp:366-369 */
                    const uint32_t* evs = evk_at(c, bk, 0, 0, n);
                    for (uint32_t j = 0; j < ell; ++j)
                        for (uint32_t x = 0; x < N; ++x) {
                            size_t o0 = ((size_t)j * 2 + 0) * N + x, o1 = ((size_t)j * 2 + 1) * N + x;
                            d[(size_t)j * N + x] = addmod(evs[o0], mulmod(ev1[o0], mono[x], Q), Q);
                            f[(size_t)j * N + x] = addmod(evs[o1], mulmod(ev1[o1], mono[x], Q), Q);
                        }
                } else { /* mk-acc-xzw_B.cpp:306-309 */
                    for (uint32_t j = 0; j < ell; ++j)
                        for (uint32_t x = 0; x < N; ++x) {
                            d[(size_t)j * N + x] = ev1[((size_t)j * 2 + 0) * N + x];
                            f[(size_t)j * N + x] = ev1[((size_t)j * 2 + 1) * N + x];
                        }
                }
            }
            if (u == 0 && i == 0) {
                hbprod(c, d, f, u, pkey, acc, work); /* acc <- HbProd(acc) */
            } else {
                for (uint32_t v = 0; v < k; ++v)
                    for (uint32_t x = 0; x < N; ++x) tmp[(size_t)v * N + x] = mulmod(acc[(size_t)v * N + x], mono[x], Q);
                hbprod(c, d, f, u, pkey, tmp, work);
                for (size_t x = 0; x < (size_t)k * N; ++x) acc[x] = addmod(acc[x], tmp[x], Q);
            }
        }
    }
    free(mono);
    free(d);
    free(tmp);
    free(work);
}

/* ---------- epilogue pieces ---------- */
uint32_t orc_round_qQ(uint32_t v, uint32_t q, uint32_t Q) { /* mntru-pke.cpp:12-17 */
    volatile double prod = (double)v * (double)q; /* volatile: forbid fused contraction, keep IEEE double steps */
    volatile double quo = prod / (double)Q;
    double r = floor(0.5 + quo);
    return (uint32_t)(((uint64_t)r) % q);
}

void orc_transpose_intt(const orc_ctx* c, uint32_t* a) {
    /* AutomorphismTransform(2N-1) in EVALUATION format, poly-impl.h:349-356 */
    const uint32_t N = c->p.N, logn = c->logN, mask = N - 1, kk = 2 * N - 1;
    uint32_t* r = (uint32_t*)malloc(sizeof(uint32_t) * N);
    uint32_t jk = kk;
    for (uint32_t j = 0; j < N; ++j, jk += 2 * kk) {
        uint32_t jrev = bitrev(j, logn), idxrev = bitrev((jk >> 1) & mask, logn);
        r[jrev] = a[idxrev];
    }
    memcpy(a, r, sizeof(uint32_t) * N);
    free(r);
    orc_ntt(c, a, 1);
}

/* MNTRUEncryptionScheme::KeySwitch2, mntru-pke.cpp:763-823, with KSK2[u][digit][l] = digit*KSK2[u][1][l] (:743-752) */
void orc_keyswitch_mntru(const orc_ctx* c, const uint32_t* ksk, const uint32_t* ctQN, uint32_t* out) {
    const uint32_t N = c->p.N, k = c->p.k, n = c->p.n, dKS = c->p.dKS, B = c->p.baseKS, qKS = c->p.qKS;
    for (uint32_t u = 0; u < k; ++u) {
        uint32_t* o = out + (size_t)u * n;
        memset(o, 0, sizeof(uint32_t) * n);
        for (uint32_t i = 0; i < N; ++i) {
            uint32_t ctmp = ctQN[(size_t)u * N + i];
            for (uint32_t j = 0; j < dKS; ++j) {
                uint32_t dig = ctmp % B;
                ctmp /= B;
                if (dig == 0) continue;
                const uint32_t* row = ksk + ((size_t)u * N * dKS + (size_t)i * dKS + j) * n;
                for (uint32_t x = 0; x < n; ++x) o[x] = addmod(o[x], mulmod(row[x], dig, qKS), qKS);
            }
        }
    }
}

/* MKLWEEncryptionScheme::KeySwitch, mklwe-pke.cpp:260-290 (zero digits are NOT skipped) */
void orc_keyswitch_mklwe(const orc_ctx* c, const uint32_t* A, const uint32_t* Bk, const uint32_t* ctQN, uint32_t b,
                         uint32_t* out, uint32_t* bout) {
    const uint32_t N = c->p.N, k = c->p.k, n = c->p.n, dKS = c->p.dKS, B = c->p.baseKS, qKS = c->p.qKS;
    memset(out, 0, sizeof(uint32_t) * k * n);
    for (uint32_t u = 0; u < k; ++u)
        for (uint32_t i = 0; i < N; ++i) {
            uint32_t atmp = ctQN[(size_t)u * N + i];
            for (uint32_t j = 0; j < dKS; ++j) {
                uint32_t a0 = atmp % B;
                atmp /= B;
                size_t row = (((size_t)u * N + i) * B + a0) * dKS + j;
                b = submod(b, Bk[row], qKS);
                const uint32_t* ar = A + row * n;
                for (uint32_t x = 0; x < n; ++x) out[(size_t)u * n + x] = submod(out[(size_t)u * n + x], ar[x], qKS);
            }
        }
    *bout = b;
}

/* ---------- EvalBinGate(NAND), binfhe-base-scheme.cpp:380-515 + BootstrapGateCore :1004-1130 ---------- */
void orc_eval_nand(const orc_ctx* c, const uint32_t* bk, const uint32_t* pkey, const uint32_t* kskA, const uint32_t* kskB,
                   const uint32_t* ctnand, const uint32_t* ct1, uint32_t b1, const uint32_t* ct2, uint32_t b2, uint32_t* out,
                   uint32_t* bout, uint32_t* trace_acc) {
    const uint32_t N = c->p.N, Q = c->p.Q, k = c->p.k, n = c->p.n, q = c->p.q, qKS = c->p.qKS;
    uint32_t* rot = (uint32_t*)malloc(sizeof(uint32_t) * k * n);
    uint32_t* acc = (uint32_t*)calloc((size_t)k * N, sizeof(uint32_t));
    const uint32_t Q2p = Q / 8 + 1, Q2pNeg = Q - Q2p; /* p = 4: Q/(2p)+1, :1018-1019, :1086-1087 */
    uint32_t rot_mod;
    if (c->p.scheme == 0) {
        /* ct_temp = ctNAND - (ct1 + ct2) mod q, :485-490 */
        for (size_t x = 0; x < (size_t)k * n; ++x) rot[x] = submod(ctnand[x], addmod(ct1[x], ct2[x], q), q);
        for (uint32_t j = 0; j < N; ++j) acc[j] = j < N / 2 ? Q2pNeg : Q2p; /* :1092-1094 */
        rot_mod = q;
    } else {
        /* ct_temp = (0, 5q/8) - (ct1 + ct2), :393-401; ModSwitch to 2N, :405-406; b_hat rotation of the test vector, :1022-1043 */
        uint32_t M = 2 * N;
        uint32_t bt = submod((5 * q / 8) % q, addmod(b1, b2, q), q);
        uint32_t bMS = orc_round_qQ(bt, M, q);
        for (size_t x = 0; x < (size_t)k * n; ++x) {
            uint32_t a = submod(0, addmod(ct1[x], ct2[x], q), q);
            uint32_t aMS = orc_round_qQ(a, M, q);
            rot[x] = submod(0, aMS, M); /* GetAneg(), mklwe-ciphertext.h:86-96 */
        }
        uint32_t b_hat = (uint32_t)(((uint64_t)bMS * 2 * N) / M); /* b*2*N/q with q = 2N here */
        for (uint32_t j = 0; j < N; ++j) {
            uint32_t Rx = j < N / 2 ? Q2p : Q2pNeg, index = b_hat + j;
            if (index >= N && index < 2 * N)
                acc[index % N] = Q - Rx;
            else
                acc[index % N] = Rx;
        }
        rot_mod = M;
    }
    orc_ntt(c, acc, 0); /* res[0].SetFormat(EVALUATION) */
    orc_eval_acc(c, bk, pkey, acc, rot, rot_mod);
    if (trace_acc) memcpy(trace_acc, acc, sizeof(uint32_t) * k * N);
    /* extract: Transpose + INTT per party (:441-449 / :498-506), then ModSwitch(qKS) */
    for (uint32_t u = 0; u < k; ++u) {
        orc_transpose_intt(c, acc + (size_t)u * N);
        for (uint32_t x = 0; x < N; ++x) acc[(size_t)u * N + x] = orc_round_qQ(acc[(size_t)u * N + x], qKS, Q);
    }
    if (c->p.scheme == 0) {
        orc_keyswitch_mntru(c, kskA, acc, out);
        if (bout) *bout = 0;
    } else {
        uint32_t bq = orc_round_qQ((Q >> 3) + 1, qKS, Q); /* :450-458 */
        orc_keyswitch_mklwe(c, kskA, kskB, acc, bq, out, bout);
    }
    free(rot);
    free(acc);
}

/* ---------- decryption ---------- */
/* ---------- key generation with the randomness as an input ----------
 * One uni-encryption evk = (d_j, f_j), j < ell, of the bit m under party secret s (s_inv = its inverse, EVALUATION):
 *   KeyGenXZW    mk-acc-xzw.cpp:132-186 (== mk-acc-xzw_B.cpp KeyGenXZW):  f_j = (NTT(e1_j) + g_j * R) * s_inv
 *                                                                         d_j = NTT(e0_j) + [m] g_j + R[j] * CRS_j
 *   KDMKeyGenXZW mk-acc-xzw.cpp:188-228:                                  d_j = NTT(e0_j) + [m] g_j * s_inv + R[j] * CRS_j
 * with g_j = B^(j+1) (Gpower[j+1]), R = NTT(r), and -- as written in the reference -- "skrPoly[i] * CRS[i]" multiplying
 * CRS_j by the SCALAR R[j] (slot j of R), not by the polynomial r.  r, e: coefficient-domain residues in [0, Q).
 * e layout [ell][2][N]: e[j][0] = noise of d_j (sampled first), e[j][1] = noise of f_j.  out layout [ell][2][N] (d | f). */
void orc_keygen_evk(const orc_ctx* c, const uint32_t* sinv, const uint32_t* crs, const uint32_t* r, const uint32_t* e, int m,
                    int kdm, uint32_t* out) {
    const uint32_t N = c->p.N, Q = c->p.Q, ell = c->ell;
    uint32_t* R = (uint32_t*)malloc(sizeof(uint32_t) * N);
    memcpy(R, r, sizeof(uint32_t) * N);
    orc_ntt(c, R, 0);
    uint32_t g = 1;
    for (uint32_t j = 0; j < ell; ++j) {
        g = mulmod(g, 1u << c->p.baseG_log2, Q);
        uint32_t* d = out + (size_t)(2 * j) * N;
        uint32_t* f = d + N;
        memcpy(d, e + (size_t)(2 * j) * N, sizeof(uint32_t) * N);
        memcpy(f, e + (size_t)(2 * j + 1) * N, sizeof(uint32_t) * N);
        orc_ntt(c, d, 0);
        orc_ntt(c, f, 0);
        for (uint32_t x = 0; x < N; ++x) {
            f[x] = mulmod(addmod(f[x], mulmod(g, R[x], Q), Q), sinv[x], Q);
            if (m) d[x] = addmod(d[x], kdm ? mulmod(g, sinv[x], Q) : g, Q);
            d[x] = addmod(d[x], mulmod(R[j], crs[(size_t)j * N + x], Q), Q);
        }
    }
    free(R);
}
/* Pkey[u][j] = NTT(e_j) - CRS_j * f_u  (binfhe-base-scheme.cpp:250-269); s = f_u in EVALUATION form; e, out: [ell][N] */
void orc_keygen_pkey(const orc_ctx* c, const uint32_t* s, const uint32_t* crs, const uint32_t* e, uint32_t* out) {
    const uint32_t N = c->p.N, Q = c->p.Q;
    for (uint32_t j = 0; j < c->ell; ++j) {
        uint32_t* o = out + (size_t)j * N;
        memcpy(o, e + (size_t)j * N, sizeof(uint32_t) * N);
        orc_ntt(c, o, 0);
        for (uint32_t x = 0; x < N; ++x) o[x] = submod(o[x], mulmod(crs[(size_t)j * N + x], s[x], Q), Q);
    }
}

/* Key-switch key generation with the randomness as an input, one party.
 * MNTRU  KeySwitchGen2 (mntru-pke.cpp:624-760), the independent digit-1 slice:  KSK = centred(E + gadget(s)) x centred(F^-1)
 *   mod qKS, where gadget(s) adds s_i * baseKS^j to column 0 of row i*dKS + j (:652-671), products accumulated in 64 bits
 *   and reduced once (mod_q, :408-418).  s_q [N]: the uni-encryption secret after SwitchModulus(qKS); e [N*dKS][n];
 *   finv [n][n] row-major; out [N*dKS][n].  (The reference fills E with copies of ONE sampled row, :646: the caller
 *   replicates it.) */
static int64_t centre(uint32_t v, uint32_t q) { return v > q / 2 ? (int64_t)v - (int64_t)q : (int64_t)v; }
void orc_keygen_ksk_mntru(const orc_ctx* c, const uint32_t* s_q, const uint32_t* e, const uint32_t* finv, uint32_t* out) {
    const uint32_t N = c->p.N, n = c->p.n, d = c->p.dKS, q = c->p.qKS;
    int64_t* row = (int64_t*)malloc(sizeof(int64_t) * n);
    int64_t* acc = (int64_t*)malloc(sizeof(int64_t) * n);
    for (uint32_t i = 0; i < N; ++i) {
        uint32_t pw = s_q[i] % q;
        for (uint32_t j = 0; j < d; ++j) {
            const size_t l = (size_t)i * d + j;
            for (uint32_t x = 0; x < n; ++x) row[x] = centre(x == 0 ? addmod(e[l * n] % q, pw, q) : e[l * n + x] % q, q);
            memset(acc, 0, sizeof(int64_t) * n);
            for (uint32_t x = 0; x < n; ++x)
                for (uint32_t y = 0; y < n; ++y) acc[y] += row[x] * centre(finv[(size_t)x * n + y], q);
            for (uint32_t y = 0; y < n; ++y) {
                int64_t v = acc[y] % (int64_t)q;
                out[l * n + y] = (uint32_t)(v < 0 ? v + q : v);
            }
            pw = mulmod(pw, c->p.baseKS % q, q);
        }
    }
    free(row);
    free(acc);
}
/* MKLWE KeySwitchGen (mklwe-pke.cpp:176-256), one party: B[i][a0][j] = e + sN_i * (a0 * baseKS^j) + <A[i][a0][j], s>  mod qKS.
 * A [N*baseKS*dKS][n] (the sampled vectors, part of the key as they are), e [rows], s [n] (LWE secret), sN_q [N]. */
void orc_keygen_ksk_mklwe(const orc_ctx* c, const uint32_t* A, const uint32_t* e, const uint32_t* s, const uint32_t* sN_q,
                          uint32_t* Bout) {
    const uint32_t N = c->p.N, n = c->p.n, d = c->p.dKS, q = c->p.qKS, base = c->p.baseKS;
    for (uint32_t i = 0; i < N; ++i)
        for (uint32_t a0 = 0; a0 < base; ++a0) {
            uint32_t pw = 1;
            for (uint32_t j = 0; j < d; ++j) {
                const size_t row = ((size_t)i * base + a0) * d + j;
                uint32_t b = addmod(e[row] % q, mulmod(sN_q[i] % q, mulmod(a0 % q, pw, q), q), q);
                for (uint32_t x = 0; x < n; ++x) b = addmod(b, mulmod(A[row * n + x] % q, s[x] % q, q), q);
                Bout[row] = b;
                pw = mulmod(pw, base % q, q);
            }
        }
}

int orc_decrypt(const orc_ctx* c, const uint32_t* sk, const uint32_t* ct, uint32_t b) {
    const uint32_t k = c->p.k, n = c->p.n, q = c->p.q, p = 4;
    uint32_t inner = 0;
    for (size_t x = 0; x < (size_t)k * n; ++x) inner = addmod(inner, mulmod(ct[x], sk[x] % q, q), q);
    if (c->p.scheme == 0) { /* mntru-pke.cpp:306-341 */
        inner = addmod(inner, q / p, q);
        return (int)(((uint64_t)p * inner) / q);
    }
    uint32_t r = submod(b, inner, q); /* mklwe-pke.cpp Decrypt */
    r = addmod(r, q / (p * 2), q);
    return (int)(((uint64_t)p * r) / q);
}

/* ---------- CPU baseline ---------- */
typedef struct {
    const orc_ctx* c;
    const uint32_t *bk, *pkey, *kskA, *kskB, *ctnand, *ct1, *ct2;
    uint32_t b1, b2;
    uint32_t* next;
    uint32_t gates;
    pthread_mutex_t* mu;
} bench_arg;

static void* bench_worker(void* v) {
    bench_arg* a = (bench_arg*)v;
    uint32_t* out = (uint32_t*)malloc(sizeof(uint32_t) * a->c->p.k * a->c->p.n);
    uint32_t bout;
    for (;;) {
        pthread_mutex_lock(a->mu);
        uint32_t g = (*a->next)++;
        pthread_mutex_unlock(a->mu);
        if (g >= a->gates) break;
        orc_eval_nand(a->c, a->bk, a->pkey, a->kskA, a->kskB, a->ctnand, a->ct1, a->b1, a->ct2, a->b2, out, &bout, NULL);
    }
    free(out);
    return NULL;
}

double orc_bench_gates(const orc_ctx* c, const uint32_t* bk, const uint32_t* pkey, const uint32_t* kskA, const uint32_t* kskB,
                       const uint32_t* ctnand, const uint32_t* ct1, uint32_t b1, const uint32_t* ct2, uint32_t b2,
                       uint32_t gates, uint32_t threads) {
    pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    uint32_t next = 0;
    bench_arg a = {c, bk, pkey, kskA, kskB, ctnand, ct1, ct2, b1, b2, &next, gates, &mu};
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * threads);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (uint32_t t = 0; t < threads; ++t) pthread_create(&th[t], NULL, bench_worker, &a);
    for (uint32_t t = 0; t < threads; ++t) pthread_join(th[t], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(th);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

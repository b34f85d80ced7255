/*
 * mkfhe_oracle.h -- plain-C restatement of the reference's multi-key bootstrapping path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under mkfhe_b200/ (the product) may include, link or call this; it is the checker
 * that tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg compare the CUDA path against.
 *
 * It follows the reference's algorithm step by step (EVALUATION-domain accumulator, monomial multiplication, HbProd as
 * written), NOT the reformulation the CUDA kernels use, so that agreement between the two is evidence for both.
 * Parity status: PINNED -- tests/test_oracle_vs_reference.py checks every function here bit-for-bit against the compiled
 * reference (oracle/_ref, built in place from /root/reference) and against the committed fixtures in tests/golden/
 * (generated from the compiled reference by tests/golden/make_golden.py); the NTT is additionally checked against the
 * reference's own known-answer test (src/core/unittest/UnitTestTransform.cpp:59-94).
 *
 * file:line citations are into /root/reference.
 */
#ifndef MKFHE_ORACLE_H
#define MKFHE_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_params {
    uint32_t k, N, Q, psi, q, qKS, n, baseG_log2, digitsG, baseKS, dKS;
    uint32_t variant; /* 0 = XZW (mk-acc-xzw.cpp), 1 = XZW_B (mk-acc-xzw_B.cpp) */
    uint32_t scheme;  /* 0 = MNTRU ciphertexts, 1 = MKLWE ciphertexts */
} orc_params;

typedef struct orc_ctx orc_ctx;

/* number theory used to derive Q and psi (binfhecontext.cpp:157-158; nbtheory-impl.h:183-231,334-379) */
uint32_t orc_first_prime(uint32_t nbits, uint32_t m);
uint32_t orc_previous_prime(uint32_t q, uint32_t m);
uint32_t orc_root_of_unity(uint32_t m, uint32_t Q); /* minimal primitive m-th root of unity mod Q */

orc_ctx* orc_create(const orc_params* p);
void orc_destroy(orc_ctx* c);

/* transformnat-impl.h:299-354 (forward, natural -> bit-reversed) and :491-552 (inverse); in place on N words */
void orc_ntt(const orc_ctx* c, uint32_t* poly, int inverse);
/* mk-acc.cpp:54-80; out is [digitsG-1][N] and is OVERWRITTEN (the reference accumulates into zero-initialised outputs) */
void orc_decompose(const orc_ctx* c, const uint32_t* poly, uint32_t* out);
/* mk-cryptoparameters.cpp:53-70: EVALUATION form of X^m - 1 (m < N) or -X^(m-N) - 1 (N <= m < 2N) */
void orc_monomial(const orc_ctx* c, uint32_t m, uint32_t* out);

/* UniEncAccumulatorXZW{,_B}::EvalAcc (mk-acc-xzw.cpp:89-130 / mk-acc-xzw_B.cpp:103-133) with the key layouts of
 * include/mkfhe_b200.h: bk [u][s][i][j][c][N] + evk*, pkey [u][j][N]; acc [k][N] EVALUATION in/out; rot [k][n]. */
void orc_eval_acc(const orc_ctx* c, const uint32_t* bk, const uint32_t* pkey, uint32_t* acc, const uint32_t* rot,
                  uint32_t rot_mod);

/* BinFHEScheme::EvalBinGate(NAND) (binfhe-base-scheme.cpp:380-515) for one gate.
 * MNTRU: ksk = [u][l < N*dKS][n] (digit-1 slice), ctnand [k][n], b* ignored.
 * MKLWE: kskA [u][i][a0][j][n], kskB [u][i][a0][j]; ctnand ignored.
 * If trace_acc != NULL it receives the accumulator after EvalAcc ([k][N], EVALUATION). */
void orc_eval_nand(const orc_ctx* c, const uint32_t* bk, const uint32_t* pkey, const uint32_t* kskA, const uint32_t* kskB,
                   const uint32_t* ctnand, const uint32_t* ct1, uint32_t b1, const uint32_t* ct2, uint32_t b2, uint32_t* out,
                   uint32_t* bout, uint32_t* trace_acc);

/* pieces of the epilogue, exposed for unit parity */
uint32_t orc_round_qQ(uint32_t v, uint32_t q, uint32_t Q); /* mntru-pke.cpp:12-17 == mklwe-pke.cpp:12-17 */
/* Transpose() + SetFormat(COEFFICIENT): poly-interface.h:443-450 -> poly-impl.h:312-365, in place on EVALUATION input */
void orc_transpose_intt(const orc_ctx* c, uint32_t* poly);
void orc_keyswitch_mntru(const orc_ctx* c, const uint32_t* ksk, const uint32_t* ctQN /*[k][N] mod qKS*/, uint32_t* out);
void orc_keyswitch_mklwe(const orc_ctx* c, const uint32_t* kskA, const uint32_t* kskB, const uint32_t* ctQN, uint32_t b,
                         uint32_t* out, uint32_t* bout);

/* decryption with exported secret material (mntru-pke.cpp:306-357 with F_col0 [k][n]; mklwe-pke.cpp Decrypt with s [k][n]) */
int orc_decrypt(const orc_ctx* c, const uint32_t* sk, const uint32_t* ct, uint32_t b);

/* key generation with the randomness supplied by the caller (mk-acc-xzw.cpp:132-228, binfhe-base-scheme.cpp:250-269);
 * layouts in mkfhe_oracle.c */
void orc_keygen_evk(const orc_ctx* c, const uint32_t* sinv, const uint32_t* crs, const uint32_t* r, const uint32_t* e, int m,
                    int kdm, uint32_t* out);
void orc_keygen_pkey(const orc_ctx* c, const uint32_t* s, const uint32_t* crs, const uint32_t* e, uint32_t* out);
/* key-switch key generation, one party (mntru-pke.cpp:624-760 digit-1 slice; mklwe-pke.cpp:176-256) */
void orc_keygen_ksk_mntru(const orc_ctx* c, const uint32_t* s_q, const uint32_t* e, const uint32_t* finv, uint32_t* out);
void orc_keygen_ksk_mklwe(const orc_ctx* c, const uint32_t* A, const uint32_t* e, const uint32_t* s, const uint32_t* sN_q,
                          uint32_t* Bout);

/* CPU baseline helper: run `gates` NAND bootstraps over `threads` pthreads (same inputs each), return wall seconds */
double orc_bench_gates(const orc_ctx* c, const uint32_t* bk, const uint32_t* pkey, const uint32_t* kskA, const uint32_t* kskB,
                       const uint32_t* ctnand, const uint32_t* ct1, uint32_t b1, const uint32_t* ct2, uint32_t b2,
                       uint32_t gates, uint32_t threads);

#ifdef __cplusplus
}
#endif
#endif

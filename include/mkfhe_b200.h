/*
 * mkfhe_b200.h -- C ABI of libmkfhe_b200.so: multi-key accumulator bootstrapping (blind rotation) of SKLC-FHE/MKFHE
 * on NVIDIA B200 (sm_100a).
 *
 * This is the drop-in boundary for ONE path of the reference: the body of
 *     BinFHEContext::EvalBinGate(NAND, ct1, ct2)                 src/binfhe/lib/binfhecontext.cpp:415-426
 *       -> BinFHEScheme::EvalBinGate (MNTRU / MKLWE)             src/binfhe/lib/binfhe-base-scheme.cpp:380-515
 *       -> BinFHEScheme::BootstrapGateCore                       src/binfhe/lib/binfhe-base-scheme.cpp:1004-1130
 *       -> UniEncAccumulator::EvalAcc  (the plug-in surface)     src/binfhe/include/mk-acc.h:54-80
 *            UniEncAccumulatorXZW::EvalAcc                       src/binfhe/lib/mk-acc-xzw.cpp:89-130
 *            UniEncAccumulatorXZW_B::EvalAcc                     src/binfhe/lib/mk-acc-xzw_B.cpp:103-133
 *       -> ModSwitch + KeySwitch2 / KeySwitch                    src/binfhe/lib/mntru-pke.cpp:359-374,763-823
 *                                                                src/binfhe/lib/mklwe-pke.cpp:160-171,260-290
 * (all file:line are into the reference tree).  Results are bit-identical to the reference's CPU code for the same
 * parameters, key bytes and input ciphertexts.
 *
 * Conventions
 *   - plain C: opaque handle, pointers and sizes; no C++ types and no exceptions cross this boundary.
 *   - every function returns 0 on success, a negative MKFHE_E* code on failure; mkfhe_last_error() gives the text.
 *     The reference throws OPENFHE_THROW(config_error, ...) (src/core/include/utils/exception.h:101-162); the C++
 *     accumulator class in integration/unienc_acc_b200.h turns non-zero codes back into those exceptions.
 *   - all residues are canonical uint32 (the reference's NativeInteger at NATIVE_SIZE=32, README.md:19).
 *   - polynomials in EVALUATION format use OpenFHE's slot order: in-place Cooley-Tukey output, slot j holds the value at
 *     psi^(2*bitrev(j)+1) (src/core/include/math/hal/intnat/transformnat-impl.h:299-354,714-727).
 *   - pointers named *_host_or_dev may be host or device memory (unified addressing decides); `_dev` entry points take
 *     device pointers only and enqueue on the context's stream without synchronising.
 *   - one context per GPU; calls on one context are serialised by the caller (as the reference's single-threaded path).
 *   - there is NO CPU fallback: without a CUDA device mkfhe_ctx_create fails with MKFHE_ENODEV.
 */
#ifndef MKFHE_B200_H
#define MKFHE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MKFHE_OK 0
#define MKFHE_EINVAL (-1)   /* bad argument / unsupported parameter combination */
#define MKFHE_ENODEV (-2)   /* no usable CUDA device (there is no CPU fallback)  */
#define MKFHE_ECUDA (-3)    /* CUDA runtime error; text in mkfhe_last_error       */
#define MKFHE_ENOKEY (-4)   /* a key this call needs has not been loaded          */
#define MKFHE_ENOMEM (-5)

/* accumulator variant: which UniEncAccumulator subclass is being replaced (binfhe-base-scheme.h:137-151) */
#define MKFHE_VARIANT_XZW 0u   /* UniEncAccumulatorXZW   : ternary secret, evk+ and evk- per coefficient (BINFHE_METHOD MKNTRU) */
#define MKFHE_VARIANT_XZW_B 1u /* UniEncAccumulatorXZW_B : binary secret, one evk per coefficient (MKNTRU_B, MKNTRU_LWE)       */
/* additive scheme either side of the accumulator */
#define MKFHE_SCHEME_MNTRU 0u /* matrix-NTRU ciphertexts: [k][n]      (boolean-mkntru) */
#define MKFHE_SCHEME_MKLWE 1u /* multi-key LWE ciphertexts: [k][n] + b (boolean-mklwe)  */

typedef struct mkfhe_ctx mkfhe_ctx;

/* The fields of UniEncCryptoParams / MNTRUCryptoParams / MKLWECryptoParams the path reads
 * (mk-cryptoparameters.h:124-181, mntru-cryptoparameters.h:59-83, mklwe-cryptoparameters.h:59-83). */
typedef struct mkfhe_params {
    uint32_t k;          /* parties                                                            */
    uint32_t N;          /* ring dimension (power of two, 8..2048)                             */
    uint32_t Q;          /* accumulator modulus, prime, Q = 1 mod 2N, Q < 2^30                 */
    uint32_t psi;        /* GetRootOfUnity(): minimal primitive 2N-th root of unity mod Q      */
    uint32_t q;          /* ciphertext modulus of the additive scheme                          */
    uint32_t qKS;        /* key-switching modulus (== q on both MK paths)                      */
    uint32_t n;          /* lattice dimension of the additive scheme                           */
    uint32_t baseG_log2; /* log2 of the gadget base                                            */
    uint32_t digitsG;    /* GetDigitsG() = ceil(log Q / log baseG); digitsG-1 digits are used  */
    uint32_t baseKS;     /* key-switching base (<= 256)                                        */
    uint32_t dKS;        /* ceil(log qKS / log baseKS)                                         */
    uint32_t variant;    /* MKFHE_VARIANT_*                                                    */
    uint32_t scheme;     /* MKFHE_SCHEME_*                                                     */
    int32_t device;      /* CUDA device ordinal; -1 = the calling thread's current device      */
} mkfhe_params;

/* ---- lifetime ---- */
int mkfhe_ctx_create(const mkfhe_params* p, mkfhe_ctx** out);
void mkfhe_ctx_destroy(mkfhe_ctx* ctx);
/* text of the last error on ctx (ctx == NULL: last mkfhe_ctx_create failure on this thread) */
const char* mkfhe_last_error(const mkfhe_ctx* ctx);
/* use `cuda_stream` (a cudaStream_t) for all later work of this context; NULL = the context's own stream.  Waits for the work
 * already enqueued on the previous stream (the context's workspaces and pooled scratch are ordered on one stream at a time).
 * ORDERING CONTRACT of every *_dev entry point: the call reads its inputs and writes its outputs in stream order on the
 * context's current stream, and returns without waiting.  The context's own stream is an ordinary (blocking) stream, so it is
 * ordered with the LEGACY default stream: a caller that produces inputs / consumes outputs on the legacy default stream
 * (PyTorch's default stream, plain cudaMemcpy) needs no extra synchronisation.  A caller working on any other stream -- a
 * cudaStreamNonBlocking stream, a torch side stream, cudaStreamPerThread -- must hand that stream to mkfhe_ctx_set_stream
 * (the handles cudaStreamLegacy = (void*)1 and cudaStreamPerThread = (void*)2 are accepted as they are) or order the work
 * itself with events. */
int mkfhe_ctx_set_stream(mkfhe_ctx* ctx, void* cuda_stream);
/* block until all work enqueued by this context has finished */
int mkfhe_ctx_synchronize(mkfhe_ctx* ctx);

/* ---- keys: caller owns the source buffers; each call copies/re-lays-out into HBM and returns ---- */

/* Bootstrapping key, flattened UniEncACCKeyImpl (mk-acckey.h; built at mk-acc-xzw.cpp:38-87 / mk-acc-xzw_B.cpp:38-101):
 *   [u < k][s < S][i < n][j < digitsG-1][c < 2][N]   then evk* = (*ek)[0][0][n] as [j][c][N]
 * S = 2 for XZW (s=0: evk+, s=1: evk-), 1 for XZW_B; c = 0: the "d" column, c = 1: the "f" column of UniEncEvalKeyImpl
 * (mk-evalkey.h); EVALUATION format.  words must equal (k*S*n + 1) * (digitsG-1) * 2 * N. */
int mkfhe_load_bk(mkfhe_ctx* ctx, const uint32_t* bk_host_or_dev, uint64_t words);
/* UniEncBTKey::Pkey (binfhe-base-scheme.h:65-83; built at binfhe-base-scheme.cpp:250-269): [u < k][j < digitsG-1][N], EVALUATION */
int mkfhe_load_pkey(mkfhe_ctx* ctx, const uint32_t* pkey_host_or_dev, uint64_t words);
/* MNTRUSwitchingKey2Impl digit-1 slice KSK2[u][1][l][.] (mntru-keyswitchkey2.h; KSK2[u][j] = j*KSK2[u][1] mod qKS,
 * mntru-pke.cpp:743-752): [u < k][l < N*dKS][n] */
int mkfhe_load_ksk_mntru(mkfhe_ctx* ctx, const uint32_t* ksk_host_or_dev, uint64_t words);
/* MKLWESwitchingKeyImpl (mklwe-keyswitchkey.h:28-29): A [u < k][i < N][a0 < baseKS][j < dKS][n], B [u][i][a0][j] */
int mkfhe_load_ksk_mklwe(mkfhe_ctx* ctx, const uint32_t* A_host_or_dev, uint64_t wordsA, const uint32_t* B_host_or_dev,
                         uint64_t wordsB);
/* m_ctNAND of BinFHEContext (binfhecontext.h:88; built by ctGateGen, binfhe-base-scheme.cpp:340-376): [u < k][n]; MNTRU only */
int mkfhe_load_ctnand(mkfhe_ctx* ctx, const uint32_t* ct_host_or_dev, uint64_t words);

/* Resident key images, for replicating keys between the GPUs of one box (north_star: "each GPU holds a replica of all parties'
 * bootstrapping keys ... NCCL over NVLink only to broadcast keys"; the reference has one address space and no such step).
 * An image is the key as the kernels read it from HBM -- bootstrapping key and Pkey in Montgomery form and register order,
 * key-switch key as padded u16 rows, ctNAND as loaded -- so a replica needs no conversion and the MKLWE key-switch key travels
 * as 403 MB instead of the 806 MB of its u32 boundary layout.  Rank 0 loads keys with mkfhe_load_* / mkfhe_keyfile_load; every
 * rank calls mkfhe_key_image (which allocates the image where no key is loaded yet) and broadcasts `bytes` bytes from rank 0's
 * pointer into the others' (ncclBroadcast / torch.distributed.broadcast on the raw device memory); receivers then call
 * mkfhe_key_image_commit.  Images are only meaningful between contexts created from identical mkfhe_params on the same
 * library build.  MKFHE_IMAGE_KSK exists on the fused key-switch path only (qKS < 2^16: every reference parameter set). */
enum { MKFHE_IMAGE_BK = 0, MKFHE_IMAGE_PKEY = 1, MKFHE_IMAGE_KSK = 2, MKFHE_IMAGE_CTNAND = 3 };
int mkfhe_key_image(mkfhe_ctx* ctx, int which, void** dev_ptr, uint64_t* bytes);
int mkfhe_key_image_commit(mkfhe_ctx* ctx, int which);

/* The decrypting parties' secret, only needed by mkfhe_decrypt: MNTRUPrivateKeyImpl::GetF_col0() (mntru-privatekey.h) or
 * MKLWEPrivateKeyImpl::GetElement() (mklwe-privatekey.h), [u < k][n] residues (reduced mod q at upload, as Decrypt's
 * SwitchModulus does for keys already stored mod q). */
int mkfhe_load_sk(mkfhe_ctx* ctx, const uint32_t* sk_host_or_dev, uint64_t words);

/* ---- the hot path ---- */

/* UniEncAccumulator::EvalAcc for `batch` independent accumulators (mk-acc.h:67-71).
 *   acc_in / acc_out : [batch][k][N] EVALUATION (may alias)
 *   rot              : [batch][k][n] -- the `ct` argument of EvalAcc: XZW rotates by rot*2N/rot_mod (mk-acc-xzw.cpp:110,125),
 *                      XZW_B by rot itself (mk-acc-xzw_B.cpp:120,126; rot_mod ignored)
 * Host pointers; synchronous. */
int mkfhe_eval_acc(mkfhe_ctx* ctx, const uint32_t* acc_in, const uint32_t* rot, uint32_t rot_mod, uint32_t* acc_out,
                   uint64_t batch);
/* same on device pointers, asynchronous on the context's stream */
int mkfhe_eval_acc_dev(mkfhe_ctx* ctx, const uint32_t* d_acc_in, const uint32_t* d_rot, uint32_t rot_mod, uint32_t* d_acc_out,
                       uint64_t batch);

/* BinFHEScheme::EvalBinGate(NAND) for `batch` independent gates (binfhe-base-scheme.cpp:380-515): input combine, initial
 * accumulator, blind rotation, extraction, modulus switch, key switch.
 *   ct1, ct2, out : [batch][k][n] mod q
 *   b1, b2, bout  : [batch]       (MKLWE only; pass NULL for MNTRU)
 * Host pointers; synchronous (H2D, kernels, D2H). */
int mkfhe_eval_nand(mkfhe_ctx* ctx, const uint32_t* ct1, const uint32_t* b1, const uint32_t* ct2, const uint32_t* b2,
                    uint32_t* out, uint32_t* bout, uint64_t batch);
/* same on device pointers, asynchronous on the context's stream */
int mkfhe_eval_nand_dev(mkfhe_ctx* ctx, const uint32_t* d_ct1, const uint32_t* d_b1, const uint32_t* d_ct2,
                        const uint32_t* d_b2, uint32_t* d_out, uint32_t* d_bout, uint64_t batch);

/* A circuit of NAND gates on a table of wires ("gate circuits on the batch API").  The reference has no
 * circuit entry: a caller loops over BinFHEContext::EvalBinGate (binfhecontext.cpp:415-419), one gate at a time, which on a GPU
 * would leave all but one SM idle.  Here gate g computes wires[out[g]] = NAND(wires[in1[g]], wires[in2[g]]); the library
 * levelises the netlist (level of a gate = 1 + the deepest gate feeding it) and runs every level as ONE batch of
 * mkfhe_eval_nand_dev, gathering and scattering ciphertext rows on the device.
 *   wires   : [n_wires][k][n] mod q, in/out (primary inputs filled by the caller, gate outputs written by the call)
 *   wires_b : [n_wires] (MKLWE; NULL for MNTRU)
 *   in1, in2, out : [n_gates] wire indices, HOST arrays in both variants.  Gates must be in topological order; every wire is
 *             written by at most one gate and never read before the gate that writes it; in1[g] != in2[g] (the reference
 *             rejects a gate on one ciphertext object: binfhe-base-scheme.cpp:382-383, 471-472).  Violations: MKFHE_EINVAL.
 *   n_levels: optional, receives the depth of the circuit = number of gate batches launched.
 * mkfhe_eval_circuit: host wires, synchronous.  _dev: wires in device memory, asynchronous on the context's stream (the index
 * arrays are consumed before the call returns).  mkfhe_circuit_levels: the levelisation alone (needs no GPU): level_of_gate
 * [n_gates] receives 0-based levels; returns the number of levels, or MKFHE_EINVAL for a netlist the calls above would refuse. */
int mkfhe_eval_circuit(mkfhe_ctx* ctx, uint32_t* wires, uint32_t* wires_b, uint64_t n_wires, const uint32_t* in1,
                       const uint32_t* in2, const uint32_t* out, uint64_t n_gates, uint32_t* n_levels);
int mkfhe_eval_circuit_dev(mkfhe_ctx* ctx, uint32_t* d_wires, uint32_t* d_wires_b, uint64_t n_wires, const uint32_t* in1,
                           const uint32_t* in2, const uint32_t* out, uint64_t n_gates, uint32_t* n_levels);
int64_t mkfhe_circuit_levels(uint64_t n_wires, const uint32_t* in1, const uint32_t* in2, const uint32_t* out, uint64_t n_gates,
                             uint32_t* level_of_gate);

/* mkfhe_eval_circuit_dev WITHOUT device allocation.  This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same results, same stream ordering as mkfhe_eval_circuit_dev; the per-level scratch
 * (level-sorted index arrays, gathered inputs, level outputs) lives in ONE file-scope device array of
 * MKFHE_STATIC_SCRATCH_WORDS 32-bit words (8 MiB) that every GPU loading the library holds until the process exits, instead of
 * the context's memory pool.  The call issues no cudaMalloc / cudaFree / pool allocation.
 *   Limit: mkfhe_circuit_static_words(params, n_gates, widest_level) <= MKFHE_STATIC_SCRATCH_WORDS, i.e.
 *          pad64(3 n_gates) + 3 pad64(widest k n) [+ 3 pad64(widest) for MKLWE] words (pad64 = round up to 64 words);
 *          levels of up to 695 gates at boolean-mklwe size (k = 2, n = 500), 621 / 311 / 155 / 77 gates at boolean-mkntru size for
 *          k = 2 / 4 / 8 / 16 (with 4 gates per widest-level gate in the netlist).  Larger circuits: MKFHE_EINVAL with the sizes in
 *          mkfhe_last_error -- use mkfhe_eval_circuit_dev.  Needs the fused key-switch path (qKS < 2^16: every reference set).
 *   Sharing: ALL contexts and calls on one GPU share the array and it keeps the previous call's contents (nothing in it has to
 *          start at zero: every word a call reads was written by that call first).  Calls that overlap on the GPU -- two
 *          contexts on different streams, or one context after mkfhe_ctx_set_stream without the wait it performs -- RACE on it;
 *          calls on one stream are ordered and safe. */
#define MKFHE_STATIC_SCRATCH_WORDS (2u << 20)
int mkfhe_eval_circuit_dev_static(mkfhe_ctx* ctx, uint32_t* d_wires, uint32_t* d_wires_b, uint64_t n_wires, const uint32_t* in1,
                                  const uint32_t* in2, const uint32_t* out, uint64_t n_gates, uint32_t* n_levels);
/* words of the static array a circuit needs (no GPU needed) and the array's size */
uint64_t mkfhe_circuit_static_words(const mkfhe_params* p, uint64_t n_gates, uint64_t widest_level);
uint64_t mkfhe_static_scratch_words(void);

/* MNTRUEncryptionScheme::Decrypt (mntru-pke.cpp:306-341) / MKLWEEncryptionScheme::Decrypt (mklwe-pke.cpp:66-97) for `batch`
 * ciphertexts: ct [batch][k][n] mod q, b [batch] (MKLWE; NULL for MNTRU), p = plaintext modulus (4 in both examples),
 * m [batch] receives the plaintexts.  Host pointers, synchronous; _dev: device pointers, asynchronous. */
int mkfhe_decrypt(mkfhe_ctx* ctx, const uint32_t* ct, const uint32_t* b, uint32_t p, uint32_t* m, uint64_t batch);
int mkfhe_decrypt_dev(mkfhe_ctx* ctx, const uint32_t* d_ct, const uint32_t* d_b, uint32_t p, uint32_t* d_m, uint64_t batch);

/* NativePoly::SetFormat on `batch` polynomials of N coefficients, in place (poly-impl.h:412-430 ->
 * transformnat-impl.h:299-354 forward, :491-552 inverse).  inverse = 0: COEFFICIENT -> EVALUATION. */
int mkfhe_ntt(mkfhe_ctx* ctx, uint32_t* polys, uint64_t batch, int inverse);
int mkfhe_ntt_dev(mkfhe_ctx* ctx, uint32_t* d_polys, uint64_t batch, int inverse);

/* UniEncAccumulator::SignedDigitDecompose on `batch` COEFFICIENT polynomials (mk-acc.cpp:54-80):
 * in [batch][N] -> out [batch][digitsG-1][N].  Host pointers; exposed for parity tests of that non-linear step. */
int mkfhe_decompose(mkfhe_ctx* ctx, const uint32_t* polys, uint32_t* out, uint64_t batch);

/* ---- introspection ---- */
/* number of kernel launches this context has issued since creation (bench.py's gpu_launches claim) */
uint64_t mkfhe_launch_count(const mkfhe_ctx* ctx);
/* SMs per gate of the last blind-rotation launch (mkfhe_eval_acc* / mkfhe_eval_nand* / circuits): 1, or the size of the thread-block
 * cluster (2, 4, 8) that small batches run on -- the reference has no counterpart (one thread per EvalAcc, mk-acc-xzw.cpp:89-130) */
uint32_t mkfhe_last_gate_sms(const mkfhe_ctx* ctx);
/* device time (CUDA events on the context's stream) of the last mkfhe_eval_nand[_dev] batch, split into the blind-rotation
 * kernel and the key-switch kernels; waits for that batch to finish.  bench.py's roofline numerators come from here. */
int mkfhe_last_gate_timing(mkfhe_ctx* ctx, float* rotate_ms, float* keyswitch_ms);
/* bytes of HBM held by this context (keys + tables + workspaces) */
uint64_t mkfhe_device_bytes(const mkfhe_ctx* ctx);
/* build-time identity: "mkfhe_b200 <version> sm_100a" */
/* Bootstrapping-key generation on the GPU with the randomness supplied by the caller.  Replaces
 * UniEncAccumulatorXZW{,_B}::KeyGenAcc -> KeyGenXZW / KDMKeyGenXZW (mk-acc-xzw.cpp:38-87,132-228; mk-acc-xzw_B.cpp:40-102) for a
 * host that samples (r, e0, e1) itself -- the reference samples inside KeyGenXZW (GetDggR / GetDgg) and its MKKeyGen is
 * clock-seeded, so bit-exact parity is defined on identical randomness.
 *   sinv [k][N]  : f_u^-1, EVALUATION (invskNTT, binfhe-base-scheme.cpp:214-229)      crs [ell][N]: EVALUATION
 *   r [E][N], e [E][ell][2][N] (e[.][j][0] = noise of d_j, [1] = noise of f_j): coefficient-domain residues in [0, Q)
 *   m [E]        : encrypted bit per evk (s == 1 for evk+, s == q-1 for evk-, 1 for evk*: mk-acc-xzw.cpp:61-84)
 *   E = k*S*n + 1 evks in the mkfhe_load_bk order [u][s][i] then evk*; (u, i) = (0, 0) and evk* are KDM encryptions.
 * bk_out (may be NULL when load != 0): mkfhe_load_bk layout, plain residues.  load != 0: also install it in ctx.
 * All pointers: host or device memory. */
int mkfhe_keygen_bk(mkfhe_ctx* ctx, const uint32_t* sinv, const uint32_t* crs, const uint32_t* r, const uint32_t* e,
                    const uint8_t* m, uint32_t* bk_out, int load);
/* Pkey[u][j] = NTT(e_j) - CRS_j * f_u (binfhe-base-scheme.cpp:250-269).  s [k][N]: f_u EVALUATION; e [k][ell][N] coefficient
 * residues; pkey_out: mkfhe_load_pkey layout (may be NULL when load != 0). */
int mkfhe_keygen_pkey(mkfhe_ctx* ctx, const uint32_t* s, const uint32_t* crs, const uint32_t* e, uint32_t* pkey_out, int load);

/* Key-switch key generation on the GPU with the noise supplied by the caller (all values residues mod qKS).
 * MNTRU: replaces MNTRUEncryptionScheme::KeySwitchGen2 (mntru-pke.cpp:624-760) for the independent digit-1 slice
 *   (KSK2[u][j] = j * KSK2[u][1], :743-752):  ksk[u] = centred(E_u + gadget(s_u)) x centred(F_u^-1) mod qKS.
 *   s_q [k][N]: uni-encryption secret coefficients after SwitchModulus(qKS) (:649); e [k][N*dKS][n] (the reference fills it
 *   with copies of one sampled row per party, :646); finv [k][n][n] = sk->GetF_inv(), row-major; ksk_out [k][N*dKS][n] =
 *   the mkfhe_load_ksk_mntru layout.
 * MKLWE: replaces MKLWEEncryptionScheme::KeySwitchGen (mklwe-pke.cpp:176-256):
 *   B[u][i][a0][j] = e + sN_q[u][i] * (a0 * baseKS^j) + <A[u][i][a0][j], s_u> mod qKS; A [k][N][baseKS][dKS][n] are the sampled
 *   vectors (part of the key as they are), e [k][N][baseKS][dKS], s [k][n] the LWE secret; B_out: mkfhe_load_ksk_mklwe's B.
 * load != 0 installs the key in ctx (outputs may then be NULL).  All pointers: host or device memory. */
int mkfhe_keygen_ksk_mntru(mkfhe_ctx* ctx, const uint32_t* s_q, const uint32_t* e, const uint32_t* finv, uint32_t* ksk_out, int load);
int mkfhe_keygen_ksk_mklwe(mkfhe_ctx* ctx, const uint32_t* A, const uint32_t* e, const uint32_t* s, const uint32_t* sN_q,
                           uint32_t* B_out, int load);

/* Batched Encrypt on the GPU with the noise supplied by the caller (host or device pointers; residues mod q).
 * MNTRU: replaces MNTRUEncryptionScheme::Encrypt (mntru-pke.cpp:158-205) for a batch:
 *   ct[b][u] = (e[b][u] + [u == 0] (m[b] mod pmod) (q / pmod) at index 0) x F_u^-1 mod q;  finv [k][n][n] = sk->GetF_inv(),
 *   row-major; e [B][k][n]; m [B] bytes; ct_out [B][k][n].
 * MKLWE: replaces MKLWEEncryptionScheme::Encrypt (mklwe-pke.cpp:36-64): b[g] = (m mod pmod)(q / pmod) + err[g] + sum_u <a[g][u], s_u>;
 *   the a vectors [B][k][n] are the caller's samples (the ciphertext's a part as they are); s [k][n]; err [B]; b_out [B]. */
int mkfhe_encrypt_mntru(mkfhe_ctx* ctx, const uint32_t* finv, const uint32_t* e, const uint8_t* m, uint32_t pmod, uint32_t* ct_out,
                        uint64_t batch);
int mkfhe_encrypt_mklwe(mkfhe_ctx* ctx, const uint32_t* s, const uint32_t* a, const uint32_t* err, const uint8_t* m, uint32_t pmod,
                        uint32_t* b_out, uint64_t batch);

/* Binary key file: the flattened layouts above, one checksummed section per key (format: mkfhe_b200/csrc/keyfile.h).
 * Replaces, for the GPU path, the reference's cereal save/load of the key object graphs (mk-acckey.h, mk-evalkey.h,
 * mntru-keyswitchkey2.h, mklwe-keyswitchkey.h; UniEncBTKey has no serializer of its own).  Host buffers; a NULL / zero-length
 * key is not written.  kskA = the MNTRU digit-1 slice (kskB, ctnand: MKLWE B column / MNTRU ctNAND).
 * mkfhe_keyfile_info: parameters and section sizes (words_by_tag[7], index 1 bk, 2 pkey, 3 ksk, 4 kskA, 5 kskB, 6 ctnand),
 * verifying every checksum -- needs no GPU.  mkfhe_keyfile_load: the file must match the context's parameter set. */
int mkfhe_keyfile_write(const char* path, const mkfhe_params* p, const uint32_t* bk, uint64_t bk_words, const uint32_t* pkey,
                        uint64_t pkey_words, const uint32_t* kskA, uint64_t kskA_words, const uint32_t* kskB, uint64_t kskB_words,
                        const uint32_t* ctnand, uint64_t ctnand_words);
int mkfhe_keyfile_info(const char* path, mkfhe_params* p, uint64_t* words_by_tag);
int mkfhe_keyfile_load(mkfhe_ctx* ctx, const char* path);
/* The reader does not trust the file: parameters are compared before any section is read; the section table is checked
 * against the file size and the word counts this parameter set implies (unknown / duplicate / mis-sized sections, kskA
 * without kskB: MKFHE_EINVAL); every checksum is verified BEFORE the first upload, so a refused file leaves the context's
 * keys as they were; allocation failures return MKFHE_ENOMEM. */

/* Binary ciphertext / wire-table file: n_ct ciphertexts in the batch layout of mkfhe_eval_nand / mkfhe_eval_circuit
 * (ct [n_ct][k][n] mod q, b [n_ct] for MKLWE, NULL for MNTRU) behind the same checksummed, parameter-stamped framing as the
 * key file (mkfhe_b200/csrc/keyfile.h).  Replaces, for batches, the reference's per-object cereal save/load of
 * MNTRUCiphertextImpl (mntru-ciphertext.h:116-125) and MKLWECiphertextImpl (mklwe-ciphertext.h:161-172).  No GPU needed.
 * mkfhe_ctfile_read: `expect` (may be NULL) must equal the file's parameters; at most capacity_ct ciphertexts are accepted. */
int mkfhe_ctfile_write(const char* path, const mkfhe_params* p, const uint32_t* ct, const uint32_t* b, uint64_t n_ct);
int mkfhe_ctfile_info(const char* path, mkfhe_params* p, uint64_t* n_ct);
int mkfhe_ctfile_read(const char* path, const mkfhe_params* expect, uint32_t* ct, uint32_t* b, uint64_t capacity_ct, uint64_t* n_ct);

const char* mkfhe_version(void);
/* INT32 pipe micro-benchmark (dependent IMAD chains on every SM): returns thread-level IMAD/s in *imad_per_s.
 * Used by bench.py as the denominator of the integer roofline (SURVEY.md section 8d asks for a measured peak). */
int mkfhe_measure_int32_peak(mkfhe_ctx* ctx, double* imad_per_s);
/* Modular-butterfly micro-benchmark: register-resident lazy Shoup butterflies (the instruction sequence of the transforms,
 * replacing transformnat-impl.h:299-354 inner loops) on every SM, no memory traffic; returns butterflies/s.  bench.py quotes
 * the gate kernel's modular multiplications per second (SURVEY.md section 8d figure) against this measured peak. */
int mkfhe_measure_butterfly_peak(mkfhe_ctx* ctx, double* bfly_per_s);

#ifdef __cplusplus
}
#endif
#endif /* MKFHE_B200_H */

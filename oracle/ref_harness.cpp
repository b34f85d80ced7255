// C API over the UNMODIFIED reference (compiled in place from /root/reference by oracle/Makefile into
// oracle/_ref/libmkfhe_ref.so).  TEST INFRASTRUCTURE ONLY: used by tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py as the checker and the CPU baseline; never by the product path.
//
// What it drives (all file:line into /root/reference/src/binfhe):
//   parameters      include/mk-cryptoparameters.h:124-181, include/mntru-cryptoparameters.h:59-83,
//                   include/mklwe-cryptoparameters.h:59-83, Q rule lib/binfhecontext.cpp:157-158
//   key generation  lib/binfhe-base-scheme.cpp:198-338 (MKKeyGen x2), :340-376 (ctGateGen),
//                   lib/mntru-pke.cpp:108-139 (KeyGen), lib/mklwe-pke.cpp (KeyGenBinary)
//   hot path        lib/binfhe-base-scheme.cpp:380-515 (EvalBinGate x2), :1004-1130 (BootstrapGateCore x2),
//                   lib/mk-acc-xzw.cpp:89-130, lib/mk-acc-xzw_B.cpp:103-133 (EvalAcc)
//   plug-in surface include/mk-acc.h:54-80 (UniEncAccumulator) -- ref_use_accumulator_b200() swaps in the
//                   drop-in class from integration/unienc_acc_b200.h so the reference's own EvalBinGate runs
//                   on the GPU library.
//   batched entry   ref_batch_b200_* drive integration/binfhe_batch_b200.h (EvalBinGateBatch over the reference's ciphertext
//                   and key objects) and decrypt its outputs with the reference's Decrypt.
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "binfhecontext.h"
#include "binfhe_batch_b200.h"
#include "unienc_acc_b200.h"

using namespace lbcrypto;

namespace {

// BinFHEScheme keeps its accumulator/scheme members protected; a derived class is the supported way in.
struct SchemeAccess : public BinFHEScheme {
    explicit SchemeAccess(BINFHE_METHOD m) : BinFHEScheme(m) {}
    std::shared_ptr<UniEncAccumulator>& Acc() { return UniEncACCscheme; }
    std::shared_ptr<MNTRUEncryptionScheme>& MNTRU() { return MNTRUscheme; }
    std::shared_ptr<MKLWEEncryptionScheme>& MKLWE() { return MKLWEscheme; }
};

struct RefCtx {
    int scheme{0};  // 0 = MKNTRU (XZW, matrix-NTRU ciphertexts), 1 = MKNTRU_LWE (XZW_B, MK-LWE ciphertexts)
    uint32_t k{0}, n{0}, N{0}, q{0}, qKS{0}, Q{0}, baseG{0}, baseKS{0}, digitsG{0}, dKS{0};
    BINFHE_METHOD method{MKNTRU};
    std::shared_ptr<MNTRUCryptoParams> mntru;
    std::shared_ptr<MKLWECryptoParams> mklwe;
    std::shared_ptr<UniEncCryptoParams> unienc;
    std::shared_ptr<BinFHECryptoParams> params;
    std::shared_ptr<SchemeAccess> sch;
    MNTRUPrivateKey skN;
    MKLWEPrivateKey skL;
    UniEncBTKey ek{};
    MNTRUCiphertext ctNAND;
    std::shared_ptr<BinFHEBatchB200> batch;  // integration/binfhe_batch_b200.h, created by ref_batch_b200_load
    std::string err;
};

inline const uint32_t* raw(const NativeVector& v) {
    static_assert(sizeof(NativeInteger) == sizeof(uint32_t), "reference must be built with NATIVEINT=32");
    return reinterpret_cast<const uint32_t*>(&v[0]);
}
inline void put(uint32_t* dst, const NativeVector& v) { std::memcpy(dst, raw(v), sizeof(uint32_t) * v.GetLength()); }
inline void put(uint32_t* dst, const NativePoly& p) { put(dst, p.GetValues()); }

NativeVector mkvec(const uint32_t* src, uint32_t len, uint32_t mod) {
    NativeVector v(len, NativeInteger(mod));
    for (uint32_t i = 0; i < len; ++i) v[i] = NativeInteger(src[i]);
    return v;
}

MNTRUCiphertext mk_mntru_ct(RefCtx* c, const uint32_t* ct, uint32_t mod) {
    std::vector<NativeVector> e;
    for (uint32_t u = 0; u < c->k; ++u) e.push_back(mkvec(ct + size_t(u) * c->n, c->n, mod));
    return std::make_shared<MNTRUCiphertextImpl>(std::move(e));
}
MKLWECiphertext mk_mklwe_ct(RefCtx* c, const uint32_t* ct, uint32_t b, uint32_t mod) {
    std::vector<NativeVector> a;
    for (uint32_t u = 0; u < c->k; ++u) a.push_back(mkvec(ct + size_t(u) * c->n, c->n, mod));
    return std::make_shared<MKLWECiphertextImpl>(std::move(a), NativeInteger(b));
}

template <class F>
int guarded(RefCtx* c, F&& f) {
    try {
        f();
        return 0;
    }
    catch (const std::exception& e) {
        if (c) c->err = e.what();
        return 1;
    }
    catch (...) {
        if (c) c->err = "unknown exception";
        return 2;
    }
}

}  // namespace

extern "C" {

// scheme: 0 = MKNTRU, 1 = MKNTRU_LWE.  keyDist: 0 = UNIFORM_TERNARY, 1 = BINARY (ignored: fixed per scheme as in the table).
void* ref_create(int scheme, uint32_t k, uint32_t numberBits, uint32_t cyclOrder, uint32_t n, uint32_t q, uint32_t qKS,
                 double stdDev, uint32_t baseKS, uint32_t baseG, uint32_t baseR) {
    auto* c = new RefCtx();
    int rc  = guarded(c, [&] {
        c->scheme = scheme;
        c->method = scheme == 0 ? MKNTRU : MKNTRU_LWE;
        c->k = k; c->n = n; c->N = cyclOrder / 2; c->q = q; c->qKS = qKS; c->baseG = baseG; c->baseKS = baseKS;
        // lib/binfhecontext.cpp:157-158
        NativeInteger Q(PreviousPrime<NativeInteger>(FirstPrime<NativeInteger>(numberBits, cyclOrder), cyclOrder));
        c->Q = Q.ConvertToInt<uint32_t>();
        SecretKeyDist kd = scheme == 0 ? UNIFORM_TERNARY : BINARY;
        c->unienc = std::make_shared<UniEncCryptoParams>(k, c->N, Q, NativeInteger(q), baseG, baseR, c->method, stdDev, kd,
                                                         false, 10);
        if (scheme == 0) {
            c->mntru  = std::make_shared<MNTRUCryptoParams>(k, n, c->N, NativeInteger(q), Q, NativeInteger(qKS), stdDev,
                                                           baseKS, kd);
            c->params = std::make_shared<BinFHECryptoParams>(c->mntru, c->unienc);
        }
        else {
            c->mklwe  = std::make_shared<MKLWECryptoParams>(k, n, c->N, NativeInteger(q), Q, NativeInteger(qKS), stdDev,
                                                           baseKS, kd);
            c->params = std::make_shared<BinFHECryptoParams>(c->mklwe, c->unienc);
        }
        c->sch     = std::make_shared<SchemeAccess>(c->method);
        c->digitsG = c->unienc->GetDigitsG();
        c->dKS     = static_cast<uint32_t>(std::ceil(log(double(qKS)) / log(double(baseKS))));
    });
    if (rc) {
        fprintf(stderr, "ref_create failed: %s\n", c->err.c_str());
        delete c;
        return nullptr;
    }
    return c;
}

void ref_destroy(void* h) { delete static_cast<RefCtx*>(h); }
const char* ref_last_error(void* h) { return static_cast<RefCtx*>(h)->err.c_str(); }

// out[0..9] = k, n, N, q, qKS, Q, baseG, digitsG, baseKS, dKS ; out[10] = psi (minimal primitive 2N-th root)
void ref_get_params(void* h, uint32_t* out) {
    auto* c = static_cast<RefCtx*>(h);
    out[0] = c->k; out[1] = c->n; out[2] = c->N; out[3] = c->q; out[4] = c->qKS; out[5] = c->Q;
    out[6] = c->baseG; out[7] = c->digitsG; out[8] = c->baseKS; out[9] = c->dKS;
    out[10] = c->unienc->GetPolyParams()->GetRootOfUnity().ConvertToInt<uint32_t>();
}

// Secret keys + bootstrapping keys (+ ctNAND for MKNTRU): the reference's own key generation, nothing restated.
int ref_keygen(void* h) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        if (c->scheme == 0) {
            c->skN    = c->sch->MNTRU()->KeyGen(c->k, c->n, NativeInteger(c->qKS));  // lib/binfhecontext.cpp:235-241
            c->ek     = c->sch->MKKeyGen(c->params, c->skN, SYM_ENCRYPT);
            c->ctNAND = c->sch->ctGateGen(c->params, c->skN, NAND);
        }
        else {
            c->skL = c->sch->MKLWE()->KeyGenBinary(c->k, c->n, NativeInteger(c->qKS));
            c->ek  = c->sch->MKKeyGen(c->params, c->skL, SYM_ENCRYPT);
        }
    });
}

// ---- flattened exports (layouts are the C-ABI layouts of include/mkfhe_b200.h) ----

// Bootstrapping key: [u<k][s<signs][i<n][j<ell][c<2][N] followed by evk* [j][c][N]; signs = 2 (MKNTRU) or 1 (MKNTRU_LWE).
uint64_t ref_bk_words(void* h) {
    auto* c       = static_cast<RefCtx*>(h);
    uint64_t ell  = c->digitsG - 1;
    uint64_t sgn  = c->scheme == 0 ? 2 : 1;
    return (uint64_t(c->k) * sgn * c->n + 1) * ell * 2 * c->N;
}
int ref_export_bk(void* h, uint32_t* out) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        const uint32_t ell = c->digitsG - 1, sgn = c->scheme == 0 ? 2 : 1, N = c->N;
        size_t off = 0;
        auto dump  = [&](const UniEncEvalKey& e) {
            for (uint32_t j = 0; j < ell; ++j)
                for (uint32_t col = 0; col < 2; ++col) {
                    put(out + off, (*e)[j][col]);
                    off += N;
                }
        };
        for (uint32_t u = 0; u < c->k; ++u)
            for (uint32_t s = 0; s < sgn; ++s)
                for (uint32_t i = 0; i < c->n; ++i) dump((*c->ek.BSkey)[u][s][i]);
        dump((*c->ek.BSkey)[0][0][c->n]);
    });
}
// Pkey: [u<k][j<ell][N]
int ref_export_pkey(void* h, uint32_t* out) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        size_t off = 0;
        for (uint32_t u = 0; u < c->k; ++u)
            for (uint32_t j = 0; j + 1 < c->digitsG; ++j) {
                put(out + off, c->ek.Pkey[u][j]);
                off += c->N;
            }
    });
}
// MNTRU key-switch key, digit-1 slice: [u<k][l<N*dKS][n].  Returns in *mismatch the number of (u,j,l,i) entries for
// which KSK2[u][j][l][i] != j*KSK2[u][1][l][i] mod qKS (expected 0: lib/mntru-pke.cpp:743-752).
int ref_export_ksk_mntru(void* h, uint32_t* out, uint64_t* mismatch) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        const auto& K  = c->ek.KSkey2->GetElements();
        uint64_t bad   = 0;
        const size_t L = size_t(c->N) * c->dKS;
        for (uint32_t u = 0; u < c->k; ++u)
            for (size_t l = 0; l < L; ++l) {
                uint32_t* row = out + (size_t(u) * L + l) * c->n;
                put(row, K[u][1][l]);
                for (uint32_t j = 0; j < c->baseKS; ++j) {
                    const uint32_t* kj = raw(K[u][j][l]);
                    for (uint32_t i = 0; i < c->n; ++i)
                        bad += (kj[i] != uint32_t((uint64_t(row[i]) * j) % c->qKS));
                }
            }
        if (mismatch) *mismatch = bad;
    });
}
// MK-LWE key-switch key: A [u][i<N][a0<baseKS][j<dKS][n], B [u][i][a0][j]
int ref_export_ksk_mklwe(void* h, uint32_t* A, uint32_t* B) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        const auto& KA = c->ek.LKSkey->GetElementsA();
        const auto& KB = c->ek.LKSkey->GetElementsB();
        size_t off     = 0;
        for (uint32_t u = 0; u < c->k; ++u)
            for (uint32_t i = 0; i < c->N; ++i)
                for (uint32_t a0 = 0; a0 < c->baseKS; ++a0)
                    for (uint32_t j = 0; j < c->dKS; ++j) {
                        put(A + off * c->n, KA[u][i][a0][j]);
                        B[off] = KB[u][i][a0][j].ConvertToInt<uint32_t>();
                        ++off;
                    }
    });
}
int ref_export_ctnand(void* h, uint32_t* out) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        for (uint32_t u = 0; u < c->k; ++u) put(out + size_t(u) * c->n, c->ctNAND->GetElements()[u]);
    });
}
// Secret material, for decryption by the independent C oracle: MKNTRU -> F_col0 [k][n] (mod qKS); MKNTRU_LWE -> s [k][n].
int ref_export_sk(void* h, uint32_t* out) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        for (uint32_t u = 0; u < c->k; ++u) {
            if (c->scheme == 0)
                put(out + size_t(u) * c->n, c->skN->GetF_col0()[u]);
            else
                put(out + size_t(u) * c->n, c->skL->GetElement()[u]);
        }
    });
}

// Uni-encryption secrets f_u in EVALUATION form [k][N] (UniEncBTKey::f, binfhe-base-scheme.cpp:229) and the common reference
// string CRS [ell][N] (EVALUATION, mk-cryptoparameters.h): the inputs of KeyGenAcc / Pkey generation besides the randomness.
// Used by the key-generation parity tests to recover the randomness the reference sampled (it is not exposed otherwise).
int ref_export_unienc_sk(void* h, uint32_t* out) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        for (uint32_t u = 0; u < c->k; ++u) put(out + size_t(u) * c->N, c->ek.f[u]);
    });
}
// f_u in COEFFICIENT form [k][N] mod Q (UniEncBTKey::fvec): the secret the key-switch keys encrypt
int ref_export_unienc_sk_coef(void* h, uint32_t* out) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        for (uint32_t u = 0; u < c->k; ++u) put(out + size_t(u) * c->N, c->ek.fvec[u]);
    });
}
// MNTRU secret matrices [k][n][n] mod qKS, row-major as stored (mntru-privatekey.h:52,104): inverse = 0 -> F, 1 -> F^-1
int ref_export_mntru_F(void* h, uint32_t* out, int inverse) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        if (c->scheme != 0) throw std::runtime_error("not an MNTRU context");
        const auto& F = inverse ? c->skN->GetF_inv() : c->skN->GetF();
        for (uint32_t u = 0; u < c->k; ++u)
            for (uint32_t l = 0; l < c->n; ++l) put(out + (size_t(u) * c->n + l) * c->n, F[u][l]);
    });
}
int ref_export_crs(void* h, uint32_t* out) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        const auto& crs = c->unienc->GetCRS();
        for (uint32_t j = 0; j + 1 < c->digitsG; ++j) {
            NativePoly p = crs[j];
            p.SetFormat(Format::EVALUATION);
            put(out + size_t(j) * c->N, p);
        }
    });
}

// ---- encrypt / decrypt (reference code) ----
int ref_encrypt(void* h, int m, uint32_t* ct, uint32_t* b) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        if (c->scheme == 0) {
            auto x = c->sch->MNTRU()->Encrypt(c->mntru, c->skN, m, 4, NativeInteger(c->q));
            for (uint32_t u = 0; u < c->k; ++u) put(ct + size_t(u) * c->n, x->GetElements()[u]);
            if (b) *b = 0;
        }
        else {
            auto x = c->sch->MKLWE()->Encrypt(c->mklwe, c->skL, m, 4, NativeInteger(c->q));
            for (uint32_t u = 0; u < c->k; ++u) put(ct + size_t(u) * c->n, x->GetA()[u]);
            *b = x->GetB().ConvertToInt<uint32_t>();
        }
    });
}
int ref_decrypt(void* h, const uint32_t* ct, uint32_t b, int* m) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        if (c->scheme == 0) {
            MNTRUPlaintext r = 0;
            ConstMNTRUCiphertext x = mk_mntru_ct(c, ct, c->q);
            c->sch->MNTRU()->Decrypt(c->mntru, c->skN, x, &r, 4);
            *m = int(r);
        }
        else {
            MKLWEPlaintext r = 0;
            ConstMKLWECiphertext x = mk_mklwe_ct(c, ct, b, c->q);
            c->sch->MKLWE()->Decrypt(c->mklwe, c->skL, x, &r, 4);
            *m = int(r);
        }
    });
}

// ---- hot path ----

// UniEncAccumulator::EvalAcc on caller-provided state: acc [k][N] EVALUATION in/out, rot [k][n] with modulus rot_mod
// (q for MKNTRU -> c = rot*2N/q ; 2N for MKNTRU_LWE -> c = rot).
int ref_eval_acc(void* h, uint32_t* acc, const uint32_t* rot, uint32_t rot_mod) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        auto pp = c->unienc->GetPolyParams();
        std::vector<NativePoly> el;
        for (uint32_t u = 0; u < c->k; ++u) {
            NativePoly p(pp, Format::EVALUATION, true);
            p.SetValues(mkvec(acc + size_t(u) * c->N, c->N, c->Q), Format::EVALUATION);
            el.push_back(std::move(p));
        }
        MKACCCiphertext a = std::make_shared<MKACCCiphertextImpl>(std::move(el));
        std::vector<NativeVector> ct;
        for (uint32_t u = 0; u < c->k; ++u) ct.push_back(mkvec(rot + size_t(u) * c->n, c->n, rot_mod));
        ConstUniEncACCKey bk = c->ek.BSkey;
        c->sch->Acc()->EvalAcc(c->unienc, bk, c->ek.Pkey, c->ek.f, a, ct);
        for (uint32_t u = 0; u < c->k; ++u) put(acc + size_t(u) * c->N, a->GetElements()[u]);
    });
}

// BinFHEScheme::EvalBinGate(NAND): ct1, ct2, out are [k][n]; b1, b2, bout only for MKNTRU_LWE.
int ref_eval_gate(void* h, const uint32_t* ct1, uint32_t b1, const uint32_t* ct2, uint32_t b2, uint32_t* out, uint32_t* bout) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        if (c->scheme == 0) {
            ConstMNTRUCiphertext x1 = mk_mntru_ct(c, ct1, c->q);
            ConstMNTRUCiphertext x2 = mk_mntru_ct(c, ct2, c->q);
            ConstMNTRUCiphertext nd = c->ctNAND;
            auto r = c->sch->EvalBinGate(c->params, NAND, c->ek, x1, x2, nd);
            for (uint32_t u = 0; u < c->k; ++u) put(out + size_t(u) * c->n, r->GetElements()[u]);
            if (bout) *bout = 0;
        }
        else {
            ConstMKLWECiphertext x1 = mk_mklwe_ct(c, ct1, b1, c->q);
            ConstMKLWECiphertext x2 = mk_mklwe_ct(c, ct2, b2, c->q);
            auto r = c->sch->EvalBinGate(c->params, NAND, c->ek, x1, x2);
            for (uint32_t u = 0; u < c->k; ++u) put(out + size_t(u) * c->n, r->GetA()[u]);
            *bout = r->GetB().ConvertToInt<uint32_t>();
        }
    });
}

// NativePoly::SetFormat on `batch` polynomials of length N (in place).  inverse = 0: COEFFICIENT -> EVALUATION.
int ref_ntt(void* h, uint32_t* polys, uint64_t batch, int inverse) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        auto pp = c->unienc->GetPolyParams();
        for (uint64_t b = 0; b < batch; ++b) {
            NativePoly p(pp, inverse ? Format::EVALUATION : Format::COEFFICIENT, true);
            p.SetValues(mkvec(polys + b * c->N, c->N, c->Q), inverse ? Format::EVALUATION : Format::COEFFICIENT);
            p.SetFormat(inverse ? Format::COEFFICIENT : Format::EVALUATION);
            put(polys + b * c->N, p);
        }
    });
}

// UniEncAccumulator::SignedDigitDecompose on one COEFFICIENT polynomial: out [ell][N]
int ref_decompose(void* h, const uint32_t* poly, uint32_t* out) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        auto pp = c->unienc->GetPolyParams();
        NativePoly p(pp, Format::COEFFICIENT, true);
        p.SetValues(mkvec(poly, c->N, c->Q), Format::COEFFICIENT);
        std::vector<NativePoly> d(c->digitsG - 1, NativePoly(pp, Format::COEFFICIENT, true));
        c->sch->Acc()->SignedDigitDecompose(c->unienc, p, d);
        for (uint32_t j = 0; j + 1 < c->digitsG; ++j) put(out + size_t(j) * c->N, d[j]);
    });
}

// CPU baseline: `threads` workers each run EvalBinGate(NAND) on fresh encryptions until `gates` have been completed in
// total (shared read-only keys; one warm-up gate first so the static NTT tables are filled before threads start,
// src/core/include/math/hal/intnat/transformnat-impl.h:593-596).  Returns wall seconds in *seconds.
int ref_bench_gates(void* h, uint32_t gates, uint32_t threads, double* seconds) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        const uint32_t kn = c->k * c->n;
        std::vector<uint32_t> a(kn), b(kn), o(kn);
        uint32_t ba = 0, bb = 0, bo = 0;
        ref_encrypt(h, 1, a.data(), &ba);
        ref_encrypt(h, 0, b.data(), &bb);
        if (ref_eval_gate(h, a.data(), ba, b.data(), bb, o.data(), &bo)) throw std::runtime_error(c->err);
        std::atomic<uint32_t> next{0};
        std::atomic<int> failed{0};
        auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> pool;
        for (uint32_t t = 0; t < threads; ++t)
            pool.emplace_back([&] {
                std::vector<uint32_t> out(kn);
                uint32_t bout = 0;
                while (next.fetch_add(1) < gates)
                    if (ref_eval_gate(h, a.data(), ba, b.data(), bb, out.data(), &bout)) failed = 1;
            });
        for (auto& th : pool) th.join();
        *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (failed) throw std::runtime_error("EvalBinGate failed in a worker");
    });
}

// Single-gate latency as the reference's example measures it (src/binfhe/examples/boolean-mkntru.cpp:34-38,
// boolean-mklwe.cpp:33-35: one EvalBinGate(NAND) on fresh encryptions of (1, 0), one thread): `reps` calls, each timed
// separately with steady_clock around the EvalBinGate call alone; ms[i] = wall milliseconds of call i.  One untimed call
// first (static NTT tables).
int ref_time_gate(void* h, uint32_t reps, double* ms) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        const uint32_t kn = c->k * c->n;
        std::vector<uint32_t> a(kn), b(kn);
        uint32_t ba = 0, bb = 0;
        ref_encrypt(h, 1, a.data(), &ba);
        ref_encrypt(h, 0, b.data(), &bb);
        for (uint32_t r = 0; r <= reps; ++r) {
            std::chrono::steady_clock::time_point t0, t1;
            if (c->scheme == 0) {
                ConstMNTRUCiphertext x1 = mk_mntru_ct(c, a.data(), c->q), x2 = mk_mntru_ct(c, b.data(), c->q), nd = c->ctNAND;
                t0 = std::chrono::steady_clock::now();
                auto out = c->sch->EvalBinGate(c->params, NAND, c->ek, x1, x2, nd);
                t1 = std::chrono::steady_clock::now();
            }
            else {
                ConstMKLWECiphertext x1 = mk_mklwe_ct(c, a.data(), ba, c->q), x2 = mk_mklwe_ct(c, b.data(), bb, c->q);
                t0 = std::chrono::steady_clock::now();
                auto out = c->sch->EvalBinGate(c->params, NAND, c->ek, x1, x2);
                t1 = std::chrono::steady_clock::now();
            }
            if (r > 0) ms[r - 1] = std::chrono::duration<double, std::milli>(t1 - t0).count();
        }
    });
}

// Swap the reference's CPU accumulator for the drop-in GPU one (integration/unienc_acc_b200.h); `lib` is the path of
// libmkfhe_b200.so.  After this call ref_eval_acc / ref_eval_gate run the reference's own orchestration code with the
// blind rotation executed by the B200 library through its C ABI.  restore = 1 puts the CPU accumulator back.
int ref_use_accumulator_b200(void* h, const char* lib, int restore) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        if (restore) {
            if (c->scheme == 0)
                c->sch->Acc() = std::make_shared<UniEncAccumulatorXZW>();
            else
                c->sch->Acc() = std::make_shared<UniEncAccumulatorXZW_B>();
            return;
        }
        auto acc = std::make_shared<UniEncAccumulatorB200>(
            lib, c->scheme == 0 ? UniEncAccumulatorB200::XZW : UniEncAccumulatorB200::XZW_B);
        // no LoadKeys call: the accumulator uploads the keys the first time EvalAcc sees them (what INTEGRATION.md section 2
        // asks of a maintainer is the make_shared line alone)
        c->sch->Acc() = acc;
    });
}

// ---- the batched entry a maintainer adds (integration/binfhe_batch_b200.h), driven with the reference's own objects ----

// Create the batch evaluator on `lib` (path of libmkfhe_b200.so) and upload this context's keys from the UniEncBTKey the
// reference generated (and ctNAND for MKNTRU).
int ref_batch_b200_load(void* h, const char* lib) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        c->batch = std::make_shared<BinFHEBatchB200>(lib);
        c->batch->LoadKeys(c->params, c->method, c->ek);
        if (c->scheme == 0) c->batch->LoadGateCiphertext(c->ctNAND);
    });
}

// SaveKeyFile from the reference's own key objects (no GPU involved), and a FRESH evaluator created from that file alone.
int ref_batch_b200_save_keyfile(void* h, const char* lib, const char* path) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        BinFHEBatchB200 w(lib);
        w.SaveKeyFile(path, c->params, c->method, c->ek, c->scheme == 0 ? c->ctNAND : nullptr);
    });
}
int ref_batch_b200_load_keyfile(void* h, const char* lib, const char* path) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        c->batch = std::make_shared<BinFHEBatchB200>(lib);
        c->batch->LoadKeyFile(path);
    });
}
// Ciphertext table round trip through the reference's objects: n encryptions of bits[] -> SaveCiphertexts -> LoadCiphertexts*;
// flat_out receives the loaded table, dec the reference's Decrypt of every loaded object.  Needs a loaded evaluator (for the
// parameter stamp); touches no GPU memory.
int ref_batch_b200_ct_roundtrip(void* h, const char* path, const uint8_t* bits, uint64_t n_ct, uint32_t* flat_in, uint32_t* b_in,
                                uint32_t* flat_out, uint32_t* b_out, int* dec) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        if (!c->batch) throw std::runtime_error("ref_batch_b200_load has not been called");
        const size_t kn = size_t(c->k) * c->n;
        if (c->scheme == 0) {
            std::vector<MNTRUCiphertext> in;
            for (uint64_t g = 0; g < n_ct; ++g) {
                auto x = c->sch->MNTRU()->Encrypt(c->mntru, c->skN, bits[g], 4, NativeInteger(c->q));
                for (uint32_t u = 0; u < c->k; ++u) put(flat_in + g * kn + size_t(u) * c->n, x->GetElements()[u]);
                in.push_back(x);
            }
            c->batch->SaveCiphertexts(path, in);
            auto out = c->batch->LoadCiphertextsNTRU(path);
            if (out.size() != n_ct) throw std::runtime_error("ciphertext count changed on the way through the file");
            for (uint64_t g = 0; g < n_ct; ++g) {
                for (uint32_t u = 0; u < c->k; ++u) put(flat_out + g * kn + size_t(u) * c->n, out[g]->GetElements()[u]);
                MNTRUPlaintext r = 0;
                c->sch->MNTRU()->Decrypt(c->mntru, c->skN, out[g], &r, 4);
                dec[g] = int(r);
            }
        }
        else {
            std::vector<MKLWECiphertext> in;
            for (uint64_t g = 0; g < n_ct; ++g) {
                auto x = c->sch->MKLWE()->Encrypt(c->mklwe, c->skL, bits[g], 4, NativeInteger(c->q));
                for (uint32_t u = 0; u < c->k; ++u) put(flat_in + g * kn + size_t(u) * c->n, x->GetA()[u]);
                b_in[g] = x->GetB().ConvertToInt<uint32_t>();
                in.push_back(x);
            }
            c->batch->SaveCiphertexts(path, in);
            auto out = c->batch->LoadCiphertextsLWE(path);
            if (out.size() != n_ct) throw std::runtime_error("ciphertext count changed on the way through the file");
            for (uint64_t g = 0; g < n_ct; ++g) {
                for (uint32_t u = 0; u < c->k; ++u) put(flat_out + g * kn + size_t(u) * c->n, out[g]->GetA()[u]);
                b_out[g] = out[g]->GetB().ConvertToInt<uint32_t>();
                MKLWEPlaintext r = 0;
                c->sch->MKLWE()->Decrypt(c->mklwe, c->skL, out[g], &r, 4);
                dec[g] = int(r);
            }
        }
    });
}

// EvalBinGateBatch(NAND) on B gates given as flat arrays (wrapped into reference ciphertext objects first); the outputs are
// exported flat and, when `dec` is not null, decrypted with the reference's Decrypt.
int ref_batch_b200_gates(void* h, const uint32_t* ct1, const uint32_t* b1, const uint32_t* ct2, const uint32_t* b2, uint32_t* out,
                         uint32_t* bout, uint64_t B, int* dec) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        if (!c->batch) throw std::runtime_error("ref_batch_b200_load has not been called");
        const size_t kn = size_t(c->k) * c->n;
        if (c->scheme == 0) {
            std::vector<MNTRUCiphertext> x1, x2;
            for (uint64_t g = 0; g < B; ++g) {
                x1.push_back(mk_mntru_ct(c, ct1 + g * kn, c->q));
                x2.push_back(mk_mntru_ct(c, ct2 + g * kn, c->q));
            }
            auto r = c->batch->EvalBinGateBatch(NAND, x1, x2);
            for (uint64_t g = 0; g < B; ++g) {
                for (uint32_t u = 0; u < c->k; ++u) put(out + g * kn + size_t(u) * c->n, r[g]->GetElements()[u]);
                if (dec) {
                    MNTRUPlaintext m = 0;
                    ConstMNTRUCiphertext x = r[g];
                    c->sch->MNTRU()->Decrypt(c->mntru, c->skN, x, &m, 4);
                    dec[g] = int(m);
                }
            }
        }
        else {
            std::vector<MKLWECiphertext> x1, x2;
            for (uint64_t g = 0; g < B; ++g) {
                x1.push_back(mk_mklwe_ct(c, ct1 + g * kn, b1[g], c->q));
                x2.push_back(mk_mklwe_ct(c, ct2 + g * kn, b2[g], c->q));
            }
            auto r = c->batch->EvalBinGateBatch(NAND, x1, x2);
            for (uint64_t g = 0; g < B; ++g) {
                for (uint32_t u = 0; u < c->k; ++u) put(out + g * kn + size_t(u) * c->n, r[g]->GetA()[u]);
                bout[g] = r[g]->GetB().ConvertToInt<uint32_t>();
                if (dec) {
                    MKLWEPlaintext m = 0;
                    ConstMKLWECiphertext x = r[g];
                    c->sch->MKLWE()->Decrypt(c->mklwe, c->skL, x, &m, 4);
                    dec[g] = int(m);
                }
            }
        }
    });
}

// BinFHEBatchB200::EvalCircuit on reference ciphertext objects: wires [W][k][n] (+ b [W]) flat in/out, `is_input[w]` marks the
// primary inputs (the other entries start as null pointers), gates as three index arrays.  dec (optional) receives the
// reference's Decrypt of every wire that is set after the call.
int ref_batch_b200_circuit(void* h, uint32_t* wires, uint32_t* wb, const uint8_t* is_input, uint64_t W, const uint32_t* in1,
                           const uint32_t* in2, const uint32_t* out, uint64_t G, uint32_t* levels, int* dec) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        if (!c->batch) throw std::runtime_error("ref_batch_b200_load has not been called");
        const size_t kn = size_t(c->k) * c->n;
        std::vector<BinFHEBatchB200::Gate> gates;
        for (uint64_t g = 0; g < G; ++g) gates.push_back({in1[g], in2[g], out[g]});
        if (c->scheme == 0) {
            std::vector<MNTRUCiphertext> w(W);
            for (uint64_t i = 0; i < W; ++i)
                if (is_input[i]) w[i] = mk_mntru_ct(c, wires + i * kn, c->q);
            *levels = c->batch->EvalCircuit(w, gates);
            for (uint64_t i = 0; i < W; ++i) {
                if (!w[i]) continue;
                for (uint32_t u = 0; u < c->k; ++u) put(wires + i * kn + size_t(u) * c->n, w[i]->GetElements()[u]);
                if (dec) {
                    MNTRUPlaintext m = 0;
                    ConstMNTRUCiphertext x = w[i];
                    c->sch->MNTRU()->Decrypt(c->mntru, c->skN, x, &m, 4);
                    dec[i] = int(m);
                }
            }
        }
        else {
            std::vector<MKLWECiphertext> w(W);
            for (uint64_t i = 0; i < W; ++i)
                if (is_input[i]) w[i] = mk_mklwe_ct(c, wires + i * kn, wb[i], c->q);
            *levels = c->batch->EvalCircuit(w, gates);
            for (uint64_t i = 0; i < W; ++i) {
                if (!w[i]) continue;
                for (uint32_t u = 0; u < c->k; ++u) put(wires + i * kn + size_t(u) * c->n, w[i]->GetA()[u]);
                wb[i] = w[i]->GetB().ConvertToInt<uint32_t>();
                if (dec) {
                    MKLWEPlaintext m = 0;
                    ConstMKLWECiphertext x = w[i];
                    c->sch->MKLWE()->Decrypt(c->mklwe, c->skL, x, &m, 4);
                    dec[i] = int(m);
                }
            }
        }
    });
}

// Argument checks of EvalBinGateBatch: returns a bit mask of the cases that threw the reference's config_error
// (1: gate other than NAND, 2: the same ciphertext object on both sides, 4: input vectors of different length).
int ref_batch_b200_errors(void* h, int* mask) {
    auto* c = static_cast<RefCtx*>(h);
    return guarded(c, [&] {
        if (!c->batch) throw std::runtime_error("ref_batch_b200_load has not been called");
        *mask = 0;
        std::vector<uint32_t> z(size_t(c->k) * c->n, 1u);
        auto probe = [&](int bit, auto&& f) {
            try { f(); }
            catch (const config_error&) { *mask |= bit; }
        };
        if (c->scheme == 0) {
            MNTRUCiphertext a = mk_mntru_ct(c, z.data(), c->q), b = mk_mntru_ct(c, z.data(), c->q);
            probe(1, [&] { c->batch->EvalBinGateBatch(AND, std::vector<MNTRUCiphertext>{a}, std::vector<MNTRUCiphertext>{b}); });
            probe(2, [&] { c->batch->EvalBinGateBatch(NAND, std::vector<MNTRUCiphertext>{a}, std::vector<MNTRUCiphertext>{a}); });
            probe(4, [&] { c->batch->EvalBinGateBatch(NAND, std::vector<MNTRUCiphertext>{a, b}, std::vector<MNTRUCiphertext>{b}); });
            // a circuit whose primary input (wire 1: read by the gate, written by none) holds no ciphertext
            probe(8, [&] {
                std::vector<MNTRUCiphertext> w{a, nullptr, nullptr};
                c->batch->EvalCircuit(w, std::vector<BinFHEBatchB200::Gate>{{0, 1, 2}});
            });
        }
        else {
            MKLWECiphertext a = mk_mklwe_ct(c, z.data(), 1, c->q), b = mk_mklwe_ct(c, z.data(), 1, c->q);
            probe(1, [&] { c->batch->EvalBinGateBatch(AND, std::vector<MKLWECiphertext>{a}, std::vector<MKLWECiphertext>{b}); });
            probe(2, [&] { c->batch->EvalBinGateBatch(NAND, std::vector<MKLWECiphertext>{a}, std::vector<MKLWECiphertext>{a}); });
            probe(4, [&] { c->batch->EvalBinGateBatch(NAND, std::vector<MKLWECiphertext>{a, b}, std::vector<MKLWECiphertext>{b}); });
            probe(8, [&] {
                std::vector<MKLWECiphertext> w{a, nullptr, nullptr};
                c->batch->EvalCircuit(w, std::vector<BinFHEBatchB200::Gate>{{0, 1, 2}});
            });
        }
    });
}

}  // extern "C"

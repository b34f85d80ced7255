// BinFHEBatchB200 -- the batched EvalBinGate a maintainer of the reference adds next to BinFHEContext::EvalBinGate
// (src/binfhe/include/binfhecontext.h:299-338, src/binfhe/lib/binfhecontext.cpp:415-419).
//
// The reference evaluates one gate per call: BinFHEScheme::EvalBinGate (binfhe-base-scheme.cpp:380-463 for MK-LWE
// ciphertexts, :467-515 for matrix-NTRU ciphertexts) runs input combine, BootstrapGateCore, extraction, ModSwitch and
// KeySwitch / KeySwitch2 on the host.  This class keeps the reference's types on both sides -- BinFHECryptoParams and
// UniEncBTKey in, vectors of MKLWECiphertext / MNTRUCiphertext in and out, OPENFHE_THROW for errors with the reference's
// messages -- and runs the whole body for a batch of gates in the GPU library through the C ABI of include/mkfhe_b200.h
// (mkfhe_eval_nand).  Header-only; depends on reference headers + <dlfcn.h>; contains no arithmetic.
//
//   BinFHEBatchB200 gpu("libmkfhe_b200.so");
//   gpu.LoadKeys(params, method, EK);            // once, after MKKeyGen (binfhe-base-scheme.cpp:198-338)
//   gpu.LoadGateCiphertext(ctNAND);              // MKNTRU only: the ctGateGen output (binfhe-base-scheme.cpp:340-376)
//   auto out = gpu.EvalBinGateBatch(NAND, ct1s, ct2s);
//   gpu.EvalCircuit(wires, gates);               // a NAND netlist on a table of ciphertexts, one batch per level
//
// Compiled in this repository only as part of oracle/_ref (the reference built in place): ref_batch_b200_* in
// oracle/ref_harness.cpp drive it, tests/test_gpu_fullsize.py compares its output objects with the reference's EvalBinGate.
#ifndef MKFHE_B200_BINFHE_BATCH_B200_H
#define MKFHE_B200_BINFHE_BATCH_B200_H

#include <dlfcn.h>

#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "binfhe-base-scheme.h"

#include "mkfhe_b200.h"

namespace lbcrypto {

class BinFHEBatchB200 {
public:
    explicit BinFHEBatchB200(const std::string& libpath, int device = -1) : m_device(device) {
        m_lib = dlopen(libpath.c_str(), RTLD_NOW | RTLD_LOCAL);
        if (!m_lib)
            OPENFHE_THROW(config_error, std::string("cannot load libmkfhe_b200: ") + dlerror());
        bind(m_create, "mkfhe_ctx_create");
        bind(m_destroy, "mkfhe_ctx_destroy");
        bind(m_error, "mkfhe_last_error");
        bind(m_load_bk, "mkfhe_load_bk");
        bind(m_load_pkey, "mkfhe_load_pkey");
        bind(m_load_ksk_mntru, "mkfhe_load_ksk_mntru");
        bind(m_load_ksk_mklwe, "mkfhe_load_ksk_mklwe");
        bind(m_load_ctnand, "mkfhe_load_ctnand");
        bind(m_eval_nand, "mkfhe_eval_nand");
        bind(m_eval_circuit, "mkfhe_eval_circuit");
    }
    BinFHEBatchB200(const BinFHEBatchB200&)            = delete;
    BinFHEBatchB200& operator=(const BinFHEBatchB200&) = delete;
    ~BinFHEBatchB200() {
        if (m_ctx)
            m_destroy(m_ctx);
        if (m_lib)
            dlclose(m_lib);
    }

    // Flatten and upload everything EvalBinGate reads from EK: BSkey, Pkey and the key-switching key of the additive scheme.
    // method: MKNTRU (ternary secrets, UniEncAccumulatorXZW, matrix-NTRU ciphertexts) or MKNTRU_LWE (binary secrets,
    // UniEncAccumulatorXZW_B, MK-LWE ciphertexts) -- the two rows of binfhe-base-scheme.h:144-147 that have a gate path.
    void LoadKeys(const std::shared_ptr<BinFHECryptoParams>& params, BINFHE_METHOD method, const UniEncBTKey& EK) {
        static_assert(sizeof(NativeInteger) == sizeof(uint32_t), "build the reference with NATIVE_SIZE=32 (README.md:19)");
        if (method != MKNTRU && method != MKNTRU_LWE)
            OPENFHE_THROW(config_error, "BinFHEBatchB200 serves the MKNTRU and MKNTRU_LWE methods");
        const auto& up = params->GetUniEncParams();
        m_lwe          = method == MKNTRU_LWE;
        mkfhe_params p{};
        if (m_lwe) {
            const auto& ap = params->GetMKLWEParams();
            p.n = ap->Getn(); p.q = u32(ap->Getq()); p.qKS = u32(ap->GetqKS()); p.baseKS = ap->GetBaseKS();
        }
        else {
            const auto& ap = params->GetMatrixNTRUParams();
            p.n = ap->Getn(); p.q = u32(ap->Getq()); p.qKS = u32(ap->GetqKS()); p.baseKS = ap->GetBaseKS();
        }
        p.k          = up->Getk();
        p.N          = up->GetN();
        p.Q          = u32(up->GetQ());
        p.psi        = u32(up->GetPolyParams()->GetRootOfUnity());
        p.baseG_log2 = static_cast<uint32_t>(__builtin_ctz(up->GetBaseG()));
        p.digitsG    = up->GetDigitsG();
        p.dKS        = static_cast<uint32_t>(std::ceil(std::log(double(p.qKS)) / std::log(double(p.baseKS))));  // mntru-pke.cpp:630, mklwe-pke.cpp:182
        p.variant    = m_lwe ? MKFHE_VARIANT_XZW_B : MKFHE_VARIANT_XZW;
        p.scheme     = m_lwe ? MKFHE_SCHEME_MKLWE : MKFHE_SCHEME_MNTRU;
        p.device     = m_device;
        if (m_ctx) {
            m_destroy(m_ctx);
            m_ctx = nullptr;
        }
        check(m_create(&p, &m_ctx), "mkfhe_ctx_create");
        m_k = p.k; m_n = p.n; m_q = p.q;

        const uint32_t k = p.k, N = p.N, n = p.n, ell = p.digitsG - 1, S = m_lwe ? 1 : 2;
        {  // UniEncACCKeyImpl: [u][s][i] evks of [j][d|f][N], then evk* (mk-acc-xzw.cpp:51,72-74)
            std::vector<uint32_t> flat((size_t(k) * S * n + 1) * ell * 2 * N);
            size_t off = 0;
            auto dump  = [&](const UniEncEvalKey& e) {
                for (uint32_t j = 0; j < ell; ++j)
                    for (uint32_t c = 0; c < 2; ++c, off += N) copy(&flat[off], (*e)[j][c].GetValues(), N);
            };
            for (uint32_t u = 0; u < k; ++u)
                for (uint32_t s = 0; s < S; ++s)
                    for (uint32_t i = 0; i < n; ++i) dump((*EK.BSkey)[u][s][i]);
            dump((*EK.BSkey)[0][0][n]);
            check(m_load_bk(m_ctx, flat.data(), flat.size()), "mkfhe_load_bk");
        }
        {
            std::vector<uint32_t> pk(size_t(k) * ell * N);
            for (uint32_t u = 0; u < k; ++u)
                for (uint32_t j = 0; j < ell; ++j) copy(&pk[(size_t(u) * ell + j) * N], EK.Pkey[u][j].GetValues(), N);
            check(m_load_pkey(m_ctx, pk.data(), pk.size()), "mkfhe_load_pkey");
        }
        if (m_lwe) {  // MKLWESwitchingKeyImpl: A[u][i][a0][j] vectors of n, B[u][i][a0][j] (mklwe-pke.cpp:176-256)
            const auto& KA = EK.LKSkey->GetElementsA();
            const auto& KB = EK.LKSkey->GetElementsB();
            const size_t rows = size_t(k) * N * p.baseKS * p.dKS;
            std::vector<uint32_t> A(rows * n), B(rows);
            size_t r = 0;
            for (uint32_t u = 0; u < k; ++u)
                for (uint32_t i = 0; i < N; ++i)
                    for (uint32_t a0 = 0; a0 < p.baseKS; ++a0)
                        for (uint32_t j = 0; j < p.dKS; ++j, ++r) {
                            copy(&A[r * n], KA[u][i][a0][j], n);
                            B[r] = u32(KB[u][i][a0][j]);
                        }
            check(m_load_ksk_mklwe(m_ctx, A.data(), A.size(), B.data(), B.size()), "mkfhe_load_ksk_mklwe");
        }
        else {  // MNTRUSwitchingKey2Impl: the digit-1 slice (the others are its multiples, mntru-pke.cpp:743-752)
            const auto& K  = EK.KSkey2->GetElements();
            const size_t L = size_t(N) * p.dKS;
            std::vector<uint32_t> flat(size_t(k) * L * n);
            for (uint32_t u = 0; u < k; ++u)
                for (size_t l = 0; l < L; ++l) copy(&flat[(size_t(u) * L + l) * n], K[u][1][l], n);
            check(m_load_ksk_mntru(m_ctx, flat.data(), flat.size()), "mkfhe_load_ksk_mntru");
        }
    }

    // MKNTRU: the encryption of 5q/8 that BinFHEContext keeps as m_ctNAND (binfhecontext.h:88, binfhecontext.cpp:573-575)
    void LoadGateCiphertext(ConstMNTRUCiphertext ctNAND) {
        need_ctx();
        std::vector<uint32_t> flat(size_t(m_k) * m_n);
        for (uint32_t u = 0; u < m_k; ++u) copy(&flat[size_t(u) * m_n], ctNAND->GetElements()[u], m_n);
        check(m_load_ctnand(m_ctx, flat.data(), flat.size()), "mkfhe_load_ctnand");
        m_have_nand = true;
    }

    // MK-LWE ciphertexts (method MKNTRU_LWE): element g of the result is EvalBinGate(params, gate, EK, ct1[g], ct2[g])
    // of binfhe-base-scheme.cpp:380-463, bit for bit.
    std::vector<MKLWECiphertext> EvalBinGateBatch(BINGATE gate, const std::vector<MKLWECiphertext>& ct1,
                                                  const std::vector<MKLWECiphertext>& ct2) const {
        need_ctx();
        if (!m_lwe)
            OPENFHE_THROW(config_error, "keys were loaded for matrix-NTRU ciphertexts");
        const size_t B = precheck(gate, ct1, ct2), kn = size_t(m_k) * m_n;
        std::vector<uint32_t> a1(B * kn), a2(B * kn), out(B * kn), b1(B), b2(B), bo(B);
        for (size_t g = 0; g < B; ++g) {
            shape(ct1[g]->GetA());
            shape(ct2[g]->GetA());
            for (uint32_t u = 0; u < m_k; ++u) {
                copy(&a1[g * kn + size_t(u) * m_n], ct1[g]->GetA()[u], m_n);
                copy(&a2[g * kn + size_t(u) * m_n], ct2[g]->GetA()[u], m_n);
            }
            b1[g] = u32(ct1[g]->GetB());
            b2[g] = u32(ct2[g]->GetB());
        }
        check(m_eval_nand(m_ctx, a1.data(), b1.data(), a2.data(), b2.data(), out.data(), bo.data(), B), "mkfhe_eval_nand");
        std::vector<MKLWECiphertext> res;
        res.reserve(B);
        for (size_t g = 0; g < B; ++g)
            res.push_back(std::make_shared<MKLWECiphertextImpl>(rows(&out[g * kn]), NativeInteger(bo[g])));
        return res;
    }

    // matrix-NTRU ciphertexts (method MKNTRU): EvalBinGate(params, gate, EK, ct1[g], ct2[g], ctNAND) of :467-515
    std::vector<MNTRUCiphertext> EvalBinGateBatch(BINGATE gate, const std::vector<MNTRUCiphertext>& ct1,
                                                  const std::vector<MNTRUCiphertext>& ct2) const {
        need_ctx();
        if (m_lwe)
            OPENFHE_THROW(config_error, "keys were loaded for MK-LWE ciphertexts");
        if (!m_have_nand)
            OPENFHE_THROW(config_error, "LoadGateCiphertext has not been called");
        const size_t B = precheck(gate, ct1, ct2), kn = size_t(m_k) * m_n;
        std::vector<uint32_t> a1(B * kn), a2(B * kn), out(B * kn);
        for (size_t g = 0; g < B; ++g) {
            shape(ct1[g]->GetElements());
            shape(ct2[g]->GetElements());
            for (uint32_t u = 0; u < m_k; ++u) {
                copy(&a1[g * kn + size_t(u) * m_n], ct1[g]->GetElements()[u], m_n);
                copy(&a2[g * kn + size_t(u) * m_n], ct2[g]->GetElements()[u], m_n);
            }
        }
        check(m_eval_nand(m_ctx, a1.data(), nullptr, a2.data(), nullptr, out.data(), nullptr, B), "mkfhe_eval_nand");
        std::vector<MNTRUCiphertext> res;
        res.reserve(B);
        for (size_t g = 0; g < B; ++g) res.push_back(std::make_shared<MNTRUCiphertextImpl>(rows(&out[g * kn])));
        return res;
    }

    // A circuit of NAND gates: what a caller of the reference writes as a loop of EvalBinGate calls
    // (binfhecontext.cpp:415-419).  `wires` holds the primary inputs (any entry a gate writes may be null on entry); gate g
    // sets wires[gates[g][2]] = NAND(wires[gates[g][0]], wires[gates[g][1]]).  Gates in topological order; the library runs
    // each level of the netlist as one batch.  Returns the number of levels.
    struct Gate { uint32_t in1, in2, out; };
    uint32_t EvalCircuit(std::vector<MKLWECiphertext>& wires, const std::vector<Gate>& gates) const {
        need_ctx();
        if (!m_lwe)
            OPENFHE_THROW(config_error, "keys were loaded for matrix-NTRU ciphertexts");
        const size_t W = wires.size(), kn = size_t(m_k) * m_n;
        std::vector<uint32_t> flat(W * kn, 0), b(W, 0), i1, i2, o;
        split(gates, W, i1, i2, o);
        for (size_t w = 0; w < W; ++w) {
            if (!wires[w])
                continue;
            shape(wires[w]->GetA());
            for (uint32_t u = 0; u < m_k; ++u) copy(&flat[w * kn + size_t(u) * m_n], wires[w]->GetA()[u], m_n);
            b[w] = u32(wires[w]->GetB());
        }
        uint32_t levels = 0;
        check(m_eval_circuit(m_ctx, flat.data(), b.data(), W, i1.data(), i2.data(), o.data(), gates.size(), &levels), "mkfhe_eval_circuit");
        for (uint32_t w : o) wires[w] = std::make_shared<MKLWECiphertextImpl>(rows(&flat[size_t(w) * kn]), NativeInteger(b[w]));
        return levels;
    }
    uint32_t EvalCircuit(std::vector<MNTRUCiphertext>& wires, const std::vector<Gate>& gates) const {
        need_ctx();
        if (m_lwe)
            OPENFHE_THROW(config_error, "keys were loaded for MK-LWE ciphertexts");
        if (!m_have_nand)
            OPENFHE_THROW(config_error, "LoadGateCiphertext has not been called");
        const size_t W = wires.size(), kn = size_t(m_k) * m_n;
        std::vector<uint32_t> flat(W * kn, 0), i1, i2, o;
        split(gates, W, i1, i2, o);
        for (size_t w = 0; w < W; ++w) {
            if (!wires[w])
                continue;
            shape(wires[w]->GetElements());
            for (uint32_t u = 0; u < m_k; ++u) copy(&flat[w * kn + size_t(u) * m_n], wires[w]->GetElements()[u], m_n);
        }
        uint32_t levels = 0;
        check(m_eval_circuit(m_ctx, flat.data(), nullptr, W, i1.data(), i2.data(), o.data(), gates.size(), &levels), "mkfhe_eval_circuit");
        for (uint32_t w : o) wires[w] = std::make_shared<MNTRUCiphertextImpl>(rows(&flat[size_t(w) * kn]));
        return levels;
    }

private:
    // index arrays of a netlist; the reference's independence check (binfhe-base-scheme.cpp:382-383) and the range check that
    // must precede touching `wires` happen here, the remaining structure checks in the library
    static void split(const std::vector<Gate>& gates, size_t W, std::vector<uint32_t>& i1, std::vector<uint32_t>& i2,
                      std::vector<uint32_t>& o) {
        for (const Gate& g : gates) {
            if (g.in1 == g.in2)
                OPENFHE_THROW(config_error, "Input ciphertexts should be independant");
            if (g.in1 >= W || g.in2 >= W || g.out >= W)
                OPENFHE_THROW(config_error, "EvalCircuit: wire index out of range");
            i1.push_back(g.in1);
            i2.push_back(g.in2);
            o.push_back(g.out);
        }
    }
    static uint32_t u32(const NativeInteger& x) { return x.ConvertToInt<uint32_t>(); }
    static void copy(uint32_t* dst, const NativeVector& v, uint32_t len) { std::memcpy(dst, &v[0], sizeof(uint32_t) * len); }
    std::vector<NativeVector> rows(const uint32_t* src) const {
        std::vector<NativeVector> r;
        r.reserve(m_k);
        for (uint32_t u = 0; u < m_k; ++u) {
            NativeVector v(m_n, NativeInteger(m_q));
            std::memcpy(&v[0], src + size_t(u) * m_n, sizeof(uint32_t) * m_n);
            r.push_back(std::move(v));
        }
        return r;
    }
    // the reference's own argument checks, with its messages (binfhe-base-scheme.cpp:342 and :383 / :472)
    template <class CT>
    size_t precheck(BINGATE gate, const std::vector<CT>& ct1, const std::vector<CT>& ct2) const {
        if (gate != NAND)
            OPENFHE_THROW(config_error, "Support NAND gate Only");
        if (ct1.size() != ct2.size())
            OPENFHE_THROW(config_error, "EvalBinGateBatch: the two input vectors differ in length");
        for (size_t g = 0; g < ct1.size(); ++g)
            if (ct1[g] == ct2[g])
                OPENFHE_THROW(config_error, "Input ciphertexts should be independant");
        return ct1.size();
    }
    void shape(const std::vector<NativeVector>& a) const {
        if (a.size() != m_k || a[0].GetLength() != m_n || u32(a[0].GetModulus()) != m_q)
            OPENFHE_THROW(config_error, "EvalBinGateBatch: ciphertext shape or modulus differs from the loaded keys");
    }
    void need_ctx() const {
        if (!m_ctx)
            OPENFHE_THROW(config_error, "BinFHEBatchB200: LoadKeys has not been called");
    }
    template <class F>
    void bind(F& f, const char* name) {
        f = reinterpret_cast<F>(dlsym(m_lib, name));
        if (!f)
            OPENFHE_THROW(config_error, std::string("libmkfhe_b200 lacks symbol ") + name);
    }
    void check(int rc, const char* what) const {
        if (rc != 0)
            OPENFHE_THROW(openfhe_error, std::string(what) + " failed: " + (m_error ? m_error(m_ctx) : "?"));
    }

    int m_device;
    void* m_lib{nullptr};
    mkfhe_ctx* m_ctx{nullptr};
    bool m_lwe{false}, m_have_nand{false};
    uint32_t m_k{0}, m_n{0}, m_q{0};
    decltype(&mkfhe_ctx_create) m_create{nullptr};
    decltype(&mkfhe_ctx_destroy) m_destroy{nullptr};
    decltype(&mkfhe_last_error) m_error{nullptr};
    decltype(&mkfhe_load_bk) m_load_bk{nullptr};
    decltype(&mkfhe_load_pkey) m_load_pkey{nullptr};
    decltype(&mkfhe_load_ksk_mntru) m_load_ksk_mntru{nullptr};
    decltype(&mkfhe_load_ksk_mklwe) m_load_ksk_mklwe{nullptr};
    decltype(&mkfhe_load_ctnand) m_load_ctnand{nullptr};
    decltype(&mkfhe_eval_nand) m_eval_nand{nullptr};
    decltype(&mkfhe_eval_circuit) m_eval_circuit{nullptr};
};

}  // namespace lbcrypto
#endif

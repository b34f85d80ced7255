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
//   gpu.SaveKeyFile(path, params, method, EK, ctNAND);   gpu2.LoadKeyFile(path);   // keys: generate once, reuse everywhere
//   gpu.SaveCiphertexts(path, cts);   auto cts2 = gpu.LoadCiphertextsLWE(path);     // ciphertext tables on disk
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
        bind(m_keyfile_write, "mkfhe_keyfile_write");
        bind(m_keyfile_info, "mkfhe_keyfile_info");
        bind(m_keyfile_load, "mkfhe_keyfile_load");
        bind(m_ctfile_write, "mkfhe_ctfile_write");
        bind(m_ctfile_info, "mkfhe_ctfile_info");
        bind(m_ctfile_read, "mkfhe_ctfile_read");
    }
    BinFHEBatchB200(const BinFHEBatchB200&)            = delete;
    BinFHEBatchB200& operator=(const BinFHEBatchB200&) = delete;
    ~BinFHEBatchB200() {
        if (m_ctx)
            m_destroy(m_ctx);
        if (m_lib)
            dlclose(m_lib);
    }

    // Everything EvalBinGate reads from EK, flattened to the layouts of include/mkfhe_b200.h: BSkey, Pkey and the
    // key-switching key of the additive scheme.  method: MKNTRU (ternary secrets, UniEncAccumulatorXZW, matrix-NTRU
    // ciphertexts) or MKNTRU_LWE (binary secrets, UniEncAccumulatorXZW_B, MK-LWE ciphertexts) -- the two rows of
    // binfhe-base-scheme.h:144-147 that have a gate path.
    struct FlatKeys {
        mkfhe_params p{};
        std::vector<uint32_t> bk, pkey, kskA, kskB;  // kskA: MKLWE A vectors or the MNTRU digit-1 slice; kskB: MKLWE only
    };
    static FlatKeys Flatten(const std::shared_ptr<BinFHECryptoParams>& params, BINFHE_METHOD method, const UniEncBTKey& EK,
                            int device = -1) {
        static_assert(sizeof(NativeInteger) == sizeof(uint32_t), "build the reference with NATIVE_SIZE=32 (README.md:19)");
        if (method != MKNTRU && method != MKNTRU_LWE)
            OPENFHE_THROW(config_error, "BinFHEBatchB200 serves the MKNTRU and MKNTRU_LWE methods");
        const auto& up = params->GetUniEncParams();
        const bool lwe = method == MKNTRU_LWE;
        FlatKeys K;
        mkfhe_params& p = K.p;
        if (lwe) {
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
        p.variant    = lwe ? MKFHE_VARIANT_XZW_B : MKFHE_VARIANT_XZW;
        p.scheme     = lwe ? MKFHE_SCHEME_MKLWE : MKFHE_SCHEME_MNTRU;
        p.device     = device;

        const uint32_t k = p.k, N = p.N, n = p.n, ell = p.digitsG - 1, S = lwe ? 1 : 2;
        {  // UniEncACCKeyImpl: [u][s][i] evks of [j][d|f][N], then evk* (mk-acc-xzw.cpp:51,72-74)
            K.bk.resize((size_t(k) * S * n + 1) * ell * 2 * N);
            size_t off = 0;
            auto dump  = [&](const UniEncEvalKey& e) {
                for (uint32_t j = 0; j < ell; ++j)
                    for (uint32_t c = 0; c < 2; ++c, off += N) copy(&K.bk[off], (*e)[j][c].GetValues(), N);
            };
            for (uint32_t u = 0; u < k; ++u)
                for (uint32_t s = 0; s < S; ++s)
                    for (uint32_t i = 0; i < n; ++i) dump((*EK.BSkey)[u][s][i]);
            dump((*EK.BSkey)[0][0][n]);
        }
        K.pkey.resize(size_t(k) * ell * N);
        for (uint32_t u = 0; u < k; ++u)
            for (uint32_t j = 0; j < ell; ++j) copy(&K.pkey[(size_t(u) * ell + j) * N], EK.Pkey[u][j].GetValues(), N);
        if (lwe) {  // MKLWESwitchingKeyImpl: A[u][i][a0][j] vectors of n, B[u][i][a0][j] (mklwe-pke.cpp:176-256)
            const auto& KA = EK.LKSkey->GetElementsA();
            const auto& KB = EK.LKSkey->GetElementsB();
            const size_t rows = size_t(k) * N * p.baseKS * p.dKS;
            K.kskA.resize(rows * n);
            K.kskB.resize(rows);
            size_t r = 0;
            for (uint32_t u = 0; u < k; ++u)
                for (uint32_t i = 0; i < N; ++i)
                    for (uint32_t a0 = 0; a0 < p.baseKS; ++a0)
                        for (uint32_t j = 0; j < p.dKS; ++j, ++r) {
                            copy(&K.kskA[r * n], KA[u][i][a0][j], n);
                            K.kskB[r] = u32(KB[u][i][a0][j]);
                        }
        }
        else {  // MNTRUSwitchingKey2Impl: the digit-1 slice (the others are its multiples, mntru-pke.cpp:743-752)
            const auto& KS = EK.KSkey2->GetElements();
            const size_t L = size_t(N) * p.dKS;
            K.kskA.resize(size_t(k) * L * n);
            for (uint32_t u = 0; u < k; ++u)
                for (size_t l = 0; l < L; ++l) copy(&K.kskA[(size_t(u) * L + l) * n], KS[u][1][l], n);
        }
        return K;
    }

    // Flatten and upload: once, after MKKeyGen (binfhe-base-scheme.cpp:198-338)
    void LoadKeys(const std::shared_ptr<BinFHECryptoParams>& params, BINFHE_METHOD method, const UniEncBTKey& EK) {
        const FlatKeys K = Flatten(params, method, EK, m_device);
        open_context(K.p);
        check(m_load_bk(m_ctx, K.bk.data(), K.bk.size()), "mkfhe_load_bk");
        check(m_load_pkey(m_ctx, K.pkey.data(), K.pkey.size()), "mkfhe_load_pkey");
        if (m_lwe)
            check(m_load_ksk_mklwe(m_ctx, K.kskA.data(), K.kskA.size(), K.kskB.data(), K.kskB.size()), "mkfhe_load_ksk_mklwe");
        else
            check(m_load_ksk_mntru(m_ctx, K.kskA.data(), K.kskA.size()), "mkfhe_load_ksk_mntru");
    }

    // ---- key file: generate once with the reference, reuse across runs and GPUs ----------------------
    // The reference saves keys through cereal, object by object (mk-acckey.h:128-137, mk-evalkey.h:107-116); UniEncBTKey has
    // no serializer of its own.  SaveKeyFile writes what LoadKeys would upload (plus ctNAND for MKNTRU) as one checksummed
    // file in the layouts of include/mkfhe_b200.h -- no GPU involved; LoadKeyFile creates the GPU context from the file's own
    // parameter header and reads the keys straight into HBM, so a later process needs neither the key objects nor the
    // BinFHECryptoParams they were generated with.
    void SaveKeyFile(const std::string& path, const std::shared_ptr<BinFHECryptoParams>& params, BINFHE_METHOD method,
                     const UniEncBTKey& EK, ConstMNTRUCiphertext ctNAND = nullptr) const {
        const FlatKeys K = Flatten(params, method, EK);
        std::vector<uint32_t> nand;
        if (method == MKNTRU) {
            if (!ctNAND)
                OPENFHE_THROW(config_error, "SaveKeyFile: the MKNTRU gate path needs ctNAND (ctGateGen)");
            nand.resize(size_t(K.p.k) * K.p.n);
            for (uint32_t u = 0; u < K.p.k; ++u) copy(&nand[size_t(u) * K.p.n], ctNAND->GetElements()[u], K.p.n);
        }
        if (m_keyfile_write(path.c_str(), &K.p, K.bk.data(), K.bk.size(), K.pkey.data(), K.pkey.size(), K.kskA.data(), K.kskA.size(),
                            K.kskB.empty() ? nullptr : K.kskB.data(), K.kskB.size(), nand.empty() ? nullptr : nand.data(), nand.size()) != 0)
            OPENFHE_THROW(openfhe_error, "mkfhe_keyfile_write failed for " + path);
    }
    void LoadKeyFile(const std::string& path) {
        mkfhe_params p{};
        uint64_t words[7] = {0};
        if (m_keyfile_info(path.c_str(), &p, words) != 0)
            OPENFHE_THROW(openfhe_error, path + ": unreadable, not a key file, or checksum mismatch");
        p.device = m_device;
        open_context(p);
        check(m_keyfile_load(m_ctx, path.c_str()), "mkfhe_keyfile_load");
        m_have_nand = words[6] != 0;
    }

    // ---- ciphertext tables on disk: the batch layout behind a checksummed, parameter-stamped header ------------------------
    // (the reference: one cereal archive per object, mklwe-ciphertext.h:161-172, mntru-ciphertext.h:116-125)
    void SaveCiphertexts(const std::string& path, const std::vector<MKLWECiphertext>& cts) const {
        need_ctx();
        if (!m_lwe)
            OPENFHE_THROW(config_error, "keys were loaded for matrix-NTRU ciphertexts");
        const size_t B = cts.size(), kn = size_t(m_k) * m_n;
        std::vector<uint32_t> a(B * kn), b(B);
        for (size_t g = 0; g < B; ++g) {
            shape(cts[g]->GetA());
            for (uint32_t u = 0; u < m_k; ++u) copy(&a[g * kn + size_t(u) * m_n], cts[g]->GetA()[u], m_n);
            b[g] = u32(cts[g]->GetB());
        }
        if (m_ctfile_write(path.c_str(), &m_params, a.data(), b.data(), B) != 0)
            OPENFHE_THROW(openfhe_error, "mkfhe_ctfile_write failed for " + path);
    }
    void SaveCiphertexts(const std::string& path, const std::vector<MNTRUCiphertext>& cts) const {
        need_ctx();
        if (m_lwe)
            OPENFHE_THROW(config_error, "keys were loaded for MK-LWE ciphertexts");
        const size_t B = cts.size(), kn = size_t(m_k) * m_n;
        std::vector<uint32_t> a(B * kn);
        for (size_t g = 0; g < B; ++g) {
            shape(cts[g]->GetElements());
            for (uint32_t u = 0; u < m_k; ++u) copy(&a[g * kn + size_t(u) * m_n], cts[g]->GetElements()[u], m_n);
        }
        if (m_ctfile_write(path.c_str(), &m_params, a.data(), nullptr, B) != 0)
            OPENFHE_THROW(openfhe_error, "mkfhe_ctfile_write failed for " + path);
    }
    std::vector<MKLWECiphertext> LoadCiphertextsLWE(const std::string& path) const {
        need_ctx();
        std::vector<uint32_t> a, b;
        const size_t B = read_table(path, true, a, b), kn = size_t(m_k) * m_n;
        std::vector<MKLWECiphertext> res;
        res.reserve(B);
        for (size_t g = 0; g < B; ++g) res.push_back(std::make_shared<MKLWECiphertextImpl>(rows(&a[g * kn]), NativeInteger(b[g])));
        return res;
    }
    std::vector<MNTRUCiphertext> LoadCiphertextsNTRU(const std::string& path) const {
        need_ctx();
        std::vector<uint32_t> a, b;
        const size_t B = read_table(path, false, a, b), kn = size_t(m_k) * m_n;
        std::vector<MNTRUCiphertext> res;
        res.reserve(B);
        for (size_t g = 0; g < B; ++g) res.push_back(std::make_shared<MNTRUCiphertextImpl>(rows(&a[g * kn])));
        return res;
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
        need_inputs(wires, i1, i2, o);
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
        need_inputs(wires, i1, i2, o);
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
    // a wire that some gate reads and no gate writes is a primary input: it must hold a ciphertext (a null entry would be
    // evaluated as the all-zero ciphertext)
    template <class CT>
    static void need_inputs(const std::vector<CT>& wires, const std::vector<uint32_t>& i1, const std::vector<uint32_t>& i2,
                            const std::vector<uint32_t>& o) {
        std::vector<uint8_t> written(wires.size(), 0);
        for (uint32_t w : o) written[w] = 1;
        for (size_t g = 0; g < i1.size(); ++g)
            for (uint32_t w : {i1[g], i2[g]})
                if (!written[w] && !wires[w])
                    OPENFHE_THROW(config_error, "EvalCircuit: wire " + std::to_string(w) + " is read by a gate, written by none, and holds no ciphertext");
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
    void open_context(const mkfhe_params& p) {
        if (m_ctx) {
            m_destroy(m_ctx);
            m_ctx = nullptr;
        }
        check(m_create(&p, &m_ctx), "mkfhe_ctx_create");
        m_params    = p;
        m_lwe       = p.scheme == MKFHE_SCHEME_MKLWE;
        m_have_nand = false;
        m_k = p.k; m_n = p.n; m_q = p.q;
    }
    size_t read_table(const std::string& path, bool lwe, std::vector<uint32_t>& a, std::vector<uint32_t>& b) const {
        if (lwe != m_lwe)
            OPENFHE_THROW(config_error, "the loaded keys are for the other ciphertext type");
        mkfhe_params fp{};
        uint64_t n = 0;
        if (m_ctfile_info(path.c_str(), &fp, &n) != 0)
            OPENFHE_THROW(openfhe_error, path + ": unreadable, not a ciphertext file, or checksum mismatch");
        a.resize(size_t(n) * m_k * m_n);
        b.resize(lwe ? size_t(n) : 0);
        if (fp.k != m_k || fp.n != m_n || m_ctfile_read(path.c_str(), &m_params, a.data(), lwe ? b.data() : nullptr, n, &n) != 0)
            OPENFHE_THROW(config_error, path + " was written for a different parameter set");
        return size_t(n);
    }
    void need_ctx() const {
        if (!m_ctx)
            OPENFHE_THROW(config_error, "BinFHEBatchB200: LoadKeys / LoadKeyFile has not been called");
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
    decltype(&mkfhe_keyfile_write) m_keyfile_write{nullptr};
    decltype(&mkfhe_keyfile_info) m_keyfile_info{nullptr};
    decltype(&mkfhe_keyfile_load) m_keyfile_load{nullptr};
    decltype(&mkfhe_ctfile_write) m_ctfile_write{nullptr};
    decltype(&mkfhe_ctfile_info) m_ctfile_info{nullptr};
    decltype(&mkfhe_ctfile_read) m_ctfile_read{nullptr};
    mkfhe_params m_params{};
};

}  // namespace lbcrypto
#endif

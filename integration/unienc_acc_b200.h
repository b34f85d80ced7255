// UniEncAccumulatorB200 -- the drop-in accumulator a maintainer of the reference adds to src/binfhe/include/.
//
// It implements the reference's own plug-in surface, `class UniEncAccumulator` (src/binfhe/include/mk-acc.h:54-80), and is
// selected exactly where the reference selects UniEncAccumulatorXZW / UniEncAccumulatorXZW_B
// (src/binfhe/include/binfhe-base-scheme.h:137-151).  KeyGenAcc delegates to the reference's CPU implementation (keys must
// be the reference's bytes); EvalAcc flattens its arguments and calls the GPU library through the C ABI of
// include/mkfhe_b200.h.  This header depends only on reference headers + <dlfcn.h>; it contains no arithmetic.
//
// Compiled in this repository only as part of oracle/_ref (the reference built in place), where it lets the tests run the
// reference's unmodified EvalBinGate orchestration on top of the GPU accumulator.
#ifndef MKFHE_B200_UNIENC_ACC_B200_H
#define MKFHE_B200_UNIENC_ACC_B200_H

#include <dlfcn.h>

#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "mk-acc-xzw.h"
#include "mk-acc-xzw_B.h"
#include "mk-acc.h"

#include "mkfhe_b200.h"

namespace lbcrypto {

class UniEncAccumulatorB200 final : public UniEncAccumulator {
public:
    enum Variant { XZW = MKFHE_VARIANT_XZW, XZW_B = MKFHE_VARIANT_XZW_B };

    UniEncAccumulatorB200(const std::string& libpath, Variant v) : m_variant(v) {
        m_lib = dlopen(libpath.c_str(), RTLD_NOW | RTLD_LOCAL);
        if (!m_lib)
            OPENFHE_THROW(config_error, std::string("cannot load libmkfhe_b200: ") + dlerror());
        bind(m_create, "mkfhe_ctx_create");
        bind(m_destroy, "mkfhe_ctx_destroy");
        bind(m_error, "mkfhe_last_error");
        bind(m_load_bk, "mkfhe_load_bk");
        bind(m_load_pkey, "mkfhe_load_pkey");
        bind(m_eval_acc, "mkfhe_eval_acc");
        if (v == XZW)
            m_cpu = std::make_shared<UniEncAccumulatorXZW>();
        else
            m_cpu = std::make_shared<UniEncAccumulatorXZW_B>();
    }
    ~UniEncAccumulatorB200() {  // the base has no virtual destructor; hold it by make_shared<UniEncAccumulatorB200>
        if (m_ctx)
            m_destroy(m_ctx);
        if (m_lib)
            dlclose(m_lib);
    }

    // Key generation stays on the host and is the reference's (mk-acc-xzw.cpp:38-87, mk-acc-xzw_B.cpp:38-101).
    UniEncACCKey KeyGenAcc(const std::shared_ptr<UniEncCryptoParams>& params, const std::vector<NativePoly>& invskNTT,
                           const ConstMNTRUPrivateKey& sk, const std::vector<NativePoly>& CRS) const override {
        return m_cpu->KeyGenAcc(params, invskNTT, sk, CRS);
    }
    UniEncACCKey KeyGenAcc(const std::shared_ptr<UniEncCryptoParams>& params, const std::vector<NativePoly>& invskNTT,
                           const ConstMKLWEPrivateKey& sk, const std::vector<NativePoly>& CRS) const override {
        return m_cpu->KeyGenAcc(params, invskNTT, sk, CRS);
    }

    // Upload BSkey + Pkey (they are constant across EvalAcc calls).  OPTIONAL: EvalAcc uploads lazily the first time it sees a
    // key object (see below), so the integration is the two make_shared lines of INTEGRATION.md section 2 and nothing else;
    // call this to move the one-time upload (flatten + H2D, ~0.1 s for 70 MiB) out of the first gate.  n and q are the
    // additive scheme's dimension and modulus (the key layout needs n; XZW scales rotations by 2N/q, mk-acc-xzw.cpp:110,
    // with the modulus EvalAcc reads from ct itself).
    void LoadKeys(const std::shared_ptr<UniEncCryptoParams>& params, ConstUniEncACCKey& ek,
                  const std::vector<std::vector<NativePoly>>& Pkey, uint32_t n, uint32_t q) const {
        static_assert(sizeof(NativeInteger) == sizeof(uint32_t), "build the reference with NATIVE_SIZE=32 (README.md:19)");
        const uint32_t k = params->Getk(), N = params->GetN(), ell = params->GetDigitsG() - 1;
        const uint32_t S = m_variant == XZW ? 2 : 1;
        mkfhe_params p{};
        p.k = k; p.N = N; p.n = n; p.q = q; p.qKS = q;
        p.Q          = params->GetQ().ConvertToInt<uint32_t>();
        p.psi        = params->GetPolyParams()->GetRootOfUnity().ConvertToInt<uint32_t>();
        p.baseG_log2 = static_cast<uint32_t>(__builtin_ctz(params->GetBaseG()));
        p.digitsG    = params->GetDigitsG();
        p.baseKS = 2; p.dKS = 1;  // key switching is not part of the accumulator plug-in
        p.variant = m_variant;
        p.scheme  = m_variant == XZW ? MKFHE_SCHEME_MNTRU : MKFHE_SCHEME_MKLWE;
        p.device  = -1;
        if (m_ctx) { m_destroy(m_ctx); m_ctx = nullptr; m_loaded = nullptr; }
        check(m_create(&p, &m_ctx), "mkfhe_ctx_create");

        std::vector<uint32_t> flat((size_t(k) * S * n + 1) * ell * 2 * N);
        size_t off = 0;
        auto dump  = [&](const UniEncEvalKey& e) {
            for (uint32_t j = 0; j < ell; ++j)
                for (uint32_t c = 0; c < 2; ++c) {
                    std::memcpy(&flat[off], &(*e)[j][c].GetValues()[0], sizeof(uint32_t) * N);
                    off += N;
                }
        };
        for (uint32_t u = 0; u < k; ++u)
            for (uint32_t s = 0; s < S; ++s)
                for (uint32_t i = 0; i < n; ++i) dump((*ek)[u][s][i]);
        dump((*ek)[0][0][n]);  // evk*, mk-acc-xzw.cpp:72-74
        check(m_load_bk(m_ctx, flat.data(), flat.size()), "mkfhe_load_bk");

        std::vector<uint32_t> pk(size_t(k) * ell * N);
        for (uint32_t u = 0; u < k; ++u)
            for (uint32_t j = 0; j < ell; ++j)
                std::memcpy(&pk[(size_t(u) * ell + j) * N], &Pkey[u][j].GetValues()[0], sizeof(uint32_t) * N);
        check(m_load_pkey(m_ctx, pk.data(), pk.size()), "mkfhe_load_pkey");
        m_k = k; m_N = N; m_n = n;
        m_loaded = ek.get();
    }

    // Same signature and semantics as the reference's EvalAcc; `skf` is unused by the reference as well (it is a debugging
    // leftover, mk-acc-xzw.cpp:89-130).
    void EvalAcc(const std::shared_ptr<UniEncCryptoParams>& params, ConstUniEncACCKey& ek,
                 std::vector<std::vector<NativePoly>> Pkey, std::vector<NativePoly> skf, MKACCCiphertext& acc,
                 const std::vector<NativeVector>& ct) const override {
        // The reference hands the keys to every EvalAcc call (mk-acc.h:67-71); they are uploaded the first time a key object is
        // seen (identity of the UniEncACCKeyImpl the shared_ptr owns) and stay resident for all later calls.
        if (ct.empty() || ct.size() != params->Getk())
            OPENFHE_THROW(config_error, "UniEncAccumulatorB200: one rotation vector per party expected");
        if (!m_ctx || m_loaded != ek.get())
            LoadKeys(params, ek, Pkey, ct[0].GetLength(), ct[0].GetModulus().ConvertToInt<uint32_t>());
        if (ct.size() != m_k || ct[0].GetLength() != m_n)
            OPENFHE_THROW(config_error, "UniEncAccumulatorB200: ciphertext shape differs from the loaded keys");
        std::vector<uint32_t> a(size_t(m_k) * m_N), r(size_t(m_k) * m_n);
        auto& el = acc->GetElements();
        for (uint32_t u = 0; u < m_k; ++u) {
            if (el[u].GetFormat() != Format::EVALUATION)
                OPENFHE_THROW(config_error, "UniEncAccumulatorB200: accumulator must be in EVALUATION format");
            std::memcpy(&a[size_t(u) * m_N], &el[u].GetValues()[0], sizeof(uint32_t) * m_N);
            std::memcpy(&r[size_t(u) * m_n], &ct[u][0], sizeof(uint32_t) * m_n);
        }
        const uint32_t mod = ct[0].GetModulus().ConvertToInt<uint32_t>();
        check(m_eval_acc(m_ctx, a.data(), r.data(), mod, a.data(), 1), "mkfhe_eval_acc");
        for (uint32_t u = 0; u < m_k; ++u) {
            NativeVector v(m_N, params->GetQ());
            std::memcpy(&v[0], &a[size_t(u) * m_N], sizeof(uint32_t) * m_N);
            el[u].SetValues(std::move(v), Format::EVALUATION);
        }
    }

private:
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

    Variant m_variant;
    void* m_lib{nullptr};
    mutable mkfhe_ctx* m_ctx{nullptr};           // EvalAcc is const in the interface; the resident copy of the keys is a cache
    mutable const void* m_loaded{nullptr};       // the UniEncACCKeyImpl whose keys are resident
    mutable uint32_t m_k{0}, m_N{0}, m_n{0};
    std::shared_ptr<UniEncAccumulator> m_cpu;
    decltype(&mkfhe_ctx_create) m_create{nullptr};
    decltype(&mkfhe_ctx_destroy) m_destroy{nullptr};
    decltype(&mkfhe_last_error) m_error{nullptr};
    decltype(&mkfhe_load_bk) m_load_bk{nullptr};
    decltype(&mkfhe_load_pkey) m_load_pkey{nullptr};
    decltype(&mkfhe_eval_acc) m_eval_acc{nullptr};
};

}  // namespace lbcrypto
#endif

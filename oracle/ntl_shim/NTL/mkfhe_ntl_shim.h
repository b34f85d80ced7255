// Minimal stand-in for the parts of NTL that the reference's KEY GENERATION uses.
//
// TEST INFRASTRUCTURE ONLY.  NTL/GMP headers are not installable in this image (no network), and the
// reference includes <NTL/ZZ_pX.h>, <NTL/mat_ZZ_p.h>, <NTL/mat_ZZ.h> from
//   src/binfhe/include/binfhe-base-scheme.h:30-32 and src/binfhe/include/mntru-pke.h:16-18.
// NTL symbols are used in exactly two key-generation helpers:
//   Get_invertible_NativeVector  (src/binfhe/lib/binfhe-base-scheme.cpp:105-195)  -> ZZ_pX, SetCoeff, InvMod, deg, conv<long>
//   Get_invertible_Matrix        (src/binfhe/lib/mntru-pke.cpp:19-106)            -> mat_ZZ_p, Vec<ZZ_p>, inv, conv<long>
// Both results are mathematically unique (ring inverse / matrix inverse mod a prime), so any correct
// implementation yields the same keys as NTL for the same sampled input.  The evaluation hot path
// (EvalBinGate / EvalAcc / ModSwitch / KeySwitch) never touches NTL, so this shim is not part of the
// parity surface.
#ifndef MKFHE_B200_NTL_SHIM_H
#define MKFHE_B200_NTL_SHIM_H

#include <cstdint>
#include <stdexcept>
#include <vector>

namespace NTL {

struct ZZ {
    long v{0};
    ZZ() = default;
    explicit ZZ(long x) : v(x) {}
};

namespace detail {
inline long& modulus() {
    static thread_local long q = 2;
    return q;
}
inline long norm(long a) {
    long q = modulus();
    a %= q;
    return a < 0 ? a + q : a;
}
inline long mulmod(long a, long b) {
    return static_cast<long>((static_cast<unsigned __int128>(a) * static_cast<unsigned __int128>(b)) %
                             static_cast<unsigned long>(modulus()));
}
inline long invmod(long a) {  // modulus is prime on every call site; extended Euclid, throws on 0
    long q = modulus();
    long t = 0, nt = 1, r = q, nr = norm(a);
    while (nr != 0) {
        long qq = r / nr;
        long tmp = t - qq * nt; t = nt; nt = tmp;
        tmp = r - qq * nr; r = nr; nr = tmp;
    }
    if (r != 1) throw std::runtime_error("NTL shim: element not invertible");
    return t < 0 ? t + q : t;
}
}  // namespace detail

class ZZ_p {
    long m_v{0};
public:
    ZZ_p() = default;
    static void init(const ZZ& q) { detail::modulus() = q.v; }
    ZZ_p& operator=(long a) { m_v = detail::norm(a); return *this; }  // double arguments convert to long (truncation), as with NTL
    long rep() const { return m_v; }
    static ZZ_p from_rep(long r) { ZZ_p x; x.m_v = r; return x; }
};

template <class T> T conv(const ZZ_p& a);
template <> inline long conv<long>(const ZZ_p& a) { return a.rep(); }

template <class T>
class Vec {
    std::vector<T> m_d;
public:
    Vec() = default;
    explicit Vec(size_t n) : m_d(n) {}
    void SetLength(size_t n) { m_d.resize(n); }
    long length() const { return static_cast<long>(m_d.size()); }
    T& operator[](long i) { return m_d[i]; }
    const T& operator[](long i) const { return m_d[i]; }
};

// ---- polynomials over Z_p ----
class ZZ_pX {
public:
    std::vector<long> c;  // canonical residues, c.back() != 0 (normalised)
    void normalize() { while (!c.empty() && c.back() == 0) c.pop_back(); }
    ZZ_p operator[](long i) const { return ZZ_p::from_rep(i < static_cast<long>(c.size()) ? c[i] : 0); }
};
inline long deg(const ZZ_pX& a) { return static_cast<long>(a.c.size()) - 1; }
inline void SetCoeff(ZZ_pX& x, long i, long a) {
    if (i >= static_cast<long>(x.c.size())) x.c.resize(i + 1, 0);
    x.c[i] = detail::norm(a);
    x.normalize();
}
inline void SetCoeff(ZZ_pX& x, long i, const ZZ_p& a) { SetCoeff(x, i, a.rep()); }

// x = a^{-1} mod f ; throws if gcd(a, f) != 1 (NTL raises an error there; the reference wraps it in try/catch).
inline void InvMod(ZZ_pX& x, const ZZ_pX& a, const ZZ_pX& f) {
    using detail::mulmod; using detail::invmod; using detail::modulus;
    const long q = modulus();
    // extended Euclid on (r0, r1) = (f, a mod f) tracking the cofactor of a only
    std::vector<long> r0 = f.c, r1 = a.c, t0, t1{1};
    auto trim = [](std::vector<long>& v) { while (!v.empty() && v.back() == 0) v.pop_back(); };
    trim(r0); trim(r1);
    if (r1.empty()) throw std::runtime_error("NTL shim: InvMod of zero");
    while (!r1.empty()) {
        // r0 = quo * r1 + rem ; t_new = t0 - quo * t1
        std::vector<long> quo, rem = r0;
        long d1 = static_cast<long>(r1.size()) - 1;
        long lead_inv = invmod(r1.back());
        if (static_cast<long>(rem.size()) - 1 >= d1) quo.assign(rem.size() - r1.size() + 1, 0);
        for (long i = static_cast<long>(rem.size()) - 1; i >= d1; --i) {
            long coef = mulmod(rem[i], lead_inv);
            if (coef == 0) continue;
            quo[i - d1] = coef;
            for (long j = 0; j <= d1; ++j) {
                long s = rem[i - d1 + j] - mulmod(coef, r1[j]);
                rem[i - d1 + j] = s < 0 ? s + q : s;
            }
        }
        trim(rem);
        std::vector<long> tn = t0;
        if (!quo.empty() && !t1.empty()) {
            if (tn.size() < quo.size() + t1.size() - 1) tn.resize(quo.size() + t1.size() - 1, 0);
            for (size_t i = 0; i < quo.size(); ++i) {
                if (quo[i] == 0) continue;
                for (size_t j = 0; j < t1.size(); ++j) {
                    long s = tn[i + j] - mulmod(quo[i], t1[j]);
                    tn[i + j] = s < 0 ? s + q : s;
                }
            }
        }
        trim(tn);
        r0.swap(r1); r1.swap(rem);
        t0.swap(t1); t1.swap(tn);
    }
    if (r0.size() != 1) throw std::runtime_error("NTL shim: InvMod: not invertible");
    long ginv = invmod(r0[0]);
    x.c.assign(t0.size(), 0);
    for (size_t i = 0; i < t0.size(); ++i) x.c[i] = mulmod(t0[i], ginv);
    x.normalize();
}

// ---- matrices over Z_p ----
struct INIT_SIZE_TYPE {};
static const INIT_SIZE_TYPE INIT_SIZE{};

class mat_ZZ_p {
    std::vector<Vec<ZZ_p>> m_rows;
    long m_r{0}, m_c{0};
public:
    mat_ZZ_p() = default;
    mat_ZZ_p(INIT_SIZE_TYPE, long r, long c) { SetDims(r, c); }
    void SetDims(long r, long c) { m_r = r; m_c = c; m_rows.assign(r, Vec<ZZ_p>(c)); }
    long NumRows() const { return m_r; }
    long NumCols() const { return m_c; }
    Vec<ZZ_p>& operator[](long i) { return m_rows[i]; }
    const Vec<ZZ_p>& operator[](long i) const { return m_rows[i]; }
};

// X = A^{-1}; throws when A is singular (Gauss-Jordan over the prime field).
inline void inv(mat_ZZ_p& X, const mat_ZZ_p& A) {
    using detail::invmod; using detail::modulus;
    const long n = A.NumRows();
    const unsigned long q = static_cast<unsigned long>(modulus());
    std::vector<unsigned long> M(static_cast<size_t>(n) * 2 * n, 0);
    for (long i = 0; i < n; ++i) {
        for (long j = 0; j < n; ++j) M[i * 2 * n + j] = static_cast<unsigned long>(A[i][j].rep());
        M[i * 2 * n + n + i] = 1;
    }
    for (long col = 0; col < n; ++col) {
        long piv = -1;
        for (long r = col; r < n; ++r) if (M[r * 2 * n + col] != 0) { piv = r; break; }
        if (piv < 0) throw std::runtime_error("NTL shim: inv: singular matrix");
        if (piv != col) for (long j = 0; j < 2 * n; ++j) std::swap(M[piv * 2 * n + j], M[col * 2 * n + j]);
        unsigned long pinv = static_cast<unsigned long>(invmod(static_cast<long>(M[col * 2 * n + col])));
        for (long j = 0; j < 2 * n; ++j) M[col * 2 * n + j] = (M[col * 2 * n + j] * pinv) % q;
        for (long r = 0; r < n; ++r) {
            if (r == col) continue;
            unsigned long fct = M[r * 2 * n + col];
            if (fct == 0) continue;
            unsigned long neg = q - fct;
            unsigned long* rr = &M[r * 2 * n];
            const unsigned long* pc = &M[col * 2 * n];
            for (long j = 0; j < 2 * n; ++j) rr[j] = (rr[j] + neg * pc[j]) % q;
        }
    }
    X.SetDims(n, n);
    for (long i = 0; i < n; ++i)
        for (long j = 0; j < n; ++j) X[i][j] = ZZ_p::from_rep(static_cast<long>(M[i * 2 * n + n + j]));
}

}  // namespace NTL
#endif

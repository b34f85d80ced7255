"""Key-generation parity helpers (test infrastructure).

The reference's MKKeyGen samples its randomness internally (clock-seeded) and does not expose it.  It can be
recovered exactly from the keys it produced together with the secrets: with s = f_u (EVALUATION) and g_j = B^(j+1),
    f_j * s            = NTT(e1_j) + g_j * NTT(r)      ->  r = round(centred(INTT(f_0 s)) / B),  e1_j = the rest
    d_j - R[j] CRS_j - [m] g_j (* s_inv if KDM) = NTT(e0_j)                       (m is the one that leaves a small e0_j)
    Pkey_j + CRS_j s   = NTT(e_j)
(mk-acc-xzw.cpp:132-228, binfhe-base-scheme.cpp:250-269).  Feeding the recovered randomness to a key generator must give the
reference's key bytes back, which pins the generator's formulas, transform ordering and layouts against the reference.
"""
import numpy as np


def slot_inverse(f, Q):
    return np.array([[pow(int(x), Q - 2, Q) for x in row] for row in f], dtype=np.uint32)


def evk_index(p, e):
    """evk number e of the boundary layout [u][s][i] + evk*  ->  (u, s, i, kdm)."""
    S = 2 if p.variant == 0 else 1
    E = p.k * S * p.n + 1
    if e == E - 1:
        return 0, 0, p.n, True
    u, rem = divmod(e, S * p.n)
    s, i = divmod(rem, p.n)
    return u, s, i, (u == 0 and i == 0)


def extract_randomness(p, o, bk, pkey, f, crs, small=64):
    """Returns dict(r [E][N], e [E][ell][2][N], m [E], kdm [E], party [E], pk_e [k][ell][N]) as uint32 residues mod Q."""
    Q, N, ell = p.Q, p.N, p.digitsG - 1
    B = 1 << p.baseG_log2
    bk = np.asarray(bk, dtype=np.uint32).reshape(-1, ell, 2, N).astype(np.int64)
    pkey = np.asarray(pkey, dtype=np.uint32).reshape(p.k, ell, N).astype(np.int64)
    f = np.asarray(f, dtype=np.uint32).reshape(p.k, N).astype(np.int64)
    crs = np.asarray(crs, dtype=np.uint32).reshape(ell, N).astype(np.int64)
    inv = slot_inverse(f, Q).astype(np.int64)
    E = bk.shape[0]

    def intt(x):
        return o.ntt(np.asarray(x % Q, dtype=np.uint32), True).astype(np.int64)

    def ntt(x):
        return o.ntt(np.asarray(x % Q, dtype=np.uint32), False).astype(np.int64)

    def cen(x):
        x = x % Q
        return np.where(x > Q // 2, x - Q, x)

    out = dict(r=np.zeros((E, N), np.uint32), e=np.zeros((E, ell, 2, N), np.uint32), m=np.zeros(E, np.uint8),
               kdm=np.zeros(E, np.uint8), party=np.zeros(E, np.uint32), pk_e=np.zeros((p.k, ell, N), np.uint32))
    for e in range(E):
        u, _, _, kdm = evk_index(p, e)
        c = [cen(intt(bk[e, j, 1] * f[u] % Q)) for j in range(ell)]
        rr = np.rint(c[0] / B).astype(np.int64)
        R = ntt(rr)
        g = 1
        for j in range(ell):
            g = g * B % Q
            e1 = c[j] - g * rr
            assert np.abs(e1).max() < small, "f-noise not small: wrong secret or layout"
            out["e"][e, j, 1] = (e1 % Q).astype(np.uint32)
        found = None
        for m in (0, 1):
            e0s, g = [], 1
            for j in range(ell):
                g = g * B % Q
                d = bk[e, j, 0] - R[j] * crs[j]
                if m:
                    d = d - (g * inv[u] if kdm else g)
                e0s.append(cen(intt(d % Q)))
            if max(np.abs(x).max() for x in e0s) < small:
                found = (m, e0s)
                break
        assert found is not None, "d-noise not small for either message"
        out["m"][e], out["kdm"][e], out["party"][e] = found[0], int(kdm), u
        out["r"][e] = (rr % Q).astype(np.uint32)
        for j in range(ell):
            out["e"][e, j, 0] = (found[1][j] % Q).astype(np.uint32)
    for u in range(p.k):
        for j in range(ell):
            ee = cen(intt((pkey[u, j] + crs[j] * f[u]) % Q))
            assert np.abs(ee).max() < small
            out["pk_e"][u, j] = (ee % Q).astype(np.uint32)
    return out


def expected_messages(p, sk):
    """m per evk from the LWE secret as KeyGenAcc derives it (mk-acc-xzw.cpp:61-84, mk-acc-xzw_B.cpp:48-62)."""
    S = 2 if p.variant == 0 else 1
    sk = np.asarray(sk).reshape(p.k, p.n)
    E = p.k * S * p.n + 1
    m = np.zeros(E, np.uint8)
    for e in range(E - 1):
        u, s, i, _ = evk_index(p, e)
        v = int(sk[u, i])
        m[e] = int(v == 1) if s == 0 else int(v == p.qKS - 1)
    m[E - 1] = 1
    return m


# ---- key-switch keys ----
def _cen(x, q):
    x = np.asarray(x).astype(np.int64) % q
    return np.where(x > q // 2, x - q, x)


def switch_modulus(f_coef, Q, q):
    """NativeVector::SwitchModulus(qKS) of the uni-encryption secret coefficients: centred lift, then mod q."""
    return (_cen(f_coef, Q) % q).astype(np.uint32)


def extract_ksk_mntru_noise(p, ksk, F, s_q, small=64):
    """KSK = centred(E + gadget(s)) x centred(F^-1)  =>  E = KSK x F - gadget(s)  (mntru-pke.cpp:624-760).  Returns E
    [k][N*dKS][n] mod qKS; the reference fills it with copies of one sampled row per party (:646)."""
    q, N, n, d, k = p.qKS, p.N, p.n, p.dKS, p.k
    ksk = np.asarray(ksk).reshape(k, N * d, n).astype(np.int64)
    out = np.zeros((k, N * d, n), np.uint32)
    for u in range(k):
        G = (ksk[u] @ np.asarray(F[u]).astype(np.int64)) % q
        for j in range(d):
            G[j::d, 0] = (G[j::d, 0] - s_q[u].astype(np.int64) * pow(p.baseKS, j, q)) % q
        E = _cen(G, q)
        assert np.abs(E).max() < small and (E == E[0]).all()
        out[u] = (E % q).astype(np.uint32)
    return out


def extract_ksk_mklwe_noise(p, A, B, s, sN_q, small=64):
    """e = B - <A, s> - sN_i a0 baseKS^j  (mklwe-pke.cpp:220-235).  Returns e [k][N*baseKS*dKS] mod qKS."""
    q, N, n, d, k, base = p.qKS, p.N, p.n, p.dKS, p.k, p.baseKS
    A = np.asarray(A).reshape(k, N, base, d, n).astype(np.int64)
    B = np.asarray(B).reshape(k, N, base, d).astype(np.int64)
    s = np.asarray(s).reshape(k, n).astype(np.int64)
    out = np.zeros((k, N * base * d), np.uint32)
    for u in range(k):
        msg = np.zeros((N, base, d), np.int64)
        for a0 in range(base):
            for j in range(d):
                msg[:, a0, j] = sN_q[u].astype(np.int64) * (a0 * pow(base, j, q) % q) % q
        e = _cen(B[u] - A[u] @ s[u] - msg, q)
        assert np.abs(e).max() < small
        out[u] = (e.reshape(-1) % q).astype(np.uint32)
    return out


# ---- Encrypt: numpy restatements, the checker for the GPU's batched Encrypt ----
def encrypt_mntru_np(p, finv, e, m, pmod=4):
    """mntru-pke.cpp:158-205: c[u] = (e[u] + [u == 0] (m mod p)(q/p) e_0) x F_u^-1 mod q.  e [B][k][n], m [B]."""
    q = p.q
    e = np.asarray(e, dtype=np.int64).reshape(-1, p.k, p.n) % q
    e[:, 0, 0] = (e[:, 0, 0] + (np.asarray(m, dtype=np.int64) % pmod) * (q // pmod)) % q
    finv = np.asarray(finv, dtype=np.int64).reshape(p.k, p.n, p.n) % q
    return np.stack([(e[:, u, :] @ finv[u]) % q for u in range(p.k)], axis=1).astype(np.uint32)


def encrypt_mklwe_np(p, s, a, err, m, pmod=4):
    """mklwe-pke.cpp:36-64: b = (m mod p)(q/p) + err + sum_u <a_u, s_u> mod q.  a [B][k][n], err [B], m [B]."""
    q = p.q
    a = np.asarray(a, dtype=np.int64).reshape(-1, p.k * p.n) % q
    s = np.asarray(s, dtype=np.int64).reshape(p.k * p.n) % q
    return (((np.asarray(m, dtype=np.int64) % pmod) * (q // pmod) + np.asarray(err, dtype=np.int64) + a @ s) % q).astype(np.uint32)

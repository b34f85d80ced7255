// mkfhe_b200.cu -- context, key upload and the C ABI of include/mkfhe_b200.h.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC (see mkfhe_b200/build.py).
//
// There is no CPU fallback in this library: every entry point that computes launches CUDA kernels.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/mkfhe_b200.h"
#include "fast_kernels.cuh"
#include "cluster_kernels.cuh"
#include "generic_kernels.cuh"
#include "keyfile.h"

using namespace mkfhe;

#define MKFHE_VERSION_STR "mkfhe_b200 0.1.0 sm_100a"

struct mkfhe_ctx {
    mkfhe_params p{};
    DevCtx d{};
    int device{0};
    cudaStream_t stream{nullptr};
    cudaStream_t own_stream{nullptr};
    std::string err;
    uint64_t launches{0};
    uint32_t last_sms_per_gate{1};  // SMs per gate of the last blind-rotation launch (cluster size, else 1)
    uint64_t dev_bytes{0};
    // tables
    uint32_t *tw{nullptr}, *twi{nullptr}, *psipow{nullptr};
    // keys
    uint32_t *bk{nullptr}, *pkey{nullptr}, *ksk_mntru{nullptr}, *kskA{nullptr}, *kskB{nullptr}, *ctnand{nullptr}, *sk{nullptr};
    uint16_t* ksk16{nullptr};  // fused key switch: u16 rows (generic_kernels.cuh, DevCtx::ksk16)
    bool fuse_ks{false};
    // workspaces (grown on demand)
    uint32_t* ws_ms{nullptr};
    size_t ws_ms_words{0};
    uint32_t* ws_raw{nullptr};
    size_t ws_raw_words{0};
    uint32_t* ws_io{nullptr};
    size_t ws_io_words{0};
    uint32_t r2{0};  // 2^64 mod Q
    FastState fast{};
    bool use_fast{false};
    bool ntt_half{false};  // N = 1024: batched transforms on the two-polynomials-per-pass path of fast_kernels.cuh
    cudaEvent_t ev[3]{nullptr, nullptr, nullptr};  // before blind rotation | after it | after key switch (last gate batch)
    bool ev_valid{false};
    // stream-ordered pool behind the per-call scratch of mkfhe_keygen_* / mkfhe_encrypt_* (TmpBuf): repeated calls reuse the same
    // device memory instead of paying cudaMalloc + cudaFree (about 2 ms a call, more than the kernels take)
    cudaMemPool_t pool{nullptr};
};

static thread_local std::string g_create_error;

static int fail(mkfhe_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c)
        c->err = buf;
    else
        g_create_error = buf;
    return code;
}

#define CU(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess)                                                                           \
            return fail(ctx, MKFHE_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// ---- host number theory (table construction only) ----
static uint32_t h_mulmod(uint32_t a, uint32_t b, uint32_t m) { return uint32_t((uint64_t(a) * b) % m); }
static uint32_t h_powmod(uint32_t a, uint64_t e, uint32_t m) {
    uint64_t r = 1, b = a % m;
    while (e) {
        if (e & 1) r = (r * b) % m;
        b = (b * b) % m;
        e >>= 1;
    }
    return uint32_t(r);
}
static uint32_t h_bitrev(uint32_t x, uint32_t bits) {
    uint32_t r = 0;
    for (uint32_t i = 0; i < bits; ++i) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}
static bool h_is_prime(uint32_t n) {
    if (n < 2) return false;
    for (uint64_t d = 2; d * d <= n; ++d)
        if (n % d == 0) return false;
    return true;
}

template <class T>
static int dev_alloc(mkfhe_ctx* ctx, T** p, size_t count) {
    CU(cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T)));
    ctx->dev_bytes += count * sizeof(T);
    return MKFHE_OK;
}
static int grow(mkfhe_ctx* ctx, uint32_t** p, size_t* have, size_t want) {
    if (*have >= want) return MKFHE_OK;
    if (*p) {
        CU(cudaStreamSynchronize(ctx->stream));
        CU(cudaFree(*p));
        ctx->dev_bytes -= *have * sizeof(uint32_t);
        *p = nullptr;
        *have = 0;
    }
    CU(cudaMalloc(reinterpret_cast<void**>(p), want * sizeof(uint32_t)));
    ctx->dev_bytes += want * sizeof(uint32_t);
    *have = want;
    return MKFHE_OK;
}

static size_t bk_words(const mkfhe_params& p) {
    const size_t S = p.variant == MKFHE_VARIANT_XZW ? 2 : 1;
    return (size_t(p.k) * S * p.n + 1) * (p.digitsG - 1) * 2 * p.N;
}

// upload `words` residues (host or device source) into *dst, converting to Montgomery form mod Q when mont != 0
static int upload(mkfhe_ctx* ctx, uint32_t** dst, const uint32_t* src, uint64_t words, bool mont) {
    if (*dst == nullptr) {
        int rc = dev_alloc(ctx, dst, words);
        if (rc) return rc;
    }
    CU(cudaMemcpyAsync(*dst, src, words * sizeof(uint32_t), cudaMemcpyDefault, ctx->stream));
    if (mont) {
        const unsigned blocks = unsigned(std::min<uint64_t>((words + 255) / 256, 148 * 16));
        to_montgomery<<<blocks, 256, 0, ctx->stream>>>(*dst, *dst, words, ctx->d.mod, ctx->r2);
        ctx->launches++;
        CU(cudaGetLastError());
    }
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}

// key-switch key -> u16 rows in HBM (fused key switch).  src/bcol: host or device; staged through a temporary device copy.
static int upload_ksk16(mkfhe_ctx* ctx, const uint32_t* src, uint64_t words, const uint32_t* bcol, uint64_t rows) {
    const uint32_t n = ctx->p.n, npad = ctx->d.npad;
    uint32_t *tmp = nullptr, *tmpb = nullptr;
    CU(cudaMalloc(&tmp, words * 4));
    if (bcol) CU(cudaMalloc(&tmpb, rows * 4));
    if (!ctx->ksk16) {
        CU(cudaMalloc(&ctx->ksk16, rows * npad * sizeof(uint16_t)));
        ctx->dev_bytes += rows * npad * sizeof(uint16_t);
    }
    CU(cudaMemcpyAsync(tmp, src, words * 4, cudaMemcpyDefault, ctx->stream));
    if (bcol) CU(cudaMemcpyAsync(tmpb, bcol, rows * 4, cudaMemcpyDefault, ctx->stream));
    pack_ksk16<<<148 * 16, 256, 0, ctx->stream>>>(ctx->ksk16, tmp, tmpb, rows, n, npad);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(ctx->stream));
    cudaFree(tmp);
    if (tmpb) cudaFree(tmpb);
    ctx->d.ksk16 = ctx->ksk16;
    return MKFHE_OK;
}

// ---- circuits of gates: rows of the wire table <-> the contiguous batch of one level ----
// One CTA per gate of the level; kn = k * n words per ciphertext.  HBM-bound copies of a few KB per gate next to a gate kernel
// that works tens of milliseconds on each: their cost is the launch.
__global__ void circuit_gather(const uint32_t* __restrict__ wires, const uint32_t* __restrict__ wb, const uint32_t* __restrict__ in1,
                               const uint32_t* __restrict__ in2, uint32_t* __restrict__ c1, uint32_t* __restrict__ c2,
                               uint32_t* __restrict__ b1, uint32_t* __restrict__ b2, uint32_t kn) {
    const uint32_t g = blockIdx.x;
    const uint32_t* s1 = wires + size_t(in1[g]) * kn;
    const uint32_t* s2 = wires + size_t(in2[g]) * kn;
    for (uint32_t e = threadIdx.x; e < kn; e += blockDim.x) {
        c1[size_t(g) * kn + e] = s1[e];
        c2[size_t(g) * kn + e] = s2[e];
    }
    if (wb != nullptr && threadIdx.x == 0) {
        b1[g] = wb[in1[g]];
        b2[g] = wb[in2[g]];
    }
}
__global__ void circuit_scatter(uint32_t* __restrict__ wires, uint32_t* __restrict__ wb, const uint32_t* __restrict__ out,
                                const uint32_t* __restrict__ c, const uint32_t* __restrict__ b, uint32_t kn) {
    const uint32_t g = blockIdx.x;
    uint32_t* d = wires + size_t(out[g]) * kn;
    for (uint32_t e = threadIdx.x; e < kn; e += blockDim.x) d[e] = c[size_t(g) * kn + e];
    if (wb != nullptr && threadIdx.x == 0) wb[out[g]] = b[g];
}

// Levelisation of a netlist (host).  level[g] = 0-based level of gate g; returns the number of levels or -1 with `why` set.
static int64_t circuit_levelise(uint64_t n_wires, const uint32_t* in1, const uint32_t* in2, const uint32_t* out, uint64_t n_gates,
                                std::vector<uint32_t>& level, const char** why) {
    // depth[w]: 0 = primary input (never written so far), else 1 + level of the gate that wrote it; seen[w]: read already
    std::vector<uint32_t> depth(n_wires, 0);
    std::vector<uint8_t> seen(n_wires, 0);
    level.assign(n_gates, 0);
    uint32_t levels = 0;
    for (uint64_t g = 0; g < n_gates; ++g) {
        const uint32_t a = in1[g], b = in2[g], o = out[g];
        if (a >= n_wires || b >= n_wires || o >= n_wires) { *why = "wire index out of range"; return -1; }
        if (a == b) { *why = "Input ciphertexts should be independent (in1 == in2)"; return -1; }
        if (depth[o] != 0) { *why = "a wire is written by two gates"; return -1; }
        if (seen[o] || o == a || o == b) { *why = "a wire is read before the gate that writes it (gates must be in topological order)"; return -1; }
        seen[a] = seen[b] = 1;
        const uint32_t l = std::max(depth[a], depth[b]);
        level[g] = l;
        depth[o] = l + 1;
        levels = std::max(levels, l + 1);
    }
    return levels;
}

namespace {
struct TmpBuf {  // device scratch of one call, returned to the context's pool (in stream order) on every exit path
    mkfhe_ctx* ctx;
    std::vector<void*> p;
    explicit TmpBuf(mkfhe_ctx* c) : ctx(c) {}
    TmpBuf(const TmpBuf&) = delete;
    TmpBuf& operator=(const TmpBuf&) = delete;
    ~TmpBuf() {
        for (void* x : p) {
            if (ctx->pool) cudaFreeAsync(x, ctx->stream);
            else cudaFree(x);
        }
    }
    template <class T>
    cudaError_t get(T** out, size_t count) {
        const size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
        cudaError_t e = ctx->pool ? cudaMallocFromPoolAsync(reinterpret_cast<void**>(out), bytes, ctx->pool, ctx->stream)
                                  : cudaMalloc(reinterpret_cast<void**>(out), bytes);
        if (e == cudaSuccess) p.push_back(*out);
        return e;
    }
};
}  // namespace

// exceptions (std::bad_alloc on a huge but well-formed file) must not cross the C boundary
template <class F>
static int guarded_io(mkfhe_ctx* ctx, F&& body) {
    try {
        return body();
    } catch (const std::bad_alloc&) {
        return fail(ctx, MKFHE_ENOMEM, "out of host memory");
    } catch (const std::exception& e) {
        return fail(ctx, MKFHE_EINVAL, "%s", e.what());
    }
}

static void expected_key_words(const mkfhe_params& p, uint64_t (&w)[mkfhe_keyfile::TAG_SLOTS]) {
    using namespace mkfhe_keyfile;
    memset(w, 0, sizeof w);
    w[BK] = bk_words(p);
    w[PKEY] = uint64_t(p.k) * (p.digitsG - 1) * p.N;
    if (p.scheme == MKFHE_SCHEME_MKLWE) {
        w[KSKB] = uint64_t(p.k) * p.N * p.baseKS * p.dKS;
        w[KSKA] = w[KSKB] * p.n;
    } else {
        w[KSK] = uint64_t(p.k) * p.N * p.dKS * p.n;
        w[CTNAND] = uint64_t(p.k) * p.n;
    }
}

extern "C" {

const char* mkfhe_version(void) { return MKFHE_VERSION_STR; }

const char* mkfhe_last_error(const mkfhe_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int mkfhe_ctx_create(const mkfhe_params* p, mkfhe_ctx** out) {
    mkfhe_ctx* ctx = nullptr;  // CU() reports into g_create_error while ctx == nullptr
    if (!p || !out) return fail(nullptr, MKFHE_EINVAL, "null argument");
    *out = nullptr;
    if (p->N < 8 || p->N > 2048 || (p->N & (p->N - 1))) return fail(nullptr, MKFHE_EINVAL, "N must be a power of two in [8, 2048]");
    if (p->k == 0 || p->k > 64 || p->n == 0) return fail(nullptr, MKFHE_EINVAL, "bad k or n");
    if (p->Q >= (1u << 30) || !h_is_prime(p->Q) || (p->Q - 1) % (2 * p->N) != 0)
        return fail(nullptr, MKFHE_EINVAL, "Q must be a prime < 2^30 with Q = 1 mod 2N");
    if (p->digitsG < 2 || p->digitsG > 9 || p->baseG_log2 == 0 || p->baseG_log2 > 16)
        return fail(nullptr, MKFHE_EINVAL, "unsupported gadget (digitsG in [2,9], baseG_log2 in [1,16])");
    if (p->variant > 1 || p->scheme > 1) return fail(nullptr, MKFHE_EINVAL, "bad variant/scheme");
    if (p->q < 2 || p->qKS < 2 || p->baseKS < 2 || p->baseKS > 256 || p->dKS == 0 || p->dKS > 16)
        return fail(nullptr, MKFHE_EINVAL, "bad q/qKS/baseKS/dKS");
    if (uint64_t(p->q) * 2 * p->N >= (1ull << 32) || p->qKS >= (1u << 24))
        return fail(nullptr, MKFHE_EINVAL, "q*2N must fit 32 bits (mk-acc-xzw.cpp:110 uses 32-bit NativeInteger) and qKS < 2^24");
    if (p->q != p->qKS)  // both MK paths refresh into the input modulus (binfhecontext.cpp:137-146: qKS == q in every row); Decrypt
                         // and the fused key switch rely on it
        return fail(nullptr, MKFHE_EINVAL, "q must equal qKS (gate outputs are ciphertexts mod qKS and feed the next gate mod q)");
    if (h_powmod(p->psi, p->N, p->Q) != p->Q - 1) return fail(nullptr, MKFHE_EINVAL, "psi is not a primitive 2N-th root of unity mod Q");

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(nullptr, MKFHE_ENODEV, "no CUDA device: libmkfhe_b200 has no CPU fallback");
    int dev = p->device;
    if (dev < 0) CU(cudaGetDevice(&dev));
    if (dev >= ndev) return fail(nullptr, MKFHE_EINVAL, "device ordinal out of range");
    CU(cudaSetDevice(dev));

    ctx = new mkfhe_ctx();
    ctx->p = *p;
    ctx->device = dev;
    auto bail = [&](int rc) {
        g_create_error = ctx->err;
        mkfhe_ctx_destroy(ctx);
        return rc;
    };
    // An ordinary (blocking) stream: work on it is ordered with the legacy default stream, which is where a caller that never
    // chose a stream (PyTorch's default, plain cudaMemcpy) produces the inputs of the *_dev entry points and reads their outputs.
    if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamDefault) != cudaSuccess)
        return bail(fail(ctx, MKFHE_ECUDA, "cudaStreamCreate failed"));
    ctx->stream = ctx->own_stream;
    for (auto& e : ctx->ev)
        if (cudaEventCreate(&e) != cudaSuccess) return bail(fail(ctx, MKFHE_ECUDA, "cudaEventCreate failed"));
    if (!getenv("MKFHE_NO_POOL")) {  // scratch pool: keeps up to 512 MiB between calls; without one TmpBuf uses cudaMalloc
        cudaMemPoolProps pp{};
        pp.allocType = cudaMemAllocationTypePinned;
        pp.handleTypes = cudaMemHandleTypeNone;
        pp.location.type = cudaMemLocationTypeDevice;
        pp.location.id = ctx->device;
        uint64_t keep = 512ull << 20;
        if (cudaMemPoolCreate(&ctx->pool, &pp) != cudaSuccess ||
            cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &keep) != cudaSuccess) {
            if (ctx->pool) cudaMemPoolDestroy(ctx->pool);
            ctx->pool = nullptr;
            cudaGetLastError();
        }
    }

    const uint32_t N = p->N, Q = p->Q;
    uint32_t logN = 0;
    while ((1u << logN) < N) ++logN;
    // Montgomery constants
    uint32_t inv = 1;  // Q^-1 mod 2^32 by Newton iteration
    for (int i = 0; i < 5; ++i) inv *= 2u - Q * inv;
    const uint32_t R = uint32_t((1ull << 32) % Q);
    ctx->r2 = h_mulmod(R, R, Q);
    auto to_m = [&](uint32_t x) { return h_mulmod(x, R, Q); };

    DevCtx& d = ctx->d;
    d.k = p->k; d.N = N; d.logN = logN; d.n = p->n; d.ell = p->digitsG - 1; d.gbits = p->baseG_log2;
    d.variant = p->variant; d.scheme = p->scheme; d.S = p->variant == MKFHE_VARIANT_XZW ? 2 : 1;
    d.q = p->q; d.qKS = p->qKS; d.baseKS = p->baseKS; d.dKS = p->dKS;
    d.mod.Q = Q; d.mod.qinv_neg = 0u - inv;
    d.one_mont = R;
    d.ninv_mont = to_m(h_powmod(N % Q, uint64_t(Q) - 2, Q));

    std::vector<uint32_t> tw(N), twi(N), pp(2 * N);
    const uint32_t psi_inv = h_powmod(p->psi, uint64_t(Q) - 2, Q);
    uint32_t x = 1, xi = 1;
    for (uint32_t i = 0; i < N; ++i) {
        const uint32_t r = h_bitrev(i, logN);
        tw[r] = to_m(x);
        twi[r] = to_m(xi);
        x = h_mulmod(x, p->psi, Q);
        xi = h_mulmod(xi, psi_inv, Q);
    }
    x = 1;
    for (uint32_t e = 0; e < 2 * N; ++e) {
        pp[e] = to_m(x);
        x = h_mulmod(x, p->psi, Q);
    }
    int rc;
    if ((rc = upload(ctx, &ctx->tw, tw.data(), N, false))) return bail(rc);
    if ((rc = upload(ctx, &ctx->twi, twi.data(), N, false))) return bail(rc);
    if ((rc = upload(ctx, &ctx->psipow, pp.data(), 2 * N, false))) return bail(rc);
    d.tw = ctx->tw; d.twi = ctx->twi; d.psipow = ctx->psipow;

    // specialised N = 2048 path (fast_kernels.cuh)
    ctx->use_fast = fast_supported(d);
    if (ctx->use_fast && (rc = fast_init(ctx->fast, d, p->psi, ctx->stream, &ctx->dev_bytes))) {
        return bail(fail(ctx, MKFHE_ECUDA, "fast path initialisation failed: %s", cudaGetErrorString(cudaGetLastError())));
    }
    ctx->ntt_half = !ctx->use_fast && fast_ntt_half_supported(d);
    if (ctx->ntt_half && (rc = fast_init(ctx->fast, d, p->psi, ctx->stream, &ctx->dev_bytes))) {
        return bail(fail(ctx, MKFHE_ECUDA, "N = 1024 transform tables: %s", cudaGetErrorString(cudaGetLastError())));
    }
    if (getenv("MKFHE_FORCE_GENERIC")) ctx->use_fast = ctx->ntt_half = false;
    // fused key switch: keys as u16 rows, the gate kernel finishes the gate on chip.  Needs qKS < 2^16 and N + npad words of
    // scratch behind the accumulator ((ell + 2) * N words exist in both kernels).
    d.npad = (p->n + (p->scheme == MKFHE_SCHEME_MKLWE ? 1u : 0u) + 15u) & ~15u;
    d.ks_rows = p->scheme == MKFHE_SCHEME_MKLWE ? N * p->baseKS * p->dKS : N * p->dKS;
    d.ks_shift = 0;
    if ((p->baseKS & (p->baseKS - 1)) == 0)
        while ((1u << d.ks_shift) < p->baseKS) ++d.ks_shift;
    {
        const uint64_t per = uint64_t(p->dKS) * (p->scheme == MKFHE_SCHEME_MKLWE ? 1u : p->baseKS - 1) * (p->qKS - 1);
        d.ks_block = uint32_t(std::max<uint64_t>(1, std::min<uint64_t>(N, 0xffffffffull / std::max<uint64_t>(per, 1))));
    }
    ctx->fuse_ks = p->qKS < 65536 && size_t(N) + d.npad <= size_t(d.ell + 2) * N && !getenv("MKFHE_UNFUSED_KEYSWITCH");
    *out = ctx;
    return MKFHE_OK;
}

void mkfhe_ctx_destroy(mkfhe_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->own_stream) cudaStreamSynchronize(ctx->own_stream);
    fast_destroy(ctx->fast);
    for (uint32_t* p : {ctx->tw, ctx->twi, ctx->psipow, ctx->bk, ctx->pkey, ctx->ksk_mntru, ctx->kskA, ctx->kskB, ctx->ctnand, ctx->sk,
                        ctx->ws_ms, ctx->ws_raw, ctx->ws_io})
        if (p) cudaFree(p);
    if (ctx->ksk16) cudaFree(ctx->ksk16);
    for (auto& e : ctx->ev)
        if (e) cudaEventDestroy(e);
    if (ctx->pool) cudaMemPoolDestroy(ctx->pool);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

int mkfhe_ctx_set_stream(mkfhe_ctx* ctx, void* s) {
    if (!ctx) return MKFHE_EINVAL;
    cudaStream_t next = s ? static_cast<cudaStream_t>(s) : ctx->own_stream;
    if (next != ctx->stream) {
        // work and stream-ordered scratch frees of the old stream finish before anything is enqueued on the new one
        CU(cudaSetDevice(ctx->device));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    ctx->stream = next;
    return MKFHE_OK;
}
int mkfhe_ctx_synchronize(mkfhe_ctx* ctx) {
    if (!ctx) return MKFHE_EINVAL;
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}
uint32_t mkfhe_last_gate_sms(const mkfhe_ctx* ctx) { return ctx ? ctx->last_sms_per_gate : 0u; }
uint64_t mkfhe_launch_count(const mkfhe_ctx* ctx) { return ctx ? ctx->launches : 0; }
uint64_t mkfhe_device_bytes(const mkfhe_ctx* ctx) { return ctx ? ctx->dev_bytes : 0; }

// ---- keys ----
int mkfhe_load_bk(mkfhe_ctx* ctx, const uint32_t* bk, uint64_t words) {
    if (!ctx || !bk) return MKFHE_EINVAL;
    if (words != bk_words(ctx->p)) return fail(ctx, MKFHE_EINVAL, "mkfhe_load_bk: expected %zu words, got %llu", bk_words(ctx->p), (unsigned long long)words);
    CU(cudaSetDevice(ctx->device));
    int rc = upload(ctx, &ctx->bk, bk, words, true);
    ctx->d.bk = ctx->bk;
    if (!rc && ctx->use_fast) {
        cudaError_t e = fast_load_bk(ctx->fast, ctx->d, words, ctx->stream, &ctx->dev_bytes, &ctx->launches);
        if (e != cudaSuccess) return fail(ctx, MKFHE_ECUDA, "fast_load_bk: %s", cudaGetErrorString(e));
    }
    return rc;
}
int mkfhe_load_pkey(mkfhe_ctx* ctx, const uint32_t* pk, uint64_t words) {
    if (!ctx || !pk) return MKFHE_EINVAL;
    if (words != uint64_t(ctx->p.k) * (ctx->p.digitsG - 1) * ctx->p.N) return fail(ctx, MKFHE_EINVAL, "mkfhe_load_pkey: wrong size");
    CU(cudaSetDevice(ctx->device));
    int rc = upload(ctx, &ctx->pkey, pk, words, true);
    ctx->d.pkey = ctx->pkey;
    if (!rc && ctx->use_fast) {
        cudaError_t e = fast_load_pkey(ctx->fast, ctx->d, words, ctx->stream, &ctx->dev_bytes, &ctx->launches);
        if (e != cudaSuccess) return fail(ctx, MKFHE_ECUDA, "fast_load_pkey: %s", cudaGetErrorString(e));
    }
    return rc;
}
int mkfhe_load_ksk_mntru(mkfhe_ctx* ctx, const uint32_t* ksk, uint64_t words) {
    if (!ctx || !ksk) return MKFHE_EINVAL;
    if (ctx->p.scheme != MKFHE_SCHEME_MNTRU) return fail(ctx, MKFHE_EINVAL, "context is not MNTRU");
    if (words != uint64_t(ctx->p.k) * ctx->p.N * ctx->p.dKS * ctx->p.n) return fail(ctx, MKFHE_EINVAL, "mkfhe_load_ksk_mntru: wrong size");
    CU(cudaSetDevice(ctx->device));
    if (ctx->fuse_ks) return upload_ksk16(ctx, ksk, words, nullptr, uint64_t(ctx->p.k) * ctx->d.ks_rows);
    return upload(ctx, &ctx->ksk_mntru, ksk, words, false);
}
int mkfhe_load_ksk_mklwe(mkfhe_ctx* ctx, const uint32_t* A, uint64_t wordsA, const uint32_t* B, uint64_t wordsB) {
    if (!ctx || !A || !B) return MKFHE_EINVAL;
    if (ctx->p.scheme != MKFHE_SCHEME_MKLWE) return fail(ctx, MKFHE_EINVAL, "context is not MKLWE");
    const uint64_t rows = uint64_t(ctx->p.k) * ctx->p.N * ctx->p.baseKS * ctx->p.dKS;
    if (wordsB != rows || wordsA != rows * ctx->p.n) return fail(ctx, MKFHE_EINVAL, "mkfhe_load_ksk_mklwe: wrong size");
    CU(cudaSetDevice(ctx->device));
    if (ctx->fuse_ks) return upload_ksk16(ctx, A, wordsA, B, rows);
    int rc = upload(ctx, &ctx->kskA, A, wordsA, false);
    if (rc) return rc;
    return upload(ctx, &ctx->kskB, B, wordsB, false);
}
int mkfhe_load_ctnand(mkfhe_ctx* ctx, const uint32_t* ct, uint64_t words) {
    if (!ctx || !ct) return MKFHE_EINVAL;
    if (ctx->p.scheme != MKFHE_SCHEME_MNTRU) return fail(ctx, MKFHE_EINVAL, "ctNAND exists only on the MNTRU path");
    if (words != uint64_t(ctx->p.k) * ctx->p.n) return fail(ctx, MKFHE_EINVAL, "mkfhe_load_ctnand: wrong size");
    CU(cudaSetDevice(ctx->device));
    return upload(ctx, &ctx->ctnand, ct, words, false);
}

// ---- resident key images (replication between GPUs) ----
static uint64_t ksk16_bytes(const mkfhe_ctx* ctx) { return uint64_t(ctx->p.k) * ctx->d.ks_rows * ctx->d.npad * sizeof(uint16_t); }
int mkfhe_key_image(mkfhe_ctx* ctx, int which, void** dev_ptr, uint64_t* bytes) {
    if (!ctx || !dev_ptr || !bytes) return MKFHE_EINVAL;
    CU(cudaSetDevice(ctx->device));
    const mkfhe_params& p = ctx->p;
    int rc = MKFHE_OK;
    switch (which) {
        case MKFHE_IMAGE_BK:
            *bytes = bk_words(p) * 4;
            if (!ctx->bk) rc = dev_alloc(ctx, &ctx->bk, bk_words(p));
            *dev_ptr = ctx->bk;
            break;
        case MKFHE_IMAGE_PKEY:
            *bytes = uint64_t(p.k) * (p.digitsG - 1) * p.N * 4;
            if (!ctx->pkey) rc = dev_alloc(ctx, &ctx->pkey, *bytes / 4);
            *dev_ptr = ctx->pkey;
            break;
        case MKFHE_IMAGE_KSK:
            if (!ctx->fuse_ks) return fail(ctx, MKFHE_EINVAL, "the key-switch key has a resident image only on the fused path (qKS < 2^16)");
            *bytes = ksk16_bytes(ctx);
            if (!ctx->ksk16) {
                CU(cudaMalloc(&ctx->ksk16, *bytes));
                ctx->dev_bytes += *bytes;
            }
            *dev_ptr = ctx->ksk16;
            break;
        case MKFHE_IMAGE_CTNAND:
            if (p.scheme != MKFHE_SCHEME_MNTRU) return fail(ctx, MKFHE_EINVAL, "ctNAND exists only on the MNTRU path");
            *bytes = uint64_t(p.k) * p.n * 4;
            if (!ctx->ctnand) rc = dev_alloc(ctx, &ctx->ctnand, *bytes / 4);
            *dev_ptr = ctx->ctnand;
            break;
        default: return fail(ctx, MKFHE_EINVAL, "unknown key image %d", which);
    }
    return rc;
}
int mkfhe_key_image_commit(mkfhe_ctx* ctx, int which) {
    if (!ctx) return MKFHE_EINVAL;
    switch (which) {
        case MKFHE_IMAGE_BK: if (!ctx->bk) break; ctx->d.bk = ctx->bk; return MKFHE_OK;
        case MKFHE_IMAGE_PKEY: if (!ctx->pkey) break; ctx->d.pkey = ctx->pkey; return MKFHE_OK;
        case MKFHE_IMAGE_KSK: if (!ctx->ksk16) break; ctx->d.ksk16 = ctx->ksk16; return MKFHE_OK;
        case MKFHE_IMAGE_CTNAND: if (!ctx->ctnand) break; return MKFHE_OK;
        default: return fail(ctx, MKFHE_EINVAL, "unknown key image %d", which);
    }
    return fail(ctx, MKFHE_ENOKEY, "key image %d was never obtained with mkfhe_key_image", which);
}

// the decryptor's secret (the party set that decrypts jointly): residues mod q
int mkfhe_load_sk(mkfhe_ctx* ctx, const uint32_t* sk, uint64_t words) {
    if (!ctx || !sk) return MKFHE_EINVAL;
    if (words != uint64_t(ctx->p.k) * ctx->p.n) return fail(ctx, MKFHE_EINVAL, "mkfhe_load_sk: wrong size");
    CU(cudaSetDevice(ctx->device));
    int rc = upload(ctx, &ctx->sk, sk, words, false);
    if (rc) return rc;
    reduce_mod<<<unsigned((words + 255) / 256), 256, 0, ctx->stream>>>(ctx->sk, words, ctx->p.q);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}

int mkfhe_decrypt_dev(mkfhe_ctx* ctx, const uint32_t* d_ct, const uint32_t* d_b, uint32_t p, uint32_t* d_m, uint64_t batch) {
    if (!ctx || !d_ct || !d_m) return MKFHE_EINVAL;
    if (!ctx->sk) return fail(ctx, MKFHE_ENOKEY, "secret key not loaded (mkfhe_load_sk)");
    if (ctx->p.scheme == MKFHE_SCHEME_MKLWE && !d_b) return fail(ctx, MKFHE_EINVAL, "MKLWE ciphertexts need b");
    if (p < 2 || 2 * uint64_t(p) > ctx->p.q) return fail(ctx, MKFHE_EINVAL, "bad plaintext modulus");
    if (batch == 0) return MKFHE_OK;
    CU(cudaSetDevice(ctx->device));
    const uint64_t blocks = (batch * 32 + 255) / 256;
    if (blocks > 0x7fffffffull) return fail(ctx, MKFHE_EINVAL, "batch too large");
    decrypt_batch<<<unsigned(blocks), 256, 0, ctx->stream>>>(ctx->d, ctx->sk, d_ct, d_b, p, d_m, batch);
    ctx->launches++;
    CU(cudaGetLastError());
    return MKFHE_OK;
}

int mkfhe_decrypt(mkfhe_ctx* ctx, const uint32_t* ct, const uint32_t* b, uint32_t p, uint32_t* m, uint64_t batch) {
    if (!ctx || !ct || !m) return MKFHE_EINVAL;
    const bool lwe = ctx->p.scheme == MKFHE_SCHEME_MKLWE;
    if (lwe && !b) return fail(ctx, MKFHE_EINVAL, "MKLWE ciphertexts need b");
    if (batch == 0) return MKFHE_OK;
    CU(cudaSetDevice(ctx->device));
    const size_t cw = size_t(batch) * ctx->p.k * ctx->p.n;
    int rc = grow(ctx, &ctx->ws_io, &ctx->ws_io_words, cw + 2 * batch);
    if (rc) return rc;
    uint32_t *dc = ctx->ws_io, *db = dc + cw, *dm = db + batch;
    CU(cudaMemcpyAsync(dc, ct, cw * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (lwe) CU(cudaMemcpyAsync(db, b, batch * 4, cudaMemcpyHostToDevice, ctx->stream));
    rc = mkfhe_decrypt_dev(ctx, dc, lwe ? db : nullptr, p, dm, batch);
    if (rc) return rc;
    CU(cudaMemcpyAsync(m, dm, batch * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}

// ---- launch helpers ----
static int generic_rotate_config(mkfhe_ctx* ctx, size_t* smem, unsigned* threads) {
    const DevCtx& d = ctx->d;
    size_t bytes = size_t(d.k + d.ell + 2) * d.N * 4 + ((size_t(d.k) * d.n * 2 + 15) & ~size_t(15));
    if (bytes > 227 * 1024) return fail(ctx, MKFHE_EINVAL, "parameter set needs %zu bytes of shared memory per gate (> 227 KB)", bytes);
    unsigned ctas = unsigned(std::max<size_t>(1, std::min<size_t>(8, (227 * 1024) / bytes)));
    unsigned t = 2048 / ctas / 2, p2 = 32;
    while (p2 * 2 <= t) p2 *= 2;
    p2 = std::min(p2, std::min(1024u, std::max(32u, d.N / 2)));
    *smem = bytes;
    *threads = std::max(32u, p2);
    return MKFHE_OK;
}

static int launch_rotate(mkfhe_ctx* ctx, bool gate, const GateIO& io, uint64_t batch) {
    if (!ctx->bk || !ctx->pkey) return fail(ctx, MKFHE_ENOKEY, "bootstrapping key / Pkey not loaded");
    if (batch == 0) return MKFHE_OK;
    if (batch > 0x7fffffffull) return fail(ctx, MKFHE_EINVAL, "batch too large");
    const uint32_t CL = ctx->use_fast ? cluster_size_for(ctx->fast, ctx->d, gate, uint32_t(batch)) : 0u;
    ctx->last_sms_per_gate = CL ? CL : 1u;
    if (CL) {  // few gates: 2, 4 or 8 SMs per gate
        cudaError_t e = cluster_blind_rotate(ctx->fast, ctx->d, gate, CL, io, uint32_t(batch), ctx->stream, &ctx->launches);
        if (e != cudaSuccess) return fail(ctx, MKFHE_ECUDA, "cluster_blind_rotate: %s", cudaGetErrorString(e));
        return MKFHE_OK;
    }
    if (ctx->use_fast) {
        cudaError_t e = fast_blind_rotate(ctx->fast, ctx->d, gate, io, uint32_t(batch), ctx->stream, &ctx->launches);
        if (e != cudaSuccess) return fail(ctx, MKFHE_ECUDA, "fast_blind_rotate: %s", cudaGetErrorString(e));
        return MKFHE_OK;
    }
    size_t smem;
    unsigned threads;
    int rc = generic_rotate_config(ctx, &smem, &threads);
    if (rc) return rc;
    if (gate) {
        CU(cudaFuncSetAttribute(blind_rotate_generic<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        blind_rotate_generic<true><<<unsigned(batch), threads, smem, ctx->stream>>>(ctx->d, io);
    } else {
        CU(cudaFuncSetAttribute(blind_rotate_generic<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        blind_rotate_generic<false><<<unsigned(batch), threads, smem, ctx->stream>>>(ctx->d, io);
    }
    ctx->launches++;
    CU(cudaGetLastError());
    return MKFHE_OK;
}

int mkfhe_eval_acc_dev(mkfhe_ctx* ctx, const uint32_t* d_acc_in, const uint32_t* d_rot, uint32_t rot_mod, uint32_t* d_acc_out,
                       uint64_t batch) {
    if (!ctx || !d_acc_in || !d_rot || !d_acc_out) return MKFHE_EINVAL;
    if (ctx->p.variant == MKFHE_VARIANT_XZW && rot_mod == 0) return fail(ctx, MKFHE_EINVAL, "rot_mod must be non-zero for XZW");
    CU(cudaSetDevice(ctx->device));
    GateIO io{};
    io.acc_in = d_acc_in; io.rot = d_rot; io.rot_mod = rot_mod ? rot_mod : 1; io.acc_out = d_acc_out;
    return launch_rotate(ctx, false, io, batch);
}

int mkfhe_eval_acc(mkfhe_ctx* ctx, const uint32_t* acc_in, const uint32_t* rot, uint32_t rot_mod, uint32_t* acc_out, uint64_t batch) {
    if (!ctx || !acc_in || !rot || !acc_out) return MKFHE_EINVAL;
    CU(cudaSetDevice(ctx->device));
    const size_t aw = size_t(batch) * ctx->p.k * ctx->p.N, rw = size_t(batch) * ctx->p.k * ctx->p.n;
    int rc = grow(ctx, &ctx->ws_io, &ctx->ws_io_words, aw + rw);
    if (rc) return rc;
    uint32_t* d_acc = ctx->ws_io;
    uint32_t* d_rot = ctx->ws_io + aw;
    CU(cudaMemcpyAsync(d_acc, acc_in, aw * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(d_rot, rot, rw * 4, cudaMemcpyHostToDevice, ctx->stream));
    rc = mkfhe_eval_acc_dev(ctx, d_acc, d_rot, rot_mod, d_acc, batch);
    if (rc) return rc;
    CU(cudaMemcpyAsync(acc_out, d_acc, aw * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}

int mkfhe_eval_nand_dev(mkfhe_ctx* ctx, const uint32_t* d_ct1, const uint32_t* d_b1, const uint32_t* d_ct2, const uint32_t* d_b2,
                        uint32_t* d_out, uint32_t* d_bout, uint64_t batch) {
    if (!ctx || !d_ct1 || !d_ct2 || !d_out) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    const bool lwe = p.scheme == MKFHE_SCHEME_MKLWE;
    if (lwe && (!d_b1 || !d_b2 || !d_bout)) return fail(ctx, MKFHE_EINVAL, "MKLWE gates need b1, b2 and bout");
    const bool have_ksk = ctx->fuse_ks ? ctx->ksk16 != nullptr : (lwe ? (ctx->kskA && ctx->kskB) : ctx->ksk_mntru != nullptr);
    if (!have_ksk || (!lwe && !ctx->ctnand))
        return fail(ctx, MKFHE_ENOKEY, lwe ? "MKLWE key-switch key not loaded" : "MNTRU key-switch key / ctNAND not loaded");
    if (batch == 0) return MKFHE_OK;
    CU(cudaSetDevice(ctx->device));
    int rc = ctx->fuse_ks ? MKFHE_OK : grow(ctx, &ctx->ws_ms, &ctx->ws_ms_words, size_t(batch) * p.k * p.N);
    if (rc) return rc;
    GateIO io{};
    io.ct1 = d_ct1; io.ct2 = d_ct2; io.b1 = d_b1; io.b2 = d_b2; io.ctnand = ctx->ctnand; io.ms_out = ctx->ws_ms;
    io.out = d_out; io.bout = d_bout;
    CU(cudaEventRecord(ctx->ev[0], ctx->stream));
    rc = launch_rotate(ctx, true, io, batch);
    if (rc) return rc;
    CU(cudaEventRecord(ctx->ev[1], ctx->stream));
    const uint32_t B = uint32_t(batch);
    if (ctx->fuse_ks) {
        // the gate kernel did the key switch on chip (rotate_epilogue)
    } else if (!lwe) {
        dim3 grid((p.n + 63) / 64, (B + 63) / 64, p.k);
        keyswitch_mntru_generic<<<grid, 256, 0, ctx->stream>>>(ctx->d, ctx->ws_ms, ctx->ksk_mntru, d_out, B);
        ctx->launches++;
        CU(cudaGetLastError());
    } else {
        const size_t rw = size_t(batch) * p.k * p.n + batch;
        rc = grow(ctx, &ctx->ws_raw, &ctx->ws_raw_words, rw);
        if (rc) return rc;
        CU(cudaMemsetAsync(ctx->ws_raw, 0, rw * 4, ctx->stream));
        uint32_t* braw = ctx->ws_raw + size_t(batch) * p.k * p.n;
        const uint32_t ichunk = std::min(p.N, 256u);
        dim3 grid(B, p.k, (p.N + ichunk - 1) / ichunk);
        keyswitch_mklwe_generic<<<grid, 256, 0, ctx->stream>>>(ctx->d, ctx->ws_ms, ctx->kskA, ctx->kskB, ctx->ws_raw, braw, ichunk);
        ctx->launches++;
        CU(cudaGetLastError());
        const uint64_t total = uint64_t(batch) * p.k * p.n;
        keyswitch_mklwe_finish<<<unsigned((std::max<uint64_t>(total, batch) + 255) / 256), 256, 0, ctx->stream>>>(
            ctx->d, ctx->ws_raw, braw, d_out, d_bout, total, B);
        ctx->launches++;
        CU(cudaGetLastError());
    }
    CU(cudaEventRecord(ctx->ev[2], ctx->stream));
    ctx->ev_valid = true;
    return MKFHE_OK;
}

int mkfhe_last_gate_timing(mkfhe_ctx* ctx, float* rotate_ms, float* keyswitch_ms) {
    if (!ctx || !rotate_ms || !keyswitch_ms) return MKFHE_EINVAL;
    if (!ctx->ev_valid) return fail(ctx, MKFHE_EINVAL, "no gate batch has been evaluated yet");
    CU(cudaEventSynchronize(ctx->ev[2]));
    CU(cudaEventElapsedTime(rotate_ms, ctx->ev[0], ctx->ev[1]));
    CU(cudaEventElapsedTime(keyswitch_ms, ctx->ev[1], ctx->ev[2]));
    return MKFHE_OK;
}

int mkfhe_eval_nand(mkfhe_ctx* ctx, const uint32_t* ct1, const uint32_t* b1, const uint32_t* ct2, const uint32_t* b2, uint32_t* out,
                    uint32_t* bout, uint64_t batch) {
    if (!ctx || !ct1 || !ct2 || !out) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    const bool lwe = p.scheme == MKFHE_SCHEME_MKLWE;
    if (lwe && (!b1 || !b2 || !bout)) return fail(ctx, MKFHE_EINVAL, "MKLWE gates need b1, b2 and bout");
    if (batch == 0) return MKFHE_OK;
    CU(cudaSetDevice(ctx->device));
    const size_t cw = size_t(batch) * p.k * p.n;
    int rc = grow(ctx, &ctx->ws_io, &ctx->ws_io_words, 3 * cw + 3 * batch);
    if (rc) return rc;
    uint32_t *d1 = ctx->ws_io, *d2 = d1 + cw, *dout = d2 + cw, *db1 = dout + cw, *db2 = db1 + batch, *dbo = db2 + batch;
    CU(cudaMemcpyAsync(d1, ct1, cw * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(d2, ct2, cw * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (lwe) {
        CU(cudaMemcpyAsync(db1, b1, batch * 4, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(db2, b2, batch * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    rc = mkfhe_eval_nand_dev(ctx, d1, lwe ? db1 : nullptr, d2, lwe ? db2 : nullptr, dout, lwe ? dbo : nullptr, batch);
    if (rc) return rc;
    CU(cudaMemcpyAsync(out, dout, cw * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (lwe) CU(cudaMemcpyAsync(bout, dbo, batch * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}

// ---- circuits ----

int64_t mkfhe_circuit_levels(uint64_t n_wires, const uint32_t* in1, const uint32_t* in2, const uint32_t* out, uint64_t n_gates,
                             uint32_t* level_of_gate) {
    if ((!in1 || !in2 || !out) && n_gates) return MKFHE_EINVAL;
    std::vector<uint32_t> level;
    const char* why = "";
    const int64_t L = circuit_levelise(n_wires, in1, in2, out, n_gates, level, &why);
    if (L < 0) return MKFHE_EINVAL;
    if (level_of_gate) std::copy(level.begin(), level.end(), level_of_gate);
    return L;
}

int mkfhe_eval_circuit_dev(mkfhe_ctx* ctx, uint32_t* d_wires, uint32_t* d_wires_b, uint64_t n_wires, const uint32_t* in1,
                           const uint32_t* in2, const uint32_t* out, uint64_t n_gates, uint32_t* n_levels) {
    if (!ctx || !d_wires || ((!in1 || !in2 || !out) && n_gates)) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    const bool lwe = p.scheme == MKFHE_SCHEME_MKLWE;
    if (lwe && !d_wires_b) return fail(ctx, MKFHE_EINVAL, "MKLWE circuits need wires_b");
    if (n_wires > 0xffffffffull || n_gates > 0x7fffffffull) return fail(ctx, MKFHE_EINVAL, "circuit too large");
    std::vector<uint32_t> level;
    const char* why = "";
    const int64_t L = circuit_levelise(n_wires, in1, in2, out, n_gates, level, &why);
    if (L < 0) return fail(ctx, MKFHE_EINVAL, "mkfhe_eval_circuit: %s", why);
    if (n_levels) *n_levels = uint32_t(L);
    if (n_gates == 0) return MKFHE_OK;
    // counting sort of the gates by level (stable): start[l] .. start[l + 1] are the gates of level l
    std::vector<uint64_t> start(size_t(L) + 1, 0);
    for (uint64_t g = 0; g < n_gates; ++g) start[level[g] + 1]++;
    uint64_t widest = 0;
    for (int64_t l = 0; l < L; ++l) {
        widest = std::max(widest, start[l + 1]);
        start[l + 1] += start[l];
    }
    std::vector<uint32_t> idx(3 * n_gates);  // [in1 | in2 | out], each sorted by level
    {
        std::vector<uint64_t> at(start.begin(), start.end() - 1);
        for (uint64_t g = 0; g < n_gates; ++g) {
            const uint64_t pos = at[level[g]]++;
            idx[pos] = in1[g];
            idx[n_gates + pos] = in2[g];
            idx[2 * n_gates + pos] = out[g];
        }
    }
    CU(cudaSetDevice(ctx->device));
    const uint32_t kn = p.k * p.n;
    TmpBuf tmp(ctx);
    uint32_t *d_idx, *c1, *c2, *co, *b1 = nullptr, *b2 = nullptr, *bo = nullptr;
    CU(tmp.get(&d_idx, idx.size()));
    CU(tmp.get(&c1, widest * kn));
    CU(tmp.get(&c2, widest * kn));
    CU(tmp.get(&co, widest * kn));
    if (lwe) {
        CU(tmp.get(&b1, widest));
        CU(tmp.get(&b2, widest));
        CU(tmp.get(&bo, widest));
    }
    // pageable source: the copy is staged before cudaMemcpyAsync returns, so `idx` may go out of scope
    CU(cudaMemcpyAsync(d_idx, idx.data(), idx.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    for (int64_t l = 0; l < L; ++l) {
        const uint64_t s = start[l], B = start[l + 1] - s;
        circuit_gather<<<unsigned(B), 256, 0, ctx->stream>>>(d_wires, lwe ? d_wires_b : nullptr, d_idx + s, d_idx + n_gates + s, c1, c2, b1,
                                                           b2, kn);
        CU(cudaGetLastError());
        const int rc = mkfhe_eval_nand_dev(ctx, c1, b1, c2, b2, co, bo, B);
        if (rc) return rc;
        circuit_scatter<<<unsigned(B), 256, 0, ctx->stream>>>(d_wires, lwe ? d_wires_b : nullptr, d_idx + 2 * n_gates + s, co, bo, kn);
        CU(cudaGetLastError());
        ctx->launches += 2;
    }
    return MKFHE_OK;  // scratch returns to the pool in stream order (TmpBuf)
}

int mkfhe_eval_circuit(mkfhe_ctx* ctx, uint32_t* wires, uint32_t* wires_b, uint64_t n_wires, const uint32_t* in1, const uint32_t* in2,
                       const uint32_t* out, uint64_t n_gates, uint32_t* n_levels) {
    if (!ctx || !wires) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    const bool lwe = p.scheme == MKFHE_SCHEME_MKLWE;
    if (lwe && !wires_b) return fail(ctx, MKFHE_EINVAL, "MKLWE circuits need wires_b");
    CU(cudaSetDevice(ctx->device));
    const size_t words = size_t(n_wires) * p.k * p.n;
    TmpBuf tmp(ctx);
    uint32_t *d_w, *d_b = nullptr;
    CU(tmp.get(&d_w, words));
    if (lwe) CU(tmp.get(&d_b, n_wires));
    CU(cudaMemcpyAsync(d_w, wires, words * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (lwe) CU(cudaMemcpyAsync(d_b, wires_b, n_wires * 4, cudaMemcpyHostToDevice, ctx->stream));
    const int rc = mkfhe_eval_circuit_dev(ctx, d_w, d_b, n_wires, in1, in2, out, n_gates, n_levels);
    if (rc) return rc;
    CU(cudaMemcpyAsync(wires, d_w, words * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (lwe) CU(cudaMemcpyAsync(wires_b, d_b, n_wires * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}

// ---- circuits without device allocation (no counterpart in the reference) ----
// The level batches of mkfhe_eval_circuit_dev (sorted index arrays, gathered inputs, level outputs) live in ONE file-scope
// device array instead of the context's pool, for callers that may not allocate: CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc.  Every GPU that loads the library holds the array (8 MiB) until the process exits.
// g_static_scratch is a host-side stand-in in host code, not a device address: the kernels below are given the pointer
// cudaGetSymbolAddress returns for the context's device (looked up once per device, with that device current).
static __device__ __align__(256) uint32_t g_static_scratch[MKFHE_STATIC_SCRATCH_WORDS];

static int static_scratch_ptr(mkfhe_ctx* ctx, uint32_t** out) {
    constexpr int MAXDEV = 64;
    static std::mutex mu;
    static uint32_t* by_device[MAXDEV] = {};
    if (ctx->device < 0 || ctx->device >= MAXDEV) return fail(ctx, MKFHE_EINVAL, "device ordinal %d out of range", ctx->device);
    std::lock_guard<std::mutex> lk(mu);
    if (!by_device[ctx->device]) {
        void* p = nullptr;
        CU(cudaSetDevice(ctx->device));
        CU(cudaGetSymbolAddress(&p, g_static_scratch));
        by_device[ctx->device] = static_cast<uint32_t*>(p);
    }
    *out = by_device[ctx->device];
    return MKFHE_OK;
}

uint64_t mkfhe_static_scratch_words(void) { return MKFHE_STATIC_SCRATCH_WORDS; }

uint64_t mkfhe_circuit_static_words(const mkfhe_params* p, uint64_t n_gates, uint64_t widest_level) {
    if (!p) return 0;
    auto pad = [](uint64_t w) { return (w + 63) & ~uint64_t(63); };  // every sub-buffer starts on a 256-byte boundary
    const uint64_t kn = uint64_t(p->k) * p->n;
    uint64_t w = pad(3 * n_gates) + 3 * pad(widest_level * kn);
    if (p->scheme == MKFHE_SCHEME_MKLWE) w += 3 * pad(widest_level);
    return w;
}

int mkfhe_eval_circuit_dev_static(mkfhe_ctx* ctx, uint32_t* d_wires, uint32_t* d_wires_b, uint64_t n_wires, const uint32_t* in1,
                                  const uint32_t* in2, const uint32_t* out, uint64_t n_gates, uint32_t* n_levels) {
    if (!ctx || !d_wires || ((!in1 || !in2 || !out) && n_gates)) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    const bool lwe = p.scheme == MKFHE_SCHEME_MKLWE;
    if (lwe && !d_wires_b) return fail(ctx, MKFHE_EINVAL, "MKLWE circuits need wires_b");
    if (!ctx->fuse_ks)
        return fail(ctx, MKFHE_EINVAL, "mkfhe_eval_circuit_dev_static needs the fused key-switch path (qKS < 2^16); the unfused path "
                                       "grows workspaces with cudaMalloc");
    if (n_wires > 0xffffffffull || n_gates > 0x7fffffffull) return fail(ctx, MKFHE_EINVAL, "circuit too large");
    std::vector<uint32_t> level;
    const char* why = "";
    const int64_t L = circuit_levelise(n_wires, in1, in2, out, n_gates, level, &why);
    if (L < 0) return fail(ctx, MKFHE_EINVAL, "mkfhe_eval_circuit_dev_static: %s", why);
    if (n_levels) *n_levels = uint32_t(L);
    if (n_gates == 0) return MKFHE_OK;
    std::vector<uint64_t> start(size_t(L) + 1, 0);
    for (uint64_t g = 0; g < n_gates; ++g) start[level[g] + 1]++;
    uint64_t widest = 0;
    for (int64_t l = 0; l < L; ++l) {
        widest = std::max(widest, start[l + 1]);
        start[l + 1] += start[l];
    }
    const uint64_t need = mkfhe_circuit_static_words(&p, n_gates, widest);
    if (need > MKFHE_STATIC_SCRATCH_WORDS)
        return fail(ctx, MKFHE_EINVAL, "mkfhe_eval_circuit_dev_static: %llu gates with a widest level of %llu need %llu words of scratch, "
                                       "the static array holds %llu (use mkfhe_eval_circuit_dev)",
                    (unsigned long long)n_gates, (unsigned long long)widest, (unsigned long long)need,
                    (unsigned long long)MKFHE_STATIC_SCRATCH_WORDS);
    std::vector<uint32_t> idx(3 * n_gates);  // [in1 | in2 | out], each sorted by level (host memory only)
    {
        std::vector<uint64_t> at(start.begin(), start.end() - 1);
        for (uint64_t g = 0; g < n_gates; ++g) {
            const uint64_t pos = at[level[g]]++;
            idx[pos] = in1[g];
            idx[n_gates + pos] = in2[g];
            idx[2 * n_gates + pos] = out[g];
        }
    }
    CU(cudaSetDevice(ctx->device));
    uint32_t* base = nullptr;
    int rc = static_scratch_ptr(ctx, &base);
    if (rc) return rc;
    auto pad = [](uint64_t w) { return (w + 63) & ~uint64_t(63); };
    const uint32_t kn = p.k * p.n;
    // nothing in the array has to start at zero: the index block is overwritten by the copy below and every row a level reads
    // from c1 / c2 / co (and b1 / b2 / bo) is written by that level's gather or gate kernel first
    uint32_t* d_idx = base;
    uint32_t* c1 = d_idx + pad(3 * n_gates);
    uint32_t* c2 = c1 + pad(widest * kn);
    uint32_t* co = c2 + pad(widest * kn);
    uint32_t *b1 = nullptr, *b2 = nullptr, *bo = nullptr;
    if (lwe) {
        b1 = co + pad(widest * kn);
        b2 = b1 + pad(widest);
        bo = b2 + pad(widest);
    }
    CU(cudaMemcpyAsync(d_idx, idx.data(), idx.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    for (int64_t l = 0; l < L; ++l) {
        const uint64_t s = start[l], B = start[l + 1] - s;
        circuit_gather<<<unsigned(B), 256, 0, ctx->stream>>>(d_wires, lwe ? d_wires_b : nullptr, d_idx + s, d_idx + n_gates + s, c1, c2, b1,
                                                           b2, kn);
        CU(cudaGetLastError());
        rc = mkfhe_eval_nand_dev(ctx, c1, b1, c2, b2, co, bo, B);
        if (rc) return rc;
        circuit_scatter<<<unsigned(B), 256, 0, ctx->stream>>>(d_wires, lwe ? d_wires_b : nullptr, d_idx + 2 * n_gates + s, co, bo, kn);
        CU(cudaGetLastError());
        ctx->launches += 2;
    }
    return MKFHE_OK;
}

int mkfhe_ntt_dev(mkfhe_ctx* ctx, uint32_t* d_polys, uint64_t batch, int inverse) {
    if (!ctx || !d_polys) return MKFHE_EINVAL;
    if (batch == 0) return MKFHE_OK;
    if (batch > 0x7fffffffull) return fail(ctx, MKFHE_EINVAL, "batch too large");
    CU(cudaSetDevice(ctx->device));
    if (ctx->use_fast) {
        cudaError_t e = fast_ntt(ctx->fast, ctx->d, d_polys, batch, inverse, ctx->stream, &ctx->launches);
        if (e != cudaSuccess) return fail(ctx, MKFHE_ECUDA, "fast_ntt: %s", cudaGetErrorString(e));
        return MKFHE_OK;
    }
    if (ctx->ntt_half) {  // pairs on the specialised kernel, an odd last polynomial on the generic one
        if (batch / 2) {
            cudaError_t e = fast_ntt(ctx->fast, ctx->d, d_polys, batch / 2, inverse, ctx->stream, &ctx->launches);
            if (e != cudaSuccess) return fail(ctx, MKFHE_ECUDA, "fast_ntt: %s", cudaGetErrorString(e));
        }
        if ((batch & 1) == 0) return MKFHE_OK;
        d_polys += (batch - 1) * ctx->p.N;
        batch = 1;
    }
    const unsigned threads = std::max(32u, std::min(1024u, ctx->p.N / 2));
    ntt_batch_generic<<<unsigned(batch), threads, ctx->p.N * 4, ctx->stream>>>(ctx->d, d_polys, inverse);
    ctx->launches++;
    CU(cudaGetLastError());
    return MKFHE_OK;
}
int mkfhe_ntt(mkfhe_ctx* ctx, uint32_t* polys, uint64_t batch, int inverse) {
    if (!ctx || !polys) return MKFHE_EINVAL;
    if (batch == 0) return MKFHE_OK;
    CU(cudaSetDevice(ctx->device));
    const size_t w = size_t(batch) * ctx->p.N;
    int rc = grow(ctx, &ctx->ws_io, &ctx->ws_io_words, w);
    if (rc) return rc;
    CU(cudaMemcpyAsync(ctx->ws_io, polys, w * 4, cudaMemcpyHostToDevice, ctx->stream));
    rc = mkfhe_ntt_dev(ctx, ctx->ws_io, batch, inverse);
    if (rc) return rc;
    CU(cudaMemcpyAsync(polys, ctx->ws_io, w * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}

int mkfhe_decompose(mkfhe_ctx* ctx, const uint32_t* polys, uint32_t* out, uint64_t batch) {
    if (!ctx || !polys || !out) return MKFHE_EINVAL;
    if (batch == 0) return MKFHE_OK;
    CU(cudaSetDevice(ctx->device));
    const size_t w = size_t(batch) * ctx->p.N, ell = ctx->p.digitsG - 1;
    int rc = grow(ctx, &ctx->ws_io, &ctx->ws_io_words, w * (1 + ell));
    if (rc) return rc;
    CU(cudaMemcpyAsync(ctx->ws_io, polys, w * 4, cudaMemcpyHostToDevice, ctx->stream));
    decompose_generic<<<unsigned((w + 255) / 256), 256, 0, ctx->stream>>>(ctx->d, ctx->ws_io, ctx->ws_io + w, w);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out, ctx->ws_io + w, w * ell * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}

// ---- key generation with caller-supplied randomness ----

int mkfhe_keygen_bk(mkfhe_ctx* ctx, const uint32_t* sinv, const uint32_t* crs, const uint32_t* r, const uint32_t* e, const uint8_t* m,
                    uint32_t* bk_out, int load) {
    if (!ctx || !sinv || !crs || !r || !e || !m || (!bk_out && !load)) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    const uint32_t N = p.N, ell = p.digitsG - 1, Q = p.Q;
    if (ell > 8) return fail(ctx, MKFHE_EINVAL, "mkfhe_keygen_bk: at most 8 gadget digits");
    const size_t S = p.variant == MKFHE_VARIANT_XZW ? 2 : 1, E = size_t(p.k) * S * p.n + 1, words = bk_words(p);
    const size_t smem = size_t(1 + 2 * ell) * N * 4;
    CU(cudaSetDevice(ctx->device));
    TmpBuf tmp(ctx);
    uint32_t *d_sinv, *d_crs, *d_r, *d_e, *d_out;
    uint8_t* d_m;
    CU(tmp.get(&d_sinv, size_t(p.k) * N));
    CU(tmp.get(&d_crs, size_t(ell) * N));
    CU(tmp.get(&d_r, E * N));
    CU(tmp.get(&d_e, words));
    CU(tmp.get(&d_out, words));
    CU(tmp.get(&d_m, E));
    CU(cudaMemcpyAsync(d_sinv, sinv, size_t(p.k) * N * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_crs, crs, size_t(ell) * N * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_r, r, E * N * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_e, e, words * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_m, m, E, cudaMemcpyDefault, ctx->stream));
    to_montgomery<<<unsigned((size_t(p.k) * N + 255) / 256), 256, 0, ctx->stream>>>(d_sinv, d_sinv, size_t(p.k) * N, ctx->d.mod, ctx->r2);
    to_montgomery<<<unsigned((size_t(ell) * N + 255) / 256), 256, 0, ctx->stream>>>(d_crs, d_crs, size_t(ell) * N, ctx->d.mod, ctx->r2);
    GadgetPow G{};
    const uint32_t R = uint32_t((uint64_t(1) << 32) % Q);
    uint32_t g = 1;
    for (uint32_t j = 0; j < ell; ++j) {
        g = h_mulmod(g, (1u << p.baseG_log2) % Q, Q);
        G.plain[j] = g;
        G.mont[j] = h_mulmod(g, R, Q);
    }
    if (ctx->use_fast) {  // N = 2048: register-resident transforms
        cudaError_t ce = fast_keygen_bk(ctx->fast, ctx->d, G, d_sinv, d_crs, d_r, d_e, d_m, d_out, unsigned(E), ctx->stream, &ctx->launches);
        if (ce != cudaSuccess) return fail(ctx, MKFHE_ECUDA, "fast_keygen_bk: %s", cudaGetErrorString(ce));
    } else {
        CU(cudaFuncSetAttribute(keygen_bk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        keygen_bk_kernel<<<unsigned(E), 256, smem, ctx->stream>>>(ctx->d, G, d_sinv, d_crs, d_r, d_e, d_m, d_out);
        ctx->launches++;
    }
    ctx->launches += 2;
    CU(cudaGetLastError());
    if (bk_out) CU(cudaMemcpyAsync(bk_out, d_out, words * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (load) return mkfhe_load_bk(ctx, d_out, words);  // device source: converted and re-laid out on the device
    return MKFHE_OK;
}

int mkfhe_keygen_pkey(mkfhe_ctx* ctx, const uint32_t* s, const uint32_t* crs, const uint32_t* e, uint32_t* pkey_out, int load) {
    if (!ctx || !s || !crs || !e || (!pkey_out && !load)) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    const uint32_t N = p.N, ell = p.digitsG - 1;
    const size_t words = size_t(p.k) * ell * N, smem = size_t(ell) * N * 4;
    CU(cudaSetDevice(ctx->device));
    TmpBuf tmp(ctx);
    uint32_t *d_s, *d_crs, *d_e, *d_out;
    CU(tmp.get(&d_s, size_t(p.k) * N));
    CU(tmp.get(&d_crs, size_t(ell) * N));
    CU(tmp.get(&d_e, words));
    CU(tmp.get(&d_out, words));
    CU(cudaMemcpyAsync(d_s, s, size_t(p.k) * N * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_crs, crs, size_t(ell) * N * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_e, e, words * 4, cudaMemcpyDefault, ctx->stream));
    to_montgomery<<<unsigned((size_t(p.k) * N + 255) / 256), 256, 0, ctx->stream>>>(d_s, d_s, size_t(p.k) * N, ctx->d.mod, ctx->r2);
    to_montgomery<<<unsigned((size_t(ell) * N + 255) / 256), 256, 0, ctx->stream>>>(d_crs, d_crs, size_t(ell) * N, ctx->d.mod, ctx->r2);
    CU(cudaFuncSetAttribute(keygen_pkey_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    keygen_pkey_kernel<<<p.k, 256, smem, ctx->stream>>>(ctx->d, d_s, d_crs, d_e, d_out);
    ctx->launches += 3;
    CU(cudaGetLastError());
    if (pkey_out) CU(cudaMemcpyAsync(pkey_out, d_out, words * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (load) return mkfhe_load_pkey(ctx, d_out, words);
    return MKFHE_OK;
}

int mkfhe_keygen_ksk_mntru(mkfhe_ctx* ctx, const uint32_t* s_q, const uint32_t* e, const uint32_t* finv, uint32_t* ksk_out, int load) {
    if (!ctx || !s_q || !e || !finv || (!ksk_out && !load)) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    if (p.scheme != MKFHE_SCHEME_MNTRU) return fail(ctx, MKFHE_EINVAL, "context is not MNTRU");
    const size_t L = size_t(p.N) * p.dKS, words = size_t(p.k) * L * p.n;
    CU(cudaSetDevice(ctx->device));
    TmpBuf tmp(ctx);
    uint32_t *d_s, *d_e, *d_f, *d_out;
    CU(tmp.get(&d_s, size_t(p.k) * p.N));
    CU(tmp.get(&d_e, words));
    CU(tmp.get(&d_f, size_t(p.k) * p.n * p.n));
    CU(tmp.get(&d_out, words));
    CU(cudaMemcpyAsync(d_s, s_q, size_t(p.k) * p.N * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_e, e, words * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_f, finv, size_t(p.k) * p.n * p.n * 4, cudaMemcpyDefault, ctx->stream));
    const dim3 grid((p.n + 63) / 64, unsigned((L + 63) / 64), p.k);
    keygen_ksk_mntru_kernel<<<grid, 256, 0, ctx->stream>>>(ctx->d, d_s, d_e, d_f, d_out);
    ctx->launches++;
    CU(cudaGetLastError());
    if (ksk_out) CU(cudaMemcpyAsync(ksk_out, d_out, words * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (load) return mkfhe_load_ksk_mntru(ctx, d_out, words);
    return MKFHE_OK;
}

int mkfhe_keygen_ksk_mklwe(mkfhe_ctx* ctx, const uint32_t* A, const uint32_t* e, const uint32_t* s, const uint32_t* sN_q,
                           uint32_t* B_out, int load) {
    if (!ctx || !A || !e || !s || !sN_q || (!B_out && !load)) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    if (p.scheme != MKFHE_SCHEME_MKLWE) return fail(ctx, MKFHE_EINVAL, "context is not MKLWE");
    const size_t rows = size_t(p.k) * p.N * p.baseKS * p.dKS, wordsA = rows * p.n;
    CU(cudaSetDevice(ctx->device));
    TmpBuf tmp(ctx);
    uint32_t *d_A, *d_e, *d_s, *d_sN, *d_B;
    CU(tmp.get(&d_A, wordsA));
    CU(tmp.get(&d_e, rows));
    CU(tmp.get(&d_s, size_t(p.k) * p.n));
    CU(tmp.get(&d_sN, size_t(p.k) * p.N));
    CU(tmp.get(&d_B, rows));
    CU(cudaMemcpyAsync(d_A, A, wordsA * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_e, e, rows * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_s, s, size_t(p.k) * p.n * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_sN, sN_q, size_t(p.k) * p.N * 4, cudaMemcpyDefault, ctx->stream));
    keygen_ksk_mklwe_kernel<<<unsigned((rows * 32 + 255) / 256), 256, 0, ctx->stream>>>(ctx->d, d_A, d_e, d_s, d_sN, d_B, rows);
    ctx->launches++;
    CU(cudaGetLastError());
    if (B_out) CU(cudaMemcpyAsync(B_out, d_B, rows * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (load) return mkfhe_load_ksk_mklwe(ctx, d_A, wordsA, d_B, rows);
    return MKFHE_OK;
}

// ---- batched Encrypt with caller-supplied noise ----
int mkfhe_encrypt_mntru(mkfhe_ctx* ctx, const uint32_t* finv, const uint32_t* e, const uint8_t* m, uint32_t pmod, uint32_t* ct_out,
                        uint64_t batch) {
    if (!ctx || !finv || !e || !m || !ct_out || pmod == 0) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    if (p.scheme != MKFHE_SCHEME_MNTRU) return fail(ctx, MKFHE_EINVAL, "context is not MNTRU");
    if (batch == 0) return MKFHE_OK;
    if (batch > 0x3fffffull) return fail(ctx, MKFHE_EINVAL, "batch too large");
    const size_t words = size_t(batch) * p.k * p.n;
    CU(cudaSetDevice(ctx->device));
    TmpBuf tmp(ctx);
    uint32_t *d_f, *d_e, *d_out;
    uint8_t* d_m;
    CU(tmp.get(&d_f, size_t(p.k) * p.n * p.n));
    CU(tmp.get(&d_e, words));
    CU(tmp.get(&d_out, words));
    CU(tmp.get(&d_m, size_t(batch)));
    CU(cudaMemcpyAsync(d_f, finv, size_t(p.k) * p.n * p.n * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_e, e, words * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_m, m, size_t(batch), cudaMemcpyDefault, ctx->stream));
    const dim3 grid((p.n + 63) / 64, unsigned((batch + 63) / 64), p.k);
    encrypt_mntru_kernel<<<grid, 256, 0, ctx->stream>>>(ctx->d, d_f, d_e, d_m, pmod, d_out, uint32_t(batch));
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(ct_out, d_out, words * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}

int mkfhe_encrypt_mklwe(mkfhe_ctx* ctx, const uint32_t* s, const uint32_t* a, const uint32_t* err, const uint8_t* m, uint32_t pmod,
                        uint32_t* b_out, uint64_t batch) {
    if (!ctx || !s || !a || !err || !m || !b_out || pmod == 0) return MKFHE_EINVAL;
    const mkfhe_params& p = ctx->p;
    if (p.scheme != MKFHE_SCHEME_MKLWE) return fail(ctx, MKFHE_EINVAL, "context is not MKLWE");
    if (batch == 0) return MKFHE_OK;
    const size_t words = size_t(batch) * p.k * p.n;
    CU(cudaSetDevice(ctx->device));
    TmpBuf tmp(ctx);
    uint32_t *d_s, *d_a, *d_err, *d_b;
    uint8_t* d_m;
    CU(tmp.get(&d_s, size_t(p.k) * p.n));
    CU(tmp.get(&d_a, words));
    CU(tmp.get(&d_err, size_t(batch)));
    CU(tmp.get(&d_b, size_t(batch)));
    CU(tmp.get(&d_m, size_t(batch)));
    CU(cudaMemcpyAsync(d_s, s, size_t(p.k) * p.n * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_a, a, words * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_err, err, size_t(batch) * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaMemcpyAsync(d_m, m, size_t(batch), cudaMemcpyDefault, ctx->stream));
    encrypt_mklwe_kernel<<<unsigned((batch * 32 + 255) / 256), 256, 0, ctx->stream>>>(ctx->d, d_s, d_a, d_err, d_m, pmod, d_b, batch);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(b_out, d_b, size_t(batch) * 4, cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return MKFHE_OK;
}

// ---- binary key file (format in keyfile.h) ----
int mkfhe_keyfile_write(const char* path, const mkfhe_params* p, const uint32_t* bk, uint64_t bk_words, const uint32_t* pkey,
                        uint64_t pkey_words, const uint32_t* kskA, uint64_t kskA_words, const uint32_t* kskB, uint64_t kskB_words,
                        const uint32_t* ctnand, uint64_t ctnand_words) {
    if (!path || !p) return MKFHE_EINVAL;
    using namespace mkfhe_keyfile;
    Writer w;
    if (!w.open(path, *p)) return MKFHE_EINVAL;
    const bool lwe = p->scheme == MKFHE_SCHEME_MKLWE;
    bool ok = w.section(BK, bk, bk_words) && w.section(PKEY, pkey, pkey_words) && w.section(lwe ? KSKA : KSK, kskA, kskA_words) &&
              (!lwe || w.section(KSKB, kskB, kskB_words)) && (lwe || w.section(CTNAND, ctnand, ctnand_words));
    ok = w.close() && ok;
    return ok ? MKFHE_OK : MKFHE_EINVAL;
}

int mkfhe_keyfile_info(const char* path, mkfhe_params* p, uint64_t* words_by_tag /* [7], index = section tag */) {
    if (!path) return MKFHE_EINVAL;
    return guarded_io(nullptr, [&]() -> int {
        uint64_t w[mkfhe_keyfile::TAG_SLOTS];
        const int rc = mkfhe_keyfile::read(path, mkfhe_keyfile::Expect{mkfhe_keyfile::MAGIC, nullptr}, p, &w,
                                           [](uint32_t, const uint32_t*, uint64_t) { return 0; });
        if (words_by_tag) memcpy(words_by_tag, w, sizeof(uint64_t) * (mkfhe_keyfile::TAG_MAX + 1));
        return rc ? MKFHE_EINVAL : MKFHE_OK;
    });
}

int mkfhe_keyfile_load(mkfhe_ctx* ctx, const char* path) {
    if (!ctx || !path) return MKFHE_EINVAL;
    using namespace mkfhe_keyfile;
    return guarded_io(ctx, [&]() -> int {
        // the header first: the file must describe this context's parameter set before any section is looked at
        mkfhe_params fp{};
        {
            FILE* f = fopen(path, "rb");
            Header h{};
            const bool ok = f && fread(&h, sizeof h, 1, f) == 1;
            if (f) fclose(f);
            if (!ok) return fail(ctx, MKFHE_EINVAL, "mkfhe_keyfile_load: cannot read %s", path);
            if (memcmp(h.magic, MAGIC, 8) != 0 || h.version != VERSION)
                return fail(ctx, MKFHE_EINVAL, "mkfhe_keyfile_load: %s is not a version-%u key file", path, VERSION);
            words_to_params(h.par, fp);
            uint32_t a[13], b[13];
            params_to_words(fp, a);
            params_to_words(ctx->p, b);
            if (memcmp(a, b, sizeof a) != 0)
                return fail(ctx, MKFHE_EINVAL, "mkfhe_keyfile_load: %s was written for a different parameter set", path);
        }
        // every section present must have exactly the size this parameter set implies (absent sections stay 0 = not expected)
        uint64_t want[TAG_SLOTS];
        expected_key_words(ctx->p, want);
        std::vector<uint32_t> kskA;  // MKLWE: A is held until B arrives (mkfhe_load_ksk_mklwe takes both)
        int inner = MKFHE_OK;
        // read() verifies the whole section table and every checksum BEFORE the first upload below, so a bad file leaves the
        // context's keys untouched
        const int rc = read(path, Expect{MAGIC, want}, nullptr, nullptr, [&](uint32_t tag, const uint32_t* data, uint64_t words) -> int {
            switch (tag) {
                case BK: inner = mkfhe_load_bk(ctx, data, words); break;
                case PKEY: inner = mkfhe_load_pkey(ctx, data, words); break;
                case KSK: inner = mkfhe_load_ksk_mntru(ctx, data, words); break;
                case KSKA: kskA.assign(data, data + words); break;
                case KSKB: inner = kskA.empty() ? MKFHE_EINVAL : mkfhe_load_ksk_mklwe(ctx, kskA.data(), kskA.size(), data, words); break;
                case CTNAND: inner = mkfhe_load_ctnand(ctx, data, words); break;
                default: return -4;
            }
            return inner ? -6 : 0;
        });
        switch (rc) {
            case 0: return MKFHE_OK;
            case -1: return fail(ctx, MKFHE_EINVAL, "mkfhe_keyfile_load: cannot read %s", path);
            case -2: return fail(ctx, MKFHE_EINVAL, "mkfhe_keyfile_load: %s is not a version-%u key file", path, VERSION);
            case -3: return fail(ctx, MKFHE_EINVAL, "mkfhe_keyfile_load: checksum mismatch in %s", path);
            case -6: return inner;  // mkfhe_load_* already recorded the reason
            default: return fail(ctx, MKFHE_EINVAL, "mkfhe_keyfile_load: malformed section table in %s (unknown, duplicate or mis-sized section, "
                                 "kskA without kskB, or data beyond the end of the file)", path);
        }
    });
}

// ---- ciphertext / wire-table files (no GPU involved) ----
int mkfhe_ctfile_write(const char* path, const mkfhe_params* p, const uint32_t* ct, const uint32_t* b, uint64_t n_ct) {
    if (!path || !p || !ct || n_ct == 0) return MKFHE_EINVAL;
    const bool lwe = p->scheme == MKFHE_SCHEME_MKLWE;
    if (lwe && !b) return MKFHE_EINVAL;
    using namespace mkfhe_keyfile;
    Writer w;
    if (!w.open(path, *p, MAGIC_CT)) return MKFHE_EINVAL;
    bool ok = w.section(CT_A, ct, n_ct * p->k * p->n) && (!lwe || w.section(CT_B, b, n_ct));
    ok = w.close() && ok;
    return ok ? MKFHE_OK : MKFHE_EINVAL;
}

int mkfhe_ctfile_info(const char* path, mkfhe_params* p, uint64_t* n_ct) {
    if (!path) return MKFHE_EINVAL;
    return guarded_io(nullptr, [&]() -> int {
        using namespace mkfhe_keyfile;
        mkfhe_params fp{};
        uint64_t w[TAG_SLOTS];
        if (read(path, Expect{MAGIC_CT, nullptr}, &fp, &w, [](uint32_t, const uint32_t*, uint64_t) { return 0; })) return MKFHE_EINVAL;
        const uint64_t kn = uint64_t(fp.k) * fp.n;
        if (kn == 0 || w[CT_A] == 0 || w[CT_A] % kn) return MKFHE_EINVAL;
        const uint64_t n = w[CT_A] / kn;
        if ((fp.scheme == MKFHE_SCHEME_MKLWE) != (w[CT_B] != 0) || (w[CT_B] && w[CT_B] != n)) return MKFHE_EINVAL;
        if (p) *p = fp;
        if (n_ct) *n_ct = n;
        return MKFHE_OK;
    });
}

int mkfhe_ctfile_read(const char* path, const mkfhe_params* expect, uint32_t* ct, uint32_t* b, uint64_t capacity_ct, uint64_t* n_ct) {
    if (!path || !ct) return MKFHE_EINVAL;
    return guarded_io(nullptr, [&]() -> int {
        using namespace mkfhe_keyfile;
        mkfhe_params fp{};
        uint64_t n = 0;
        int rc = mkfhe_ctfile_info(path, &fp, &n);
        if (rc) return rc;
        if (expect) {
            uint32_t a[13], e[13];
            params_to_words(fp, a);
            params_to_words(*expect, e);
            if (memcmp(a, e, sizeof a) != 0) return MKFHE_EINVAL;
        }
        const bool lwe = fp.scheme == MKFHE_SCHEME_MKLWE;
        if (n > capacity_ct || (lwe && !b)) return MKFHE_EINVAL;
        rc = read(path, Expect{MAGIC_CT, nullptr}, nullptr, nullptr, [&](uint32_t tag, const uint32_t* data, uint64_t words) -> int {
            if (tag == CT_A) memcpy(ct, data, words * 4);
            else if (tag == CT_B) memcpy(b, data, words * 4);
            else return -4;
            return 0;
        });
        if (rc) return MKFHE_EINVAL;
        if (n_ct) *n_ct = n;
        return MKFHE_OK;
    });
}

int mkfhe_measure_butterfly_peak(mkfhe_ctx* ctx, double* bfly_per_s) {
    if (!ctx || !bfly_per_s) return MKFHE_EINVAL;
    CU(cudaSetDevice(ctx->device));
    int sms = 0;
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
    uint32_t* sink = nullptr;
    CU(cudaMalloc(&sink, 4));
    const uint32_t iters = 4096, blocks = sms * 4, threads = 256, Q = ctx->p.Q < (1u << 27) ? ctx->p.Q : 134176769u;
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    butterfly_peak_kernel<<<blocks, threads, 0, ctx->stream>>>(sink, 64, Q, 0u, 1u);  // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CU(cudaEventRecord(e0, ctx->stream));
        butterfly_peak_kernel<<<blocks, threads, 0, ctx->stream>>>(sink, iters, Q, 0u, 2u + rep);
        CU(cudaEventRecord(e1, ctx->stream));
        CU(cudaEventSynchronize(e1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        best = std::min(best, ms);
        ctx->launches++;
    }
    CU(cudaGetLastError());
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    *bfly_per_s = double(blocks) * threads * double(iters) * 32 / (best * 1e-3);
    return MKFHE_OK;
}

int mkfhe_measure_int32_peak(mkfhe_ctx* ctx, double* imad_per_s) {
    if (!ctx || !imad_per_s) return MKFHE_EINVAL;
    CU(cudaSetDevice(ctx->device));
    int sms = 0;
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
    uint32_t* sink = nullptr;
    CU(cudaMalloc(&sink, 4));
    const uint32_t iters = 4096, blocks = sms * 8, threads = 256;
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    int32_peak_kernel<<<blocks, threads, 0, ctx->stream>>>(sink, 64, 7u);  // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CU(cudaEventRecord(e0, ctx->stream));
        int32_peak_kernel<<<blocks, threads, 0, ctx->stream>>>(sink, iters, 7u + rep);
        CU(cudaEventRecord(e1, ctx->stream));
        CU(cudaEventSynchronize(e1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        best = std::min(best, ms);
        ctx->launches++;
    }
    CU(cudaGetLastError());
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    *imad_per_s = double(blocks) * threads * double(iters) * 16 * 8 / (best * 1e-3);
    return MKFHE_OK;
}

}  // extern "C"

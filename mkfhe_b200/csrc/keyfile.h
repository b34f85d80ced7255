// keyfile.h -- binary wire format of the flattened keys.  Host code only.
//
// The reference serialises keys through cereal as graphs of shared_ptr<NativePoly> (mk-acckey.h, mk-evalkey.h,
// mntru-keyswitchkey2.h, mklwe-keyswitchkey.h save/load): JSON or portable binary, element by element, with the 32 redundant
// slices of KSK2.  This format stores exactly the C-ABI layouts of include/mkfhe_b200.h, so that a key set is generated once
// (by the reference or by mkfhe_keygen_*) and then read straight into HBM on every GPU of every run.
//
//   header  : "MKFHEKEY" | u32 version (1) | u32 section count | 13 x u32 parameters (mkfhe_params without `device`) | u32 0
//   section : u32 tag | u32 0 | u64 words | u64 FNV-1a-64 of the payload bytes | payload (words x u32 LE) | pad to 16 bytes
//   tags    : 1 bk, 2 pkey, 3 ksk (MNTRU digit-1 slice), 4 kskA, 5 kskB (MKLWE), 6 ctnand (MNTRU)
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "../../include/mkfhe_b200.h"

namespace mkfhe_keyfile {

constexpr char MAGIC[8] = {'M', 'K', 'F', 'H', 'E', 'K', 'E', 'Y'};
constexpr uint32_t VERSION = 1;
enum Tag : uint32_t { BK = 1, PKEY = 2, KSK = 3, KSKA = 4, KSKB = 5, CTNAND = 6, TAG_MAX = 6 };

struct Header {
    char magic[8];
    uint32_t version, nsections;
    uint32_t par[13];
    uint32_t zero;
};
struct Section {
    uint32_t tag, zero;
    uint64_t words, fnv;
};
static_assert(sizeof(Header) == 72 && sizeof(Section) == 24, "packed layouts");

inline uint64_t fnv1a(const void* data, size_t bytes, uint64_t h = 1469598103934665603ull) {
    const unsigned char* p = static_cast<const unsigned char*>(data);
    for (size_t i = 0; i < bytes; ++i) h = (h ^ p[i]) * 1099511628211ull;
    return h;
}
inline void params_to_words(const mkfhe_params& p, uint32_t (&w)[13]) {
    const uint32_t v[13] = {p.k, p.N, p.Q, p.psi, p.q, p.qKS, p.n, p.baseG_log2, p.digitsG, p.baseKS, p.dKS, p.variant, p.scheme};
    memcpy(w, v, sizeof v);
}
inline void words_to_params(const uint32_t (&w)[13], mkfhe_params& p) {
    p.k = w[0]; p.N = w[1]; p.Q = w[2]; p.psi = w[3]; p.q = w[4]; p.qKS = w[5]; p.n = w[6]; p.baseG_log2 = w[7];
    p.digitsG = w[8]; p.baseKS = w[9]; p.dKS = w[10]; p.variant = w[11]; p.scheme = w[12];
}

struct Writer {
    FILE* f{nullptr};
    uint32_t count{0};
    bool open(const char* path, const mkfhe_params& p) {
        f = fopen(path, "wb");
        if (!f) return false;
        Header h{};
        memcpy(h.magic, MAGIC, 8);
        h.version = VERSION;
        params_to_words(p, h.par);
        return fwrite(&h, sizeof h, 1, f) == 1;
    }
    bool section(uint32_t tag, const uint32_t* data, uint64_t words) {
        if (!data || words == 0) return true;  // absent sections are simply not written
        Section s{tag, 0, words, fnv1a(data, words * 4)};
        static const char pad[16] = {0};
        const size_t tail = (16 - (sizeof s + words * 4) % 16) % 16;
        ++count;
        return fwrite(&s, sizeof s, 1, f) == 1 && fwrite(data, 4, words, f) == words && fwrite(pad, 1, tail, f) == tail;
    }
    bool close() {
        bool ok = fseek(f, 12, SEEK_SET) == 0 && fwrite(&count, 4, 1, f) == 1;
        ok = (fclose(f) == 0) && ok;
        f = nullptr;
        return ok;
    }
};

// Reads the file section by section; `sink(tag, data, words)` consumes each payload (0 = ok).  Returns 0, or a negative code:
// -1 cannot open / short read, -2 bad magic or version, -3 checksum mismatch, -4 unknown tag, or the sink's value.
template <class Sink>
int read(const char* path, mkfhe_params* par, uint64_t (*words_by_tag)[TAG_MAX + 1], Sink&& sink) {
    FILE* f = fopen(path, "rb");
    if (!f) return -1;
    Header h{};
    int rc = 0;
    std::vector<uint32_t> buf;
    if (fread(&h, sizeof h, 1, f) != 1) rc = -1;
    else if (memcmp(h.magic, MAGIC, 8) != 0 || h.version != VERSION) rc = -2;
    if (!rc && par) words_to_params(h.par, *par);
    if (words_by_tag) memset(*words_by_tag, 0, sizeof *words_by_tag);
    for (uint32_t i = 0; !rc && i < h.nsections; ++i) {
        Section s{};
        if (fread(&s, sizeof s, 1, f) != 1) { rc = -1; break; }
        if (s.tag == 0 || s.tag > TAG_MAX) { rc = -4; break; }
        buf.resize(s.words);
        if (fread(buf.data(), 4, s.words, f) != s.words) { rc = -1; break; }
        if (fnv1a(buf.data(), s.words * 4) != s.fnv) { rc = -3; break; }
        if (words_by_tag) (*words_by_tag)[s.tag] = s.words;
        rc = sink(s.tag, buf.data(), s.words);
        const size_t tail = (16 - (sizeof s + s.words * 4) % 16) % 16;
        if (!rc && tail && fseek(f, long(tail), SEEK_CUR) != 0) rc = -1;
    }
    fclose(f);
    return rc;
}

}  // namespace mkfhe_keyfile

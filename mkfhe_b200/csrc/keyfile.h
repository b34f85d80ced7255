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
//
// Ciphertext / wire-table file (magic "MKFHECTX", same header and section framing): the reference serialises ciphertexts one
// object at a time through cereal (mntru-ciphertext.h:116-125, mklwe-ciphertext.h:161-172); here a whole table travels as
//   tags    : 8 a-parts [n_ct][k][n] mod q, 9 b-parts [n_ct] (MKLWE only)
// which is the layout mkfhe_eval_nand / mkfhe_eval_circuit take.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "../../include/mkfhe_b200.h"

namespace mkfhe_keyfile {

constexpr char MAGIC[8] = {'M', 'K', 'F', 'H', 'E', 'K', 'E', 'Y'};
constexpr char MAGIC_CT[8] = {'M', 'K', 'F', 'H', 'E', 'C', 'T', 'X'};  // ciphertext / wire-table file (same container)
constexpr uint32_t VERSION = 1;
enum Tag : uint32_t { BK = 1, PKEY = 2, KSK = 3, KSKA = 4, KSKB = 5, CTNAND = 6, TAG_MAX = 6, CT_A = 8, CT_B = 9, TAG_SLOTS = 10 };

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
    bool open(const char* path, const mkfhe_params& p, const char* magic = MAGIC) {
        f = fopen(path, "wb");
        if (!f) return false;
        Header h{};
        memcpy(h.magic, magic, 8);
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

// What a file must contain to be accepted: the magic of its kind, and for every tag the exact word count (0 = any).
struct Expect {
    const char* magic;
    const uint64_t* words_by_tag;  // [TAG_SLOTS] or null
};

// Reads the file section by section; `sink(tag, data, words)` consumes each payload (0 = ok).  The file is NOT trusted: the
// section table is validated against the file size (and `expect`) and every checksum is verified in a first pass, before the
// sink sees a single byte -- so a corrupt, truncated or foreign file never leaves a consumer half-updated, and no allocation
// is sized by an unchecked field.  Returns 0, or a negative code: -1 cannot open / short read, -2 bad magic or version,
// -3 checksum mismatch, -4 malformed section table (unknown or duplicate tag, size beyond the file, unexpected word count,
// kskA without kskB), or the sink's value.
template <class Sink>
int read(const char* path, const Expect& expect, mkfhe_params* par, uint64_t (*words_by_tag)[TAG_SLOTS], Sink&& sink) {
    FILE* f = fopen(path, "rb");
    if (!f) return -1;
    struct Closer {
        FILE* f;
        ~Closer() { fclose(f); }
    } closer{f};
    if (fseek(f, 0, SEEK_END) != 0) return -1;
    const long end = ftell(f);
    if (end < long(sizeof(Header)) || fseek(f, 0, SEEK_SET) != 0) return -1;
    Header h{};
    if (fread(&h, sizeof h, 1, f) != 1) return -1;
    if (memcmp(h.magic, expect.magic, 8) != 0 || h.version != VERSION) return -2;
    if (par) words_to_params(h.par, *par);
    if (words_by_tag) memset(*words_by_tag, 0, sizeof *words_by_tag);
    if (h.nsections > TAG_SLOTS) return -4;
    struct Entry {
        Section s;
        long at;
    } table[TAG_SLOTS];
    uint64_t have[TAG_SLOTS] = {0};
    std::vector<uint32_t> buf;
    constexpr size_t CHUNK = size_t(1) << 22;  // words per read: bounded scratch, whatever the section claims
    buf.resize(CHUNK);
    long pos = long(sizeof h);
    for (uint32_t i = 0; i < h.nsections; ++i) {  // pass 1: structure and checksums
        Section s{};
        if (fread(&s, sizeof s, 1, f) != 1) return -1;
        pos += long(sizeof s);
        if (s.tag == 0 || s.tag >= TAG_SLOTS || have[s.tag] || s.words == 0) return -4;
        if (s.words > uint64_t(end - pos) / 4) return -4;
        if (expect.words_by_tag && expect.words_by_tag[s.tag] != s.words) return -4;
        uint64_t hsh = 1469598103934665603ull;
        for (uint64_t done = 0; done < s.words;) {
            const size_t take = size_t(std::min<uint64_t>(CHUNK, s.words - done));
            if (fread(buf.data(), 4, take, f) != take) return -1;
            hsh = fnv1a(buf.data(), take * 4, hsh);
            done += take;
        }
        if (hsh != s.fnv) return -3;
        have[s.tag] = s.words;
        table[i] = Entry{s, pos};
        pos += long(s.words * 4);
        const long tail = long((16 - (sizeof s + s.words * 4) % 16) % 16);
        pos += tail;
        if (pos > end && i + 1 < h.nsections) return -4;
        if (tail && i + 1 < h.nsections && fseek(f, tail, SEEK_CUR) != 0) return -1;
    }
    if ((have[KSKA] != 0) != (have[KSKB] != 0)) return -4;
    if (words_by_tag) memcpy(*words_by_tag, have, sizeof have);
    for (uint32_t i = 0; i < h.nsections; ++i) {  // pass 2: hand the verified payloads to the consumer
        const Section& s = table[i].s;
        if (fseek(f, table[i].at, SEEK_SET) != 0) return -1;
        std::vector<uint32_t> payload(s.words);  // bounded by the file size (checked above)
        if (fread(payload.data(), 4, s.words, f) != s.words) return -1;
        if (const int rc = sink(s.tag, payload.data(), s.words)) return rc;
    }
    return 0;
}

}  // namespace mkfhe_keyfile

// Parallel gzip decoder, see pgzip.h.  Own RFC 1951 inflate (needed because a chunk that starts in the middle
// of a stream has to carry "unknown window byte" markers through its copies, which no zlib entry point does).
#include "pgzip.h"

#include <malloc.h>
#include <zlib.h>      // crc32, crc32_combine only
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

namespace pgz {

namespace {

#define PGZ_LIKELY(x) __builtin_expect(!!(x), 1)
#define PGZ_UNLIKELY(x) __builtin_expect(!!(x), 0)

constexpr uint32_t kWindow = 32768;

// ---------------------------------------------------------------------------------------------------------
// Bit reader: LSB-first, 64-bit buffer.  Reads past the end deliver zero bits; callers compare bitpos() with
// the size of the input at block boundaries to detect truncation.
struct Bits {
    const uint8_t *base; size_t size; size_t ip; uint64_t buf; unsigned cnt;
    Bits(const uint8_t *b, size_t n, uint64_t bitpos) : base(b), size(n), ip((size_t)(bitpos >> 3)), buf(0), cnt(0) {
        refill();
        drop((unsigned)(bitpos & 7));
    }
    inline void refill() {
        if (PGZ_LIKELY(ip + 8 <= size)) {
            uint64_t w; memcpy(&w, base + ip, 8);
            buf |= w << cnt;
            ip += (63 - cnt) >> 3;
            cnt |= 56;
        } else {
            while (cnt <= 56) {
                const uint64_t b = ip < size ? base[ip] : 0;
                buf |= b << cnt; ++ip; cnt += 8;
            }
        }
    }
    inline uint32_t peek(unsigned n) const { return (uint32_t)(buf & ((1ull << n) - 1)); }
    inline void drop(unsigned n) { buf >>= n; cnt -= n; }
    inline uint32_t take(unsigned n) { const uint32_t v = peek(n); drop(n); return v; }
    inline uint64_t bitpos() const { return (uint64_t)ip * 8 - cnt; }
    inline void align_byte() { drop(cnt & 7); }
};

// ---------------------------------------------------------------------------------------------------------
// Huffman decode tables.  Entry: bits 0-7 = code bits to drop, bits 8-15 = op, bits 16-31 = value.
constexpr uint32_t OP_LIT = 0x80, OP_EOB = 0x40, OP_SUB = 0x20, OP_BAD = 0x10;   // else: base + (op & 15) extra bits
constexpr int LIT_PB = 11, DIST_PB = 8, PRE_PB = 7;
constexpr int LIT_CAP = (1 << LIT_PB) + 2048, DIST_CAP = (1 << DIST_PB) + 1024, PRE_CAP = 1 << PRE_PB;

const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint8_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
const uint8_t kPreOrder[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

inline uint32_t lit_entry(int sym) {               // op/value part of a literal/length symbol
    if (sym < 256) return ((uint32_t)sym << 16) | (OP_LIT << 8);
    if (sym == 256) return OP_EOB << 8;
    if (sym > 285) return OP_BAD << 8;
    return ((uint32_t)kLenBase[sym - 257] << 16) | ((uint32_t)kLenExtra[sym - 257] << 8);
}
inline uint32_t dist_entry(int sym) {
    if (sym > 29) return OP_BAD << 8;
    return ((uint32_t)kDistBase[sym] << 16) | ((uint32_t)kDistExtra[sym] << 8);
}
inline uint32_t pre_entry(int sym) { return (uint32_t)sym << 16; }

inline uint32_t bitrev(uint32_t c, int len) {
    uint32_t r = 0;
    for (int i = 0; i < len; ++i) { r = (r << 1) | (c & 1); c >>= 1; }
    return r;
}

enum Kraft { K_OVER = -1, K_INCOMPLETE = 0, K_COMPLETE = 1, K_EMPTY = 2 };

// Canonical Huffman code of `n` symbols with lengths `lens` (0 = unused) into a two-level table.
template <typename EntryFn>
Kraft build_table(const uint8_t *lens, int n, int pb, uint32_t *tab, int cap, EntryFn entry) {
    int count[16] = {0};
    for (int s = 0; s < n; ++s) ++count[lens[s]];
    count[0] = 0;
    uint32_t kraft = 0; int used = 0;
    for (int l = 1; l <= 15; ++l) { kraft += (uint32_t)count[l] << (15 - l); used += count[l]; }
    if (kraft > 32768u) return K_OVER;
    uint32_t next[16]; uint32_t code = 0;
    for (int l = 1; l <= 15; ++l) { code = (code + (uint32_t)count[l - 1]) << 1; next[l] = code; }
    const int psize = 1 << pb;
    for (int i = 0; i < psize; ++i) tab[i] = (OP_BAD << 8) | 1;
    uint16_t rev[288];
    uint8_t submax[1 << LIT_PB];
    bool any_long = false;
    for (int s = 0; s < n; ++s) {
        const int l = lens[s];
        if (!l) continue;
        const uint32_t r = bitrev(next[l]++, l);
        rev[s] = (uint16_t)r;
        if (l <= pb) {
            const uint32_t e = entry(s) | (uint32_t)l;
            for (uint32_t i = r; i < (uint32_t)psize; i += 1u << l) tab[i] = e;
        } else {
            if (!any_long) { memset(submax, 0, (size_t)psize); any_long = true; }
            uint8_t &m = submax[r & (psize - 1)];
            if (l > m) m = (uint8_t)l;
        }
    }
    if (any_long) {
        int free_at = psize;
        for (int s = 0; s < n; ++s) {
            const int l = lens[s];
            if (l <= pb) continue;
            const uint32_t prefix = rev[s] & (psize - 1);
            if (!((tab[prefix] >> 8) & OP_SUB)) {
                const int sb = submax[prefix] - pb;
                if (free_at + (1 << sb) > cap) return K_OVER;
                tab[prefix] = ((uint32_t)free_at << 16) | ((OP_SUB | (uint32_t)sb) << 8) | (uint32_t)pb;
                for (int i = 0; i < (1 << sb); ++i) tab[free_at + i] = (OP_BAD << 8) | 1;
                free_at += 1 << sb;
            }
            const uint32_t off = tab[prefix] >> 16, sb = (tab[prefix] >> 8) & 15;
            const uint32_t e = entry(s) | (uint32_t)(l - pb);
            for (uint32_t i = rev[s] >> pb; i < (1u << sb); i += 1u << (l - pb)) tab[off + i] = e;
        }
    }
    if (!used) return K_EMPTY;
    return kraft == 32768u ? K_COMPLETE : K_INCOMPLETE;
}

// Second pass over the primary literal/length table: where a literal's code leaves room for another complete
// literal code inside the LIT_PB index bits, the entry yields both (op LIT | 1, value = first | second << 8).
// Sequence lines of a FASTQ file are 2-3 bit codes, so most lookups there return two bases.
void pair_literals(uint32_t *tab) {
    uint32_t copy[1 << LIT_PB];
    memcpy(copy, tab, sizeof copy);
    for (uint32_t i = 0; i < (1u << LIT_PB); ++i) {
        const uint32_t e = copy[i];
        if (!(e & (OP_LIT << 8))) continue;
        const uint32_t l1 = e & 0xff, e2 = copy[i >> l1];
        if (!(e2 & (OP_LIT << 8)) || l1 + (e2 & 0xff) > (uint32_t)LIT_PB) continue;
        tab[i] = ((e2 >> 16) << 24) | (e & 0x00ff0000u) | ((OP_LIT | 1u) << 8) | (l1 + (e2 & 0xff));
    }
}

struct Tables {
    uint32_t lit[LIT_CAP];
    uint32_t dist[DIST_CAP];
};

const Tables &fixed_tables() {
    static const Tables *t = [] {
        Tables *x = new Tables;
        uint8_t l[288];
        for (int i = 0; i < 144; ++i) l[i] = 8;
        for (int i = 144; i < 256; ++i) l[i] = 9;
        for (int i = 256; i < 280; ++i) l[i] = 7;
        for (int i = 280; i < 288; ++i) l[i] = 8;
        build_table(l, 288, LIT_PB, x->lit, LIT_CAP, lit_entry);
        uint8_t d[32];
        for (int i = 0; i < 32; ++i) d[i] = 5;
        build_table(d, 32, DIST_PB, x->dist, DIST_CAP, dist_entry);
        return x;
    }();
    return *t;
}

// Reads the header of a dynamic block (after the 3 block-type bits) and builds its tables.  `strict` is the
// block finder's mode: only headers a sane compressor writes pass (complete codes, an end-of-block code).
bool read_dynamic_header(Bits &in, Tables &t, bool strict) {
    in.refill();
    const int hlit = (int)in.take(5) + 257, hdist = (int)in.take(5) + 1, hclen = (int)in.take(4) + 4;
    if (hlit > 286 || hdist > 30) return false;
    uint8_t pre[19] = {0};
    in.refill();
    for (int i = 0; i < hclen; ++i) {
        if (in.cnt < 3) in.refill();
        pre[kPreOrder[i]] = (uint8_t)in.take(3);
    }
    uint32_t ptab[PRE_CAP];
    const Kraft pk = build_table(pre, 19, PRE_PB, ptab, PRE_CAP, pre_entry);
    if (pk == K_OVER || pk == K_EMPTY || (strict && pk != K_COMPLETE)) return false;
    uint8_t lens[286 + 30 + 138];
    int n = 0; const int total = hlit + hdist;
    while (n < total) {
        in.refill();
        const uint32_t e = ptab[in.peek(PRE_PB)];
        if ((e >> 8) & OP_BAD) return false;
        in.drop(e & 0xff);
        const int sym = (int)(e >> 16);
        if (sym < 16) { lens[n++] = (uint8_t)sym; continue; }
        int rep; uint8_t v = 0;
        if (sym == 16) { if (!n) return false; v = lens[n - 1]; rep = 3 + (int)in.take(2); }
        else if (sym == 17) rep = 3 + (int)in.take(3);
        else rep = 11 + (int)in.take(7);
        if (n + rep > total) return false;
        memset(lens + n, v, (size_t)rep); n += rep;
    }
    if (!lens[256]) return false;                                   // no end-of-block code
    const Kraft lk = build_table(lens, hlit, LIT_PB, t.lit, LIT_CAP, lit_entry);
    if (lk == K_OVER || lk == K_EMPTY || (strict && lk != K_COMPLETE)) return false;
    if (!strict) pair_literals(t.lit);
    const Kraft dk = build_table(lens + hlit, hdist, DIST_PB, t.dist, DIST_CAP, dist_entry);
    if (dk == K_OVER) return false;
    if (strict && dk == K_INCOMPLETE) {                             // zlib allows ONE distance code of length 1
        int used = 0;
        for (int i = 0; i < hdist; ++i) used += lens[hlit + i] != 0;
        if (used != 1) return false;
    }
    return true;
}

// ---------------------------------------------------------------------------------------------------------
// Output of one decode run.  `base` starts with a kWindow-symbol prefix holding the previous 32 KiB: real
// bytes when they are known (T = uint8_t), markers 0x8000 + i when they are not (T = uint16_t), so copies
// never need to know whether they reach back before the run's start.
template <typename T>
struct OutBuf {
    T *base = nullptr; size_t cap = 0, pos = kWindow, limit = 0;
    size_t floor = 0;                                   // lowest index a copy may read
    OutBuf() {}
    OutBuf(const OutBuf &) = delete;
    OutBuf &operator=(const OutBuf &) = delete;
    ~OutBuf() { free(base); }
    bool reserve(size_t want) {                         // total symbols incl. prefix
        if (want <= cap) return true;
        if (want > limit) return false;
        size_t ncap = std::max<size_t>(cap * 2, 1u << 16);
        while (ncap < want) ncap *= 2;
        ncap = std::min(ncap, limit);
        T *nb = (T *)realloc(base, ncap * sizeof(T));
        if (!nb) return false;
        base = nb; cap = ncap;
        return true;
    }
    size_t produced() const { return pos - kWindow; }
    const T *out() const { return base + kWindow; }
};

constexpr size_t kSlack = 320;                          // longest match + copy overshoot

template <typename T>
inline void copy_match(T *dst, uint32_t dist, uint32_t len) {
    constexpr uint32_t W = 16 / sizeof(T);
    const T *src = dst - dist;
    if (dist >= W) {
        for (uint32_t i = 0; i < len; i += W) memcpy(dst + i, src + i, 16);
    } else if (dist == 1) {
        const T v = src[0];
        for (uint32_t i = 0; i < len; ++i) dst[i] = v;
    } else {
        for (uint32_t i = 0; i < len; ++i) dst[i] = src[i];
    }
}

// One Huffman-coded block body.  0 = end of block reached, -1 = invalid data, -2 = output limit.
template <typename T>
int inflate_block(Bits &in, const Tables &t, OutBuf<T> &o) {
    T *base = o.base; size_t pos = o.pos, cap = o.cap;
    for (;;) {
        if (PGZ_UNLIKELY(pos + kSlack > cap)) {
            o.pos = pos;
            if (!o.reserve(pos + kSlack + (1u << 16))) return -2;
            base = o.base; cap = o.cap;
        }
        in.refill();
        if (PGZ_UNLIKELY(in.ip > in.size + 8)) { o.pos = pos; return -1; }   // ran off the end of the input
        uint32_t e = t.lit[in.peek(LIT_PB)];
        if (PGZ_UNLIKELY(e & (OP_SUB << 8))) { in.drop(e & 0xff); e = t.lit[(e >> 16) + in.peek((e >> 8) & 15)]; }
        in.drop(e & 0xff);
        if (e & (OP_LIT << 8)) {
            // up to three lookups fit the bits already buffered (3 x 15 <= 56); each yields one or two literals
            base[pos] = (T)((e >> 16) & 0xff); base[pos + 1] = (T)(e >> 24); pos += 1 + ((e >> 8) & 1);
            e = t.lit[in.peek(LIT_PB)];
            if (!(e & (OP_LIT << 8))) continue;
            in.drop(e & 0xff);
            base[pos] = (T)((e >> 16) & 0xff); base[pos + 1] = (T)(e >> 24); pos += 1 + ((e >> 8) & 1);
            e = t.lit[in.peek(LIT_PB)];
            if (!(e & (OP_LIT << 8))) continue;
            in.drop(e & 0xff);
            base[pos] = (T)((e >> 16) & 0xff); base[pos + 1] = (T)(e >> 24); pos += 1 + ((e >> 8) & 1);
            continue;
        }
        if (PGZ_UNLIKELY(e & ((OP_EOB | OP_BAD) << 8))) { o.pos = pos; return (e & (OP_EOB << 8)) ? 0 : -1; }
        const uint32_t len = (e >> 16) + in.take((e >> 8) & 15);
        uint32_t d = t.dist[in.peek(DIST_PB)];
        if (PGZ_UNLIKELY(d & (OP_SUB << 8))) { in.drop(d & 0xff); d = t.dist[(d >> 16) + in.peek((d >> 8) & 15)]; }
        in.drop(d & 0xff);
        if (PGZ_UNLIKELY(d & (OP_BAD << 8))) { o.pos = pos; return -1; }
        const uint32_t dist = (d >> 16) + in.take((d >> 8) & 15);
        if (PGZ_UNLIKELY((size_t)dist > pos - o.floor)) { o.pos = pos; return -1; }
        copy_match(base + pos, dist, len);
        pos += len;
    }
}

struct MemberEnd { size_t out_pos; uint32_t crc, isize; };
struct Target { uint64_t bit; uint8_t kind; };          // kind 0: deflate block, 1: gzip member header (byte aligned)

// gzip member header at byte `off` (RFC 1952).  Returns the offset of the deflate data, 0 if this is not a header.
size_t member_header(const uint8_t *gz, size_t size, size_t off) {
    if (off + 18 > size) return 0;
    const uint8_t *p = gz + off;
    if (p[0] != 0x1f || p[1] != 0x8b || p[2] != 8 || (p[3] & 0xe0)) return 0;
    const uint8_t flg = p[3];
    size_t q = off + 10;
    if (flg & 4) { if (q + 2 > size) return 0; q += 2 + ((size_t)gz[q] | ((size_t)gz[q + 1] << 8)); }
    if (flg & 8) { while (q < size && gz[q]) ++q; ++q; }
    if (flg & 16) { while (q < size && gz[q]) ++q; ++q; }
    if (flg & 2) q += 2;
    return q + 8 <= size ? q : 0;
}

struct RunResult {
    bool ok = false, eof = false;
    uint64_t end_bit = 0; uint8_t end_kind = 0;
    std::vector<MemberEnd> ends;
    const char *why = "";
};

// Decodes from (start_bit, start_kind) until a block boundary that is one of `targets` (sorted by bit, all
// beyond the start), or lies at/after `stop_after`, or comes after `soft_out` symbols of output, or the end of the file.
template <typename T>
void decode_run(const uint8_t *gz, size_t size, uint64_t start_bit, uint8_t start_kind, const Target *targets,
                size_t nt, uint64_t stop_after, OutBuf<T> &o, RunResult &r, size_t soft_out = SIZE_MAX) {
    const uint64_t total_bits = (uint64_t)size * 8;
    uint64_t b = start_bit; uint8_t kind = start_kind;
    bool first = true; size_t ti = 0;
    Bits in(gz, size, kind == 0 ? b : 0);
    std::unique_ptr<Tables> dyn(new Tables);
    if (!o.reserve(kWindow + (1u << 16))) { r.why = "out of memory"; return; }
    for (;;) {
        if (!first) {
            bool stop = b >= stop_after || o.produced() >= soft_out;
            while (ti < nt && targets[ti].bit < b) ++ti;
            for (size_t tj = ti; !stop && tj < nt && targets[tj].bit == b; ++tj) stop = targets[tj].kind == kind;
            if (stop) break;
        }
        if (kind == 1) {
            const size_t data = member_header(gz, size, (size_t)(b >> 3));
            if (!data) {
                if (first) { r.why = "not a gzip member"; return; }
                r.eof = true; break;                    // trailing garbage is ignored, as gzread does
            }
            first = false;
            b = (uint64_t)data * 8; kind = 0;
            in = Bits(gz, size, b);
            continue;
        }
        first = false;
        in.refill();
        const uint32_t bfinal = in.take(1), type = in.take(2);
        if (type == 0) {
            in.align_byte(); in.refill();
            const uint32_t len = in.take(16), nlen = in.take(16);
            if ((len ^ 0xffffu) != nlen) { r.why = "invalid stored block"; return; }
            const uint64_t at = in.bitpos() >> 3;
            if (at + len > size) { r.why = "truncated stored block"; return; }
            if (!o.reserve(o.pos + len + kSlack)) { r.why = "output limit"; return; }
            for (uint32_t i = 0; i < len; ++i) o.base[o.pos + i] = (T)gz[at + i];
            o.pos += len;
            in = Bits(gz, size, (at + len) * 8);
        } else if (type == 3) {
            r.why = "invalid block type"; return;
        } else {
            const Tables *t = &fixed_tables();
            if (type == 2) {
                if (!read_dynamic_header(in, *dyn, false)) { r.why = "invalid code lengths"; return; }
                t = dyn.get();
            }
            const int rc = inflate_block(in, *t, o);
            if (rc) { r.why = rc == -2 ? "output limit" : "invalid compressed data"; return; }
        }
        if (in.bitpos() > total_bits) { r.why = "truncated input"; return; }
        if (bfinal) {
            in.align_byte();
            const uint64_t at = in.bitpos() >> 3;
            if (at + 8 > size) { r.why = "truncated trailer"; return; }
            uint32_t crc, isize; memcpy(&crc, gz + at, 4); memcpy(&isize, gz + at + 4, 4);
            r.ends.push_back({o.produced(), crc, isize});
            b = (at + 8) * 8; kind = 1;
            if (at + 8 >= size) { r.eof = true; break; }
        } else {
            b = in.bitpos(); kind = 0;
        }
    }
    r.end_bit = b; r.end_kind = kind; r.ok = true;
}

// First position in [from_bit, to_bit) that parses as a gzip member header or as the header of a non-final
// dynamic-Huffman block with complete codes.
bool find_start(const uint8_t *gz, size_t size, uint64_t from_bit, uint64_t to_bit, Target &out) {
    std::unique_ptr<Tables> t(new Tables);
    to_bit = std::min<uint64_t>(to_bit, (uint64_t)size * 8);
    static const uint8_t kraft7[8] = {0, 64, 32, 16, 8, 4, 2, 1};
    for (uint64_t byte = from_bit >> 3; byte * 8 < to_bit; ++byte) {
        if (byte + 17 > size) return false;
        unsigned __int128 w; memcpy(&w, gz + byte, 16);
        if (byte * 8 >= from_bit && ((uint64_t)w & 0xe0ffffffull) == 0x00088b1full && member_header(gz, size, (size_t)byte)) {
            out = {byte * 8, 1}; return true;
        }
        const uint64_t lo = (uint64_t)w;
        for (uint32_t m = (uint32_t)(~lo & ~(lo >> 1) & (lo >> 2)) & 0xff; m; m &= m - 1) {   // BFINAL = 0, BTYPE = 2
            const unsigned s = (unsigned)__builtin_ctz(m);
            const uint64_t bit = byte * 8 + s;
            if (bit < from_bit) continue;
            if (bit >= to_bit) return false;
            unsigned __int128 x = w >> s;               // 120 valid bits: 17 of header + up to 57 of precode lengths
            const uint32_t v = (uint32_t)x;
            if ((v & 7) != 4 || ((v >> 3) & 31) > 29 || ((v >> 8) & 31) > 29) continue;
            const int hclen = (int)((v >> 13) & 15) + 4;
            x >>= 17;
            uint32_t kraft = 0;
            for (int i = 0; i < hclen; ++i) { kraft += kraft7[(uint32_t)x & 7]; x >>= 3; }
            if (kraft != 128) continue;                 // the code-length code is not complete
            Bits in(gz, size, bit + 3);
            if (read_dynamic_header(in, *t, true) && in.bitpos() < (uint64_t)size * 8) { out = {bit, 0}; return true; }
        }
    }
    return false;
}

// ---------------------------------------------------------------------------------------------------------
template <typename Fn>
void parallel_for(size_t n, int threads, Fn fn) {
    if (n == 0) return;
    const size_t nt = std::min<size_t>(n, (size_t)std::max(1, threads));
    if (nt == 1) { for (size_t i = 0; i < n; ++i) fn(i); return; }
    std::atomic<size_t> next{0};
    std::vector<std::thread> th;
    for (size_t t = 0; t < nt; ++t)
        th.emplace_back([&] { for (size_t i; (i = next++) < n;) fn(i); });
    for (auto &x : th) x.join();
}

struct Chunk {
    Target start{0, 0}; bool has_start = false;
    bool exact = false;                                 // decoded with a known window: bytes, no markers
    OutBuf<uint16_t> spec; OutBuf<uint8_t> bytes;
    RunResult res;
    std::vector<uint8_t> window;                        // the 32 KiB before this chunk (chain chunks only)
    size_t dst_off = 0;
    std::vector<uint32_t> piece_crc; std::vector<size_t> piece_len;
    size_t n_out() const { return exact ? bytes.produced() : spec.produced(); }
    void reset() {                                      // buffers keep their capacity from wave to wave
        has_start = false; exact = false; spec.pos = bytes.pos = kWindow; spec.floor = bytes.floor = 0;
        res = RunResult(); dst_off = 0;
    }
};

// window' = last 32 KiB of (window ++ output of c)
void propagate_window(std::vector<uint8_t> &w, const Chunk &c) {
    const size_t n = c.n_out();
    std::vector<uint8_t> nw(kWindow);
    const size_t keep = n >= kWindow ? 0 : kWindow - n;            // bytes of the old window that survive
    memcpy(nw.data(), w.data() + (kWindow - keep), keep);
    const size_t from = n - (kWindow - keep);
    if (c.exact) {
        memcpy(nw.data() + keep, c.bytes.out() + from, kWindow - keep);
    } else {
        const uint16_t *s = c.spec.out() + from;
        for (size_t i = 0; i < kWindow - keep; ++i) nw[keep + i] = s[i] < 0x8000 ? (uint8_t)s[i] : w[s[i] & 0x7fff];
    }
    w.swap(nw);
}

// markers -> bytes, CRC-32 of every piece between member ends
void resolve_range(Chunk &c, char *dst, size_t from, size_t n) {     // symbols [from, n)
    if (c.exact) {
        memcpy(dst + from, c.bytes.out() + from, n - from);
    } else {
        const uint16_t *s = c.spec.out(); const uint8_t *w = c.window.data();
        size_t i = from;
#if defined(__SSE2__)
        for (; i + 16 <= n; i += 16) {
            const __m128i a = _mm_loadu_si128((const __m128i *)(s + i)), b = _mm_loadu_si128((const __m128i *)(s + i + 8));
            if (PGZ_LIKELY(!(_mm_movemask_epi8(_mm_or_si128(a, b)) & 0xaaaa))) {
                _mm_storeu_si128((__m128i *)(dst + i), _mm_packus_epi16(a, b));
            } else {
                for (int j = 0; j < 16; ++j) dst[i + j] = (char)(s[i + j] < 0x8000 ? (uint8_t)s[i + j] : w[s[i + j] & 0x7fff]);
            }
        }
#endif
        for (; i < n; ++i) dst[i] = (char)(s[i] < 0x8000 ? (uint8_t)s[i] : w[s[i] & 0x7fff]);
    }
}

// CRC-32 of every piece between member ends, over the resolved bytes
void crc_pieces(Chunk &c, const char *dst, size_t from, size_t to, bool begin) {
    if (begin) { c.piece_crc.assign(c.res.ends.size() + 1, (uint32_t)crc32(0L, Z_NULL, 0)); c.piece_len.assign(c.res.ends.size() + 1, 0); }
    size_t p = 0;
    while (from < to) {
        const size_t end = p < c.res.ends.size() ? c.res.ends[p].out_pos : c.n_out();
        if (from >= end && p < c.res.ends.size()) { ++p; continue; }
        const size_t step = std::min(to, end) - from;
        c.piece_crc[p] = (uint32_t)crc32(c.piece_crc[p], (const Bytef *)dst + from, (uInt)step);
        c.piece_len[p] += step;
        from += step;
    }
}

// markers -> bytes and CRCs, in blocks small enough that the CRC reads the bytes back from cache
void resolve_chunk(Chunk &c, char *dst) {
    const size_t n = c.n_out(), block = 64 << 10;
    crc_pieces(c, dst, 0, 0, true);
    for (size_t a = 0; a < n; a += block) {
        const size_t b = std::min(n, a + block);
        resolve_range(c, dst, a, b);
        crc_pieces(c, dst, a, b, false);
    }
}

// One consumer thread taking pieces in order, at most one in flight behind the decoder.
struct Consumer {
    Sink &sink; std::mutex mu; std::condition_variable cv;
    char *buf = nullptr; size_t n = 0; bool last = false, has = false, quit = false, ok = true;
    std::vector<char *> spare;                          // consumed buffers, reused (realloc) by the decoder
    std::thread th;
    explicit Consumer(Sink &s) : sink(s), th([this] { loop(); }) {}
    void loop() {
        for (;;) {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return has || quit; });
            if (!has) return;
            char *b = buf; const size_t len = n; const bool l = last;
            lk.unlock();
            const bool r = sink.consume(b, len, l);
            lk.lock();
            spare.push_back(b);
            if (!r) ok = false;
            has = false; cv.notify_all();
        }
    }
    bool submit(char *b, size_t len, bool l) {          // takes ownership of b
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return !has; });
        if (!ok) { spare.push_back(b); return false; }
        buf = b; n = len; last = l; has = true; cv.notify_all();
        return true;
    }
    char *get_buffer(size_t bytes) {
        char *b = nullptr;
        { std::lock_guard<std::mutex> lk(mu); if (!spare.empty()) { b = spare.back(); spare.pop_back(); } }
        if (b && malloc_usable_size(b) >= bytes) return b;
        free(b);
        return (char *)malloc(bytes + bytes / 4);
    }
    ~Consumer() { for (char *b : spare) free(b); }
    bool finish() {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return !has; });
        quit = true; cv.notify_all(); lk.unlock();
        th.join();
        return ok;
    }
};

}  // namespace

bool is_gzip(const unsigned char *p, size_t n) { return n >= 18 && p[0] == 0x1f && p[1] == 0x8b && p[2] == 8; }

bool inflate_parallel(const unsigned char *gz, size_t size, int threads, Sink &sink, std::string &err,
                      Stats *stats, size_t chunk_bytes) {
    Stats st;
    if (!is_gzip(gz, size) || !member_header(gz, size, 0)) { err = "not a gzip file"; return false; }
    threads = std::max(1, threads);
    const size_t C = std::max<size_t>(chunk_bytes ? chunk_bytes : (size_t)2 << 20, 64);
    const size_t per_wave = threads == 1 ? 1 : (size_t)threads * 2;
    // A run that expands more than this is given up on (a false start decoding noise, or extremely repetitive
    // data); the chain walk then decodes that stretch sequentially without a limit.
    const size_t spec_limit = kWindow + std::max<size_t>(C * 64, (size_t)4 << 20);
    // Runs with a known window are not limited, but they hand over at the next block boundary once they have
    // produced soft_out bytes, and a wave ends early once it holds wave_out bytes: memory stays bounded on
    // inputs that expand a thousandfold (a 2 MiB chunk of a run of one base is 2 GB of output).
    static const size_t soft_env = [] { const char *e = getenv("KQ_PGZ_SOFT_OUT"); const long long v = e ? atoll(e) : 0; return v > 0 ? (size_t)v : (size_t)0; }();
    const size_t soft_out = soft_env ? soft_env : (size_t)256 << 20, wave_out = soft_out * 4;

    std::vector<uint8_t> window(kWindow, 0);
    uint64_t pos_bit = 0; uint8_t pos_kind = 1;
    uint32_t member_crc = (uint32_t)crc32(0L, Z_NULL, 0); uint64_t member_len = 0;
    std::string carry;
    Consumer consumer(sink);
    bool eof = false, good = true;
    std::vector<std::unique_ptr<Chunk>> chunks;          // reused by every wave: no fresh pages to fault in

    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    while (!eof && good) {
        double tm = now();
        const size_t wave_byte = (size_t)(pos_bit >> 3);
        const size_t nchunks = std::max<size_t>(1, std::min(per_wave, (size - wave_byte + C - 1) / C));
        const bool to_end = wave_byte + nchunks * C >= size;
        const uint64_t stop_after = to_end ? UINT64_MAX : (uint64_t)(wave_byte + nchunks * C) * 8;
        while (chunks.size() < nchunks) chunks.emplace_back(new Chunk);
        for (size_t i = 0; i < nchunks; ++i) chunks[i]->reset();
        chunks[0]->start = {pos_bit, pos_kind}; chunks[0]->has_start = true; chunks[0]->exact = true;

        // A1: starting points
        parallel_for(nchunks - 1, threads, [&](size_t j) {
            Chunk &c = *chunks[j + 1];
            const uint64_t from = std::max<uint64_t>((uint64_t)(wave_byte + (j + 1) * C) * 8, pos_bit + 1);
            const uint64_t to = (uint64_t)std::min(size, wave_byte + (j + 2) * C) * 8;
            c.has_start = from < to && find_start(gz, size, from, to, c.start);
        });
        st.t_find += now() - tm; tm = now();
        std::vector<Target> targets; std::vector<size_t> first_target(nchunks + 1, 0);
        for (size_t i = 0; i < nchunks; ++i) {
            first_target[i] = targets.size();           // chunk i stops at the starts of chunks > i
            if (i + 1 < nchunks && chunks[i + 1]->has_start) targets.push_back(chunks[i + 1]->start);
        }
        // A2: decode
        parallel_for(nchunks, threads, [&](size_t i) {
            Chunk &c = *chunks[i];
            if (!c.has_start) return;
            const Target *t = targets.data() + first_target[i]; const size_t nt = targets.size() - first_target[i];
            if (c.exact) {
                c.bytes.limit = SIZE_MAX / 2;
                if (!c.bytes.reserve(kWindow + 4 * C)) return;
                memcpy(c.bytes.base, window.data(), kWindow);
                decode_run(gz, size, c.start.bit, c.start.kind, t, nt, stop_after, c.bytes, c.res, soft_out);
            } else {
                c.spec.limit = spec_limit;
                if (!c.spec.reserve(std::min(spec_limit, kWindow + 4 * C))) return;
                for (uint32_t m = 0; m < kWindow; ++m) c.spec.base[m] = (uint16_t)(0x8000u | m);
                decode_run(gz, size, c.start.bit, c.start.kind, t, nt, stop_after, c.spec, c.res);
            }
        });
        st.chunks += nchunks;
        st.t_decode += now() - tm; tm = now();

        // B: walk the chain from the known position; every link proves the next chunk's starting point
        std::vector<Chunk *> chain;
        std::vector<std::unique_ptr<Chunk>> extra;
        uint64_t at = pos_bit; uint8_t at_kind = pos_kind;
        size_t total = 0;
        for (;;) {
            Chunk *c = nullptr; size_t idx = nchunks;
            for (size_t i = 0; i < nchunks; ++i)
                if (chunks[i]->has_start && chunks[i]->start.bit == at && chunks[i]->start.kind == at_kind) { c = chunks[i].get(); idx = i; break; }
            if (!c || !c->res.ok) {
                if (c && c->exact) { err = std::string("gzip: ") + c->res.why; good = false; break; }
                // speculation failed here (or the run ended where no chunk starts): decode with the real window
                extra.emplace_back(new Chunk);
                Chunk *f = extra.back().get();
                f->exact = true; f->start = {at, at_kind}; f->has_start = true;
                f->bytes.limit = SIZE_MAX / 2;
                size_t ft = 0;
                if (idx < nchunks) ft = first_target[idx];
                else while (ft < targets.size() && targets[ft].bit <= at) ++ft;
                if (!f->bytes.reserve(kWindow + (1u << 16))) { err = "gzip: out of memory"; good = false; break; }
                memcpy(f->bytes.base, window.data(), kWindow);
                decode_run(gz, size, at, at_kind, targets.data() + ft, targets.size() - ft, stop_after, f->bytes, f->res, soft_out);
                ++st.fallbacks;
                if (!f->res.ok) { err = std::string("gzip: ") + f->res.why; good = false; break; }
                c = f;
            }
            c->window = window;
            propagate_window(window, *c);
            chain.push_back(c); ++st.chained;
            total += c->n_out();
            at = c->res.end_bit; at_kind = c->res.end_kind;
            if (c->res.eof) { eof = true; break; }
            if (at >= stop_after || total >= wave_out) break;
        }
        if (!good) break;
        pos_bit = at; pos_kind = at_kind;
        st.t_chain += now() - tm; tm = now();

        // C: resolve into one contiguous piece behind the carried-over tail
        char *buf = consumer.get_buffer(carry.size() + total + 1);
        if (!buf) { err = "gzip: out of memory"; good = false; break; }
        memcpy(buf, carry.data(), carry.size());
        size_t off = carry.size();
        for (Chunk *c : chain) { c->dst_off = off; off += c->n_out(); }
        parallel_for(chain.size(), threads, [&](size_t i) { resolve_chunk(*chain[i], buf + chain[i]->dst_off); });
        for (Chunk *c : chain) {
            for (size_t p = 0; p < c->piece_crc.size() && good; ++p) {
                member_crc = (uint32_t)crc32_combine(member_crc, c->piece_crc[p], (z_off_t)c->piece_len[p]);
                member_len += c->piece_len[p];
                if (p < c->res.ends.size()) {
                    if (member_crc != c->res.ends[p].crc || (uint32_t)member_len != c->res.ends[p].isize) {
                        err = "gzip: CRC or length mismatch"; good = false;
                    }
                    member_crc = (uint32_t)crc32(0L, Z_NULL, 0); member_len = 0; ++st.members;
                }
            }
        }
        if (!good) { free(buf); break; }
        st.t_resolve += now() - tm; tm = now();
        st.out_bytes += total;
        const size_t have = carry.size() + total;
        size_t take = eof ? have : std::min(have, sink.split(buf, have, false));
        carry.assign(buf + take, have - take);
        if (!consumer.submit(buf, take, eof)) { err = "gzip: consumer stopped"; good = false; break; }
        st.t_hand += now() - tm;
    }
    if (good && !eof) { err = "gzip: unexpected end"; good = false; }
    const bool cons_ok = consumer.finish();
    if (good && !cons_ok) { err = "gzip: consumer stopped"; good = false; }
    if (stats) *stats = st;
    return good;
}

}  // namespace pgz

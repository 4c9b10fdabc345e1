// Per-base output files.  Same bytes as the reference's writers, built with a flat integer formatter and
// one buffered write per few MiB instead of per-value iostream calls.
//   .bed / .csvtable   DBG::printTable                   src/kreeq-output.cpp:138-241
//   .kwig              DBG::printTableCompressed         src/kreeq-output.cpp:243-303
//   .bkwig             DBG::printTableCompressedBinary   src/kreeq-output.cpp:305-399
#include "kreeq_cli.h"

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <functional>
#include <thread>

#include <unistd.h>

namespace cli {
namespace {

// Formats blocks [0, n) with fmt(block, out) on all cores, a wave at a time; the blocks of a wave are then
// written at their final offsets concurrently (pwrite), which is what a tmpfs / page-cache target rewards.
bool emit_blocks(FILE *f, size_t n, const std::function<void(size_t, std::string &)> &fmt) {
    const size_t nthreads = std::max<size_t>(1, std::min<size_t>(std::thread::hardware_concurrency(), 32));
    const size_t wave = nthreads * 2;
    std::vector<std::string> out(wave);
    std::vector<off_t> at(wave);
    if (fflush(f) != 0) return false;
    const int fd = fileno(f);
    off_t pos = lseek(fd, 0, SEEK_END);
    if (pos < 0) return false;
    std::atomic<bool> ok{true};
    for (size_t base = 0; base < n; base += wave) {
        const size_t cnt = std::min(wave, n - base);
        auto run = [&](const std::function<void(size_t)> &job) {
            std::atomic<size_t> next{0};
            auto work = [&] { for (size_t i; (i = next.fetch_add(1)) < cnt;) job(i); };
            std::vector<std::thread> th;
            for (size_t t = 1; t < std::min(nthreads, cnt); ++t) th.emplace_back(work);
            work();
            for (auto &t : th) t.join();
        };
        run([&](size_t i) { out[i].clear(); fmt(base + i, out[i]); });
        for (size_t i = 0; i < cnt; ++i) { at[i] = pos; pos += (off_t)out[i].size(); }
        run([&](size_t i) {
            const char *p = out[i].data(); size_t left = out[i].size(); off_t o = at[i];
            while (left) {
                const ssize_t w = pwrite(fd, p, left, o);
                if (w <= 0) { ok = false; return; }
                p += w; left -= (size_t)w; o += w;
            }
        });
        if (!ok) return false;
    }
    return lseek(fd, pos, SEEK_SET) >= 0;
}

// decimal digits of v at p, returns the end; two digits at a time from a table
inline char *put_num(char *p, uint64_t v) {
    static const char *const pairs =
        "0001020304050607080910111213141516171819202122232425262728293031323334353637383940414243444546474849"
        "5051525354555657585960616263646566676869707172737475767778798081828384858687888990919293949596979899";
    if (v < 10) { *p = (char)('0' + v); return p + 1; }
    if (v < 100) { memcpy(p, pairs + 2 * v, 2); return p + 2; }
    char tmp[20];
    int len = 0;
    while (v >= 100) { memcpy(tmp + len, pairs + 2 * (v % 100), 2); len += 2; v /= 100; }   // stored low pair first, digits swapped back below
    if (v >= 10) { memcpy(tmp + len, pairs + 2 * v, 2); len += 2; }
    else tmp[len++] = (char)('0' + v), tmp[len++] = 0;
    // tmp holds pairs "ab" least significant first; emit most significant first
    int i = len - 2;
    if (tmp[i + 1] == 0) { *p++ = tmp[i]; i -= 2; }
    for (; i >= 0; i -= 2) { p[0] = tmp[i]; p[1] = tmp[i + 1]; p += 2; }
    return p;
}

inline void append_num(std::string &o, uint64_t v) {
    char tmp[20];
    int len = 0;
    do { tmp[len++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (len) o.push_back(tmp[--len]);
}

constexpr uint64_t BLOCK = 1u << 20;               // bases per formatting task
struct Span { const Path *path; const Component *comp; uint64_t abs_pos, from, to; };   // bases [from, to) of one segment

std::vector<Span> spans_of(const Genome &g, uint64_t block = BLOCK) {
    std::vector<Span> v;
    for (const Path &p : g.paths) {
        uint64_t abs_pos = 0;
        for (const Component &c : p.components) {
            if (c.is_segment) {
                const uint64_t len = g.bases[c.segment].size();
                if (len == 0) v.push_back({&p, &c, abs_pos, 0, 0});
                for (uint64_t a = 0; a < len; a += block) v.push_back({&p, &c, abs_pos, a, std::min(len, a + block)});
            }
            abs_pos += c.len;
        }
    }
    return v;
}

struct Out {
    FILE *f;
    std::vector<char> buf;
    size_t n = 0;
    explicit Out(const std::string &file) : f(fopen(file.c_str(), "wb")), buf(8u << 20) {}
    ~Out() { close(); }
    bool ok() const { return f != nullptr; }
    void room(size_t need) {
        if (n + need > buf.size()) flush();
        if (need > buf.size()) buf.resize(need);
    }
    void flush() { if (f && n) fwrite(buf.data(), 1, n, f); n = 0; }
    bool close() { flush(); const bool good = f && fclose(f) == 0; f = nullptr; return good; }
    void put(char c) { room(1); buf[n++] = c; }
    void put(const char *s, size_t len) { room(len); memcpy(buf.data() + n, s, len); n += len; }
    void put(const std::string &s) { put(s.data(), s.size()); }
    void num(uint64_t v) {
        room(20);
        char tmp[20];
        int len = 0;
        do { tmp[len++] = (char)('0' + v % 10); v /= 10; } while (v);
        while (len) buf[n++] = tmp[--len];
    }
    template <typename T> void raw(T v) { put(reinterpret_cast<const char *>(&v), sizeof(T)); }
};

// What the writers print for one base: cov, then the edge support toward the next assembly base, then
// toward the previous one (kreeq-output.cpp:277, 379-383).
inline void triple(const kq_base &b, uint32_t &cov, uint32_t &F, uint32_t &B) {
    cov = b.cov;
    F = b.isFw ? b.fw : b.bw;
    B = b.isFw ? b.bw : b.fw;
}

} // namespace

bool write_kwig(const std::string &file, int k, const Genome &g) {
    FILE *f = fopen(file.c_str(), "wb");
    if (!f) return false;
    fprintf(f, "%d\n", k);
    const std::vector<Span> spans = spans_of(g);
    const bool ok = emit_blocks(f, spans.size(), [&](size_t i, std::string &o) {
        const Span &sp = spans[i];
        o.reserve((sp.to - sp.from) * 12 + 128);
        if (sp.from == 0) {
            o += "fixedStep chrom="; o += sp.path->header; o += " start="; append_num(o, sp.abs_pos); o += " step=1\n";
        }
        const kq_base *b = g.bases[sp.comp->segment].data();
        const size_t head = o.size();
        o.resize(head + (sp.to - sp.from) * 33);   // 3 x (10 digits + separator) per row at most
        char *p = &o[head];
        for (uint64_t x = sp.from; x < sp.to; ++x) {
            uint32_t cov, F, B;
            triple(b[x], cov, F, B);
            p = put_num(p, cov); *p++ = ','; p = put_num(p, F); *p++ = ','; p = put_num(p, B); *p++ = '\n';
        }
        o.resize((size_t)(p - o.data()));
    });
    return (fclose(f) == 0) && ok;
}

// Each line repeats the last k values of the three tracks (zeros before the segment start).  Consecutive lines
// share k - 1 of their k values, so a block first writes every track once as one joined text
// "v0:v1:v2:..." with the offset of every value; a line's column is then ONE slice of that text.
bool write_table(const std::string &file, const std::string &ext, int k, const Genome &g) {
    char col = ',', entry = ',';
    if (ext == "bed") { col = '\t'; entry = ':'; }
    else if (ext == "csv") { col = ','; entry = ' '; }
    FILE *f = fopen(file.c_str(), "wb");
    if (!f) return false;
    const std::vector<Span> spans = spans_of(g, 1u << 16);          // a line is ~200 bytes: 64 Ki lines per task
    const uint64_t K = (uint64_t)k;
    const bool ok = emit_blocks(f, spans.size(), [&](size_t idx, std::string &o) {
        const Span &sp = spans[idx];
        const kq_base *b = g.bases[sp.comp->segment].data();
        const uint64_t rows = sp.to - sp.from, vals = rows + K - 1;   // values from - (k - 1) .. to - 1, zeros before 0
        std::string joined[3];
        std::vector<uint32_t> off[3];
        for (int t = 0; t < 3; ++t) { joined[t].resize(vals * 11); off[t].resize(vals + 1); }
        char *w[3] = {&joined[0][0], &joined[1][0], &joined[2][0]};
        for (uint64_t i = 0; i < vals; ++i) {
            uint32_t v[3] = {0, 0, 0};
            if (sp.from + i >= K - 1) triple(b[sp.from + i - (K - 1)], v[0], v[1], v[2]);
            for (int t = 0; t < 3; ++t) {
                off[t][i] = (uint32_t)(w[t] - joined[t].data());
                w[t] = put_num(w[t], v[t]); *w[t]++ = entry;
            }
        }
        for (int t = 0; t < 3; ++t) off[t][vals] = (uint32_t)(w[t] - joined[t].data());
        size_t total = 0;
        for (uint64_t r = 0; r < rows; ++r)
            for (int t = 0; t < 3; ++t) total += off[t][r + K] - off[t][r];      // k values, k - 1 separators + the column separator
        total += rows * (sp.path->header.size() + 1 + 20 + 1);
        o.resize(total);
        char *p = &o[0];
        const std::string &name = sp.path->header;
        for (uint64_t r = 0; r < rows; ++r) {
            memcpy(p, name.data(), name.size()); p += name.size();
            *p++ = col; p = put_num(p, sp.abs_pos + sp.from + r);
            for (int t = 0; t < 3; ++t) {
                *p++ = col;
                const size_t len = off[t][r + K] - off[t][r] - 1;               // without the trailing separator
                memcpy(p, joined[t].data() + off[t][r], len); p += len;
            }
            *p++ = '\n';
        }
        o.resize((size_t)(p - o.data()));
    });
    return (fclose(f) == 0) && ok;
}

// Layout (kreeq-output.cpp:305-399): u8 k; u32 #paths; per path {u16 len, header, u32 #segments,
// per segment {u64 start, u64 len, u8 step}}; then per base three u32 (cov, F, B), segments in path order.
bool write_bkwig(const std::string &file, int k, const Genome &g) {
    Out o(file);
    if (!o.ok()) return false;
    o.raw<uint8_t>((uint8_t)k);
    o.raw<uint32_t>((uint32_t)g.paths.size());
    for (const Path &p : g.paths) {
        uint32_t nseg = 0;
        for (const Component &c : p.components) nseg += c.is_segment;
        o.raw<uint16_t>((uint16_t)p.header.size());
        o.put(p.header);
        o.raw<uint32_t>(nseg);
        uint64_t abs_pos = 0;
        for (const Component &c : p.components) {
            if (c.is_segment) {
                o.raw<uint64_t>(abs_pos);
                o.raw<uint64_t>(c.len);
                o.raw<uint8_t>(1);
            }
            abs_pos += c.len;
        }
    }
    FILE *f = o.f;
    o.flush();
    const std::vector<Span> spans = spans_of(g);
    const bool ok = emit_blocks(f, spans.size(), [&](size_t i, std::string &out) {   // 12 bytes per base, blocks in parallel
        const Span &sp = spans[i];
        const kq_base *b = g.bases[sp.comp->segment].data();
        out.resize((sp.to - sp.from) * 12);
        char *p = &out[0];
        for (uint64_t x = sp.from; x < sp.to; ++x, p += 12) {
            uint32_t v[3];
            triple(b[x], v[0], v[1], v[2]);
            memcpy(p, v, 12);
        }
    });
    return o.close() && ok;
}

} // namespace cli

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

inline void append_num(std::string &o, uint64_t v) {
    char tmp[20];
    int len = 0;
    do { tmp[len++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (len) o.push_back(tmp[--len]);
}

constexpr uint64_t BLOCK = 1u << 20;               // bases per formatting task
struct Span { const Path *path; const Component *comp; uint64_t abs_pos, from, to; };   // bases [from, to) of one segment

std::vector<Span> spans_of(const Genome &g) {
    std::vector<Span> v;
    for (const Path &p : g.paths) {
        uint64_t abs_pos = 0;
        for (const Component &c : p.components) {
            if (c.is_segment) {
                const uint64_t len = g.bases[c.segment].size();
                if (len == 0) v.push_back({&p, &c, abs_pos, 0, 0});
                for (uint64_t a = 0; a < len; a += BLOCK) v.push_back({&p, &c, abs_pos, a, std::min(len, a + BLOCK)});
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
        for (uint64_t x = sp.from; x < sp.to; ++x) {
            uint32_t cov, F, B;
            triple(b[x], cov, F, B);
            append_num(o, cov); o.push_back(','); append_num(o, F); o.push_back(','); append_num(o, B); o.push_back('\n');
        }
    });
    return (fclose(f) == 0) && ok;
}

// Each line repeats the last k values of the three tracks (zeros before the segment start), so a line is
// three joins over a ring of pre-formatted numbers.
bool write_table(const std::string &file, const std::string &ext, int k, const Genome &g) {
    char col = ',', entry = ',';
    if (ext == "bed") { col = '\t'; entry = ':'; }
    else if (ext == "csv") { col = ','; entry = ' '; }
    FILE *f = fopen(file.c_str(), "wb");
    if (!f) return false;
    struct Tok { char s[11]; uint8_t len; };
    auto fmt = [](uint32_t v, Tok &t) {
        char tmp[10];
        int len = 0;
        do { tmp[len++] = (char)('0' + v % 10); v /= 10; } while (v);
        t.len = (uint8_t)len;
        for (int i = 0; i < len; ++i) t.s[i] = tmp[len - 1 - i];
    };
    const std::vector<Span> spans = spans_of(g);
    const bool ok = emit_blocks(f, spans.size(), [&](size_t idx, std::string &o) {
        const Span &sp = spans[idx];
        const kq_base *b = g.bases[sp.comp->segment].data();
        std::vector<Tok> ring[3];
        for (auto &r : ring) { r.resize((size_t)k); for (Tok &t : r) { t.s[0] = '0'; t.len = 1; } }
        size_t head = 0;                            // slot of the oldest value
        const uint64_t warm = sp.from >= (uint64_t)(k - 1) ? sp.from - (uint64_t)(k - 1) : 0;
        for (uint64_t x = warm; x < sp.from; ++x) {   // the window a block inherits from the bases before it
            uint32_t v[3];
            triple(b[x], v[0], v[1], v[2]);
            for (int t = 0; t < 3; ++t) fmt(v[t], ring[t][head]);
            head = (head + 1) % (size_t)k;
        }
        o.reserve((sp.to - sp.from) * (size_t)(3 * k * 3 + 32));
        for (uint64_t x = sp.from; x < sp.to; ++x) {
            uint32_t v[3];
            triple(b[x], v[0], v[1], v[2]);
            for (int t = 0; t < 3; ++t) fmt(v[t], ring[t][head]);
            head = (head + 1) % (size_t)k;         // the newest value is now the last of the window
            o += sp.path->header; o.push_back(col); append_num(o, sp.abs_pos + x);
            for (int t = 0; t < 3; ++t) {
                o.push_back(col);
                for (int i = 0; i < k; ++i) {
                    const Tok &tk = ring[t][(head + (size_t)i) % (size_t)k];
                    o.append(tk.s, tk.len);
                    if (i + 1 < k) o.push_back(entry);
                }
            }
            o.push_back('\n');
        }
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
    for (const Path &p : g.paths)
        for (const Component &c : p.components) {
            if (!c.is_segment) continue;
            for (const kq_base &b : g.bases[c.segment]) {
                uint32_t v[3];
                triple(b, v[0], v[1], v[2]);
                o.put(reinterpret_cast<const char *>(v), 12);
            }
        }
    return o.close();
}

} // namespace cli

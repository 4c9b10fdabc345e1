// Per-base output files.  Same bytes as the reference's writers, built with a flat integer formatter and
// one buffered write per few MiB instead of per-value iostream calls.
//   .bed / .csvtable   DBG::printTable                   src/kreeq-output.cpp:138-241
//   .kwig              DBG::printTableCompressed         src/kreeq-output.cpp:243-303
//   .bkwig             DBG::printTableCompressedBinary   src/kreeq-output.cpp:305-399
#include "kreeq_cli.h"

#include <cstdio>
#include <cstring>

namespace cli {
namespace {

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
    Out o(file);
    if (!o.ok()) return false;
    o.num((uint64_t)k); o.put('\n');
    for (const Path &p : g.paths) {
        uint64_t abs_pos = 0;
        for (const Component &c : p.components) {
            if (c.is_segment) {
                o.put("fixedStep chrom=", 16); o.put(p.header);
                o.put(" start=", 7); o.num(abs_pos); o.put(" step=1\n", 8);
                for (const kq_base &b : g.bases[c.segment]) {
                    uint32_t cov, F, B;
                    triple(b, cov, F, B);
                    o.num(cov); o.put(','); o.num(F); o.put(','); o.num(B); o.put('\n');
                }
            }
            abs_pos += c.len;
        }
    }
    return o.close();
}

// Each line repeats the last k values of the three tracks (zeros before the segment start), so a line is
// three joins over a ring of pre-formatted numbers.
bool write_table(const std::string &file, const std::string &ext, int k, const Genome &g) {
    char col = ',', entry = ',';
    if (ext == "bed") { col = '\t'; entry = ':'; }
    else if (ext == "csv") { col = ','; entry = ' '; }
    Out o(file);
    if (!o.ok()) return false;
    struct Tok { char s[11]; uint8_t len; };
    auto fmt = [](uint32_t v, Tok &t) {
        char tmp[10];
        int len = 0;
        do { tmp[len++] = (char)('0' + v % 10); v /= 10; } while (v);
        t.len = (uint8_t)len;
        for (int i = 0; i < len; ++i) t.s[i] = tmp[len - 1 - i];
    };
    std::vector<Tok> ring[3];
    for (auto &r : ring) r.resize((size_t)k);
    for (const Path &p : g.paths) {
        uint64_t abs_pos = 0;
        for (const Component &c : p.components) {
            if (!c.is_segment) { abs_pos += c.len; continue; }
            for (auto &r : ring)
                for (Tok &t : r) { t.s[0] = '0'; t.len = 1; }
            size_t head = 0;                       // slot of the oldest value
            for (const kq_base &b : g.bases[c.segment]) {
                uint32_t v[3];
                triple(b, v[0], v[1], v[2]);
                for (int t = 0; t < 3; ++t) fmt(v[t], ring[t][head]);
                head = (head + 1) % (size_t)k;     // the newest value is now the last of the window
                o.put(p.header); o.put(col); o.num(abs_pos);
                o.room(3 * (size_t)k * 12 + 8);
                for (int t = 0; t < 3; ++t) {
                    o.buf[o.n++] = col;
                    for (int i = 0; i < k; ++i) {
                        const Tok &tk = ring[t][(head + (size_t)i) % (size_t)k];
                        memcpy(o.buf.data() + o.n, tk.s, tk.len);
                        o.n += tk.len;
                        if (i + 1 < k) o.buf[o.n++] = entry;
                    }
                }
                o.buf[o.n++] = '\n';
                ++abs_pos;
            }
        }
    }
    return o.close();
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

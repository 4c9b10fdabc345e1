// kq_db.cpp -- reader/writer of the .kreeq on-disk layout.
//
// A map file is the raw dump of phmap::parallel_flat_hash_map<uint64_t, V, std::hash, ..., N=8, NullMutex>
// (reference: gfalibs/include/parallel-hashmap/phmap_dump.h:52-67,104-124).  This file does not use
// phmap; it re-creates the bytes from the published layout rules so that the reference's
// phmap_load + find() work on our files:
//   hash      = hi64(key*K) + lo64(key*K), K = 0xde5fb9d2630458e9       (phmap_utils.h:66-77)
//   submap    = ((h>>8) ^ (h>>16) ^ (h>>24)) & 255                       (phmap.h:3840-3842)
//   ctrl byte = h & 0x7F, probe start (h>>7) & capacity, 16-byte groups   (phmap.h:357-364,144-168)
//   element sits in the first group of its probe sequence that had a free byte  (phmap.h:2172-2183)
#include "kq_host.h"

#include <cstdio>
#include <cstring>
#include <fstream>

namespace kq {

namespace {

constexpr uint64_t kVersion = 0xFFFFFFFFFFFFFFF5ULL;   // SIZE_MAX - 10, phmap_dump.h:47-48
constexpr int8_t kEmpty = -128, kSentinel = -1;
constexpr size_t kWidth = 16;                          // SSE2 Group::kWidth, phmap.h:405

struct Slot8  { uint64_t key; uint8_t fw[4], bw[4], cov; uint8_t pad[7]; };
struct Slot32 { uint64_t key; uint32_t fw[4], bw[4], cov; uint8_t pad[4]; };
static_assert(sizeof(Slot8) == 24 && sizeof(Slot32) == 48, "slot sizes of std::pair<const uint64_t, DBGkmer[32]>");

inline void fill(Slot8 &s, const Rec8Entry &e)   { std::memset(&s, 0, sizeof s); s.key = e.key; std::memcpy(s.fw, e.fw, 4); std::memcpy(s.bw, e.bw, 4); s.cov = e.cov; }
inline void fill(Slot32 &s, const Rec32Entry &e) { std::memset(&s, 0, sizeof s); s.key = e.key; std::memcpy(s.fw, e.fw, 16); std::memcpy(s.bw, e.bw, 16); s.cov = e.cov; }
inline void take(Rec8Entry &e, const Slot8 &s)   { e.key = s.key; std::memcpy(e.fw, s.fw, 4); std::memcpy(e.bw, s.bw, 4); e.cov = s.cov; }
inline void take(Rec32Entry &e, const Slot32 &s) { e.key = s.key; std::memcpy(e.fw, s.fw, 16); std::memcpy(e.bw, s.bw, 16); e.cov = s.cov; }

// smallest capacity 2^n-1 whose growth budget (capacity - capacity/8, phmap.h:576-588) holds n
inline uint64_t capacity_for(uint64_t n) {
    uint64_t cap = 1;
    while (cap - cap / 8 < n) cap = cap * 2 + 1;
    return cap;
}

template <class SlotT, class EntryT>
std::string write_map(const std::string &path, std::vector<EntryT> &entries) {
    // bucket by submap (counting sort keeps it O(n))
    std::vector<uint64_t> start(257, 0);
    std::vector<uint64_t> hashes(entries.size());
    for (size_t i = 0; i < entries.size(); ++i) { hashes[i] = phmap_hash(entries[i].key); ++start[phmap_submap(hashes[i]) + 1]; }
    for (int s = 0; s < 256; ++s) start[s + 1] += start[s];
    std::vector<uint32_t> order(entries.size());
    { std::vector<uint64_t> cur(start.begin(), start.end() - 1);
      for (size_t i = 0; i < entries.size(); ++i) order[cur[phmap_submap(hashes[i])]++] = (uint32_t)i; }

    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) return "cannot open " + path + " for writing";
    std::vector<char> out;
    auto put64 = [&](uint64_t v) { const char *p = reinterpret_cast<const char *>(&v); out.insert(out.end(), p, p + 8); };
    put64(256);                                                       // submap_count, phmap_dump.h:109-110
    std::vector<int8_t> ctrl; std::vector<SlotT> slots;
    for (int s = 0; s < 256; ++s) {
        const uint64_t n = start[s + 1] - start[s];
        put64(kVersion); put64(n);
        if (n == 0) { put64(0); continue; }                           // size_ == 0: nothing else is dumped
        const uint64_t cap = capacity_for(n);
        put64(cap);
        ctrl.assign(cap + kWidth + 1, kEmpty);                        // reset_ctrl, phmap.h:2286-2290
        ctrl[cap] = kSentinel;
        slots.assign(cap, SlotT());
        std::memset(slots.data(), 0, cap * sizeof(SlotT));
        for (uint64_t j = start[s]; j < start[s + 1]; ++j) {
            const uint64_t h = hashes[order[j]];
            uint64_t off = (h >> 7) & cap, step = 0, pos = 0;         // find_first_non_full, phmap.h:2172-2183
            for (;;) {
                bool placed = false;
                for (size_t i = 0; i < kWidth; ++i)
                    if (ctrl[off + i] == kEmpty) { pos = (off + i) & cap; placed = true; break; }
                if (placed) break;
                step += kWidth; off = (off + step) & cap;             // probe_seq::next, phmap.h:160-164
            }
            const int8_t h2 = (int8_t)(h & 0x7F);
            ctrl[pos] = h2;                                           // set_ctrl + mirrored byte, phmap.h:2264-2276
            ctrl[((pos - kWidth) & cap) + 1 + ((kWidth - 1) & cap)] = h2;
            fill(slots[pos], entries[order[j]]);
        }
        out.insert(out.end(), reinterpret_cast<char *>(ctrl.data()), reinterpret_cast<char *>(ctrl.data()) + ctrl.size());
        out.insert(out.end(), reinterpret_cast<char *>(slots.data()), reinterpret_cast<char *>(slots.data()) + cap * sizeof(SlotT));
        put64(cap - cap / 8 - n);                                     // growth_left, phmap.h:2292-2294
        if (out.size() > (8u << 20)) { if (std::fwrite(out.data(), 1, out.size(), f) != out.size()) { std::fclose(f); return "short write on " + path; } out.clear(); }
    }
    const bool ok = std::fwrite(out.data(), 1, out.size(), f) == out.size();
    if (std::fclose(f) != 0 || !ok) return "short write on " + path;
    return "";
}

template <class SlotT, class EntryT>
std::string read_map(const std::string &path, std::vector<EntryT> &outv) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) return "cannot open " + path;
    const std::streamsize sz = f.tellg();
    f.seekg(0);
    std::vector<char> buf((size_t)sz);
    if (sz && !f.read(buf.data(), sz)) return "cannot read " + path;
    size_t off = 0;
    auto get64 = [&](uint64_t &v) { if (off + 8 > buf.size()) return false; std::memcpy(&v, buf.data() + off, 8); off += 8; return true; };
    uint64_t nsub = 0;
    if (!get64(nsub) || nsub != 256) return path + ": submap count is not 256";
    for (int s = 0; s < 256; ++s) {
        uint64_t ver = 0, n = 0, cap = 0;
        if (!get64(ver)) return path + ": truncated";
        if (ver < kVersion) { n = ver; } else if (!get64(n)) return path + ": truncated";   // phmap_dump.h:77-83
        if (!get64(cap)) return path + ": truncated";
        if (n == 0) continue;
        const size_t nctrl = cap + kWidth + 1;
        if (off + nctrl + cap * sizeof(SlotT) + (ver >= kVersion ? 8 : 0) > buf.size()) return path + ": truncated submap";
        const int8_t *ctrl = reinterpret_cast<const int8_t *>(buf.data() + off); off += nctrl;
        const char *slots = buf.data() + off; off += cap * sizeof(SlotT);
        if (ver >= kVersion) off += 8;                                // growth_left
        uint64_t seen = 0;
        for (uint64_t i = 0; i < cap; ++i) {
            if (ctrl[i] < 0) continue;                                // IsFull(ctrl) <=> ctrl >= 0
            SlotT sl; std::memcpy(&sl, slots + i * sizeof(SlotT), sizeof(SlotT));
            EntryT e; take(e, sl); outv.push_back(e); ++seen;
        }
        if (seen != n) return path + ": submap size does not match its control bytes";
    }
    return "";
}

}  // namespace

uint64_t phmap_hash(uint64_t key) {
    const unsigned __int128 p = (unsigned __int128)key * 0xde5fb9d2630458e9ULL;
    return (uint64_t)(p >> 64) + (uint64_t)p;
}

std::string write_map8(const std::string &path, std::vector<Rec8Entry> &e)   { return write_map<Slot8, Rec8Entry>(path, e); }
std::string write_map32(const std::string &path, std::vector<Rec32Entry> &e) { return write_map<Slot32, Rec32Entry>(path, e); }
std::string read_map8(const std::string &path, std::vector<Rec8Entry> &o)    { return read_map<Slot8, Rec8Entry>(path, o); }
std::string read_map32(const std::string &path, std::vector<Rec32Entry> &o)  { return read_map<Slot32, Rec32Entry>(path, o); }

std::string write_index(const std::string &dir, int k) {
    std::ofstream f(dir + "/.index");
    if (!f) return "cannot write " + dir + "/.index";
    f << k << "\n" << 128 << std::endl;                               // src/kreeq-output.cpp:91
    return f ? "" : "cannot write " + dir + "/.index";
}

int read_index_k(const std::string &dir) {
    std::ifstream f(dir + "/.index");
    if (!f) return -1;
    std::string line;
    if (!std::getline(f, line)) return -1;
    try { return std::stoi(line); } catch (...) { return -1; }         // src/input.cpp:60-66
}

}  // namespace kq

// File inputs of the command line: FASTQ read batching and the -f genome (FASTA / FASTQ / GFA1).
#include "kreeq_cli.h"
#include "pgzip.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <unordered_map>

namespace cli {

std::string file_ext(const std::string &name) {     // gfalibs/include/functions.h:317-332
    const size_t dot = name.find_last_of('.');
    if (dot == std::string::npos) return "";
    const std::string ext = name.substr(dot + 1);
    if (ext == "gz") return file_ext(name.substr(0, dot)) + ".gz";
    return ext;
}

namespace {

// gzread passes plain files through untouched, so one reader serves both.
struct GzReader {
    gzFile f = nullptr;
    explicit GzReader(const std::string &path) {
        f = gzopen(path.c_str(), "rb");
        if (f) gzbuffer(f, 1u << 20);
    }
    ~GzReader() { if (f) gzclose(f); }
    long read(char *dst, size_t cap) { return gzread(f, dst, (unsigned)cap); }
};

} // namespace


namespace {

long env_long(const char *name, long dflt) { const char *e = getenv(name); const long v = e ? atol(e) : 0; return v > 0 ? v : dflt; }

struct Mapped {                                       // read-only mapping of a whole file
    const unsigned char *data = nullptr; size_t size = 0; int fd = -1;
    explicit Mapped(const std::string &path) {
        fd = open(path.c_str(), O_RDONLY);
        struct stat st;
        if (fd < 0 || fstat(fd, &st) != 0 || st.st_size <= 0) return;
        void *m = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) return;
        madvise(m, (size_t)st.st_size, MADV_SEQUENTIAL);
        data = (const unsigned char *)m; size = (size_t)st.st_size;
    }
    ~Mapped() { if (data) munmap((void *)data, size); if (fd >= 0) close(fd); }
};

// gzip files of at least KQ_CLI_PGZ_MIN bytes (default 4 MiB) go through the parallel decoder (pgzip.h),
// smaller ones through zlib.  KQ_CLI_PGZ_CHUNK sets the compressed chunk size (tests use tiny chunks).
bool wants_parallel_gz(const Mapped &m) {
    static const size_t pgz_min = (size_t)env_long("KQ_CLI_PGZ_MIN", 4 << 20);
    return m.data && m.size >= pgz_min && pgz::is_gzip(m.data, m.size);
}
size_t pgz_chunk() { static const size_t c = (size_t)env_long("KQ_CLI_PGZ_CHUNK", 0); return c; }

} // namespace

std::atomic<int> g_feed_share{1};
int feed_threads() {
    static const int forced = (int)env_long("KQ_CLI_THREADS", 0);
    if (forced) return forced;
    const int hw = (int)std::max(1u, std::min(std::thread::hardware_concurrency(), 32u));
    return std::max(2, hw / std::max(1, g_feed_share.load()));
}

bool read_file(const std::string &path, std::string &out, std::string &err) {
    {
        Mapped m(path);
        if (wants_parallel_gz(m)) {
            struct Append : pgz::Sink {
                std::string &out; explicit Append(std::string &o) : out(o) {}
                bool consume(const char *d, size_t n, bool) override { out.append(d, n); return true; }
            } sink(out);
            out.clear();
            std::string e;
            if (!pgz::inflate_parallel(m.data, m.size, feed_threads(), sink, e, nullptr, pgz_chunk())) { err = e + " in " + path; return false; }
            return true;
        }
        if (m.data && !(m.size >= 2 && m.data[0] == 0x1f && m.data[1] == 0x8b)) {   // plain file: one copy out of the mapping
            out.assign((const char *)m.data, m.size);
            return true;
        }
    }
    GzReader r(path);
    if (!r.f) { err = "cannot open " + path; return false; }
    out.clear();
    std::vector<char> buf(8u << 20);
    for (;;) {
        const long n = r.read(buf.data(), buf.size());
        if (n < 0) { err = "read error in " + path; return false; }
        if (n == 0) break;
        out.append(buf.data(), (size_t)n);
    }
    return true;
}

// getKmers (functions.h:69-122) consumes, per record: the header line, the sequence line (kept), the '+'
// line and the quality line, and only then appends the 'N' separator; if the file ends inside any of the
// last three lines the record's bases are kept but no 'N' follows.  This state machine does the same over
// arbitrary pieces of the file.
namespace {
struct FastqState {
    std::string batch; int line = 0; size_t batch_bytes; ReadSink &sink;
    FastqState(size_t bb, ReadSink &s) : batch_bytes(bb), sink(s) { fresh(); }
    void fresh() {                                  // a buffer to fill: a recycled one if the sink has any
        if (batch.capacity() < batch_bytes) {
            if (!sink.recycle(batch)) batch = std::string();
            batch.reserve(batch_bytes + (1u << 20));
        }
        batch.clear();
    }
    bool feed(const char *p, const char *end) {
        while (p < end) {
            const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
            const char *stop = nl ? nl : end;
            if (line == 1) batch.append(p, (size_t)(stop - p));
            if (!nl) break;                         // line continues in the next piece
            if (line == 3) {
                batch.push_back('N');
                if (batch.size() >= batch_bytes) {
                    if (!sink.take(std::move(batch))) return false;
                    fresh();
                }
            }
            line = (line + 1) & 3;
            p = nl + 1;
        }
        return true;
    }
    bool finish() { return batch.empty() || sink.take(std::move(batch)); }
};

// Uncompressed files are mapped and cut into one slice per core AT RECORD BOUNDARIES: newlines are counted in
// parallel first, so the line number (mod 4) of every slice start is known exactly -- no guessing where a
// FASTQ record starts -- and the slices are parsed concurrently.  Counting commutes, so slice order is free.
bool stream_plain_parallel(const char *data, size_t size, size_t batch_bytes, ReadSink &sink, std::mutex *sink_mu, size_t slice_default = (size_t)64 << 20) {
    static const size_t slice_env = [] { const char *e = getenv("KQ_CLI_SLICE"); const long v = e ? atol(e) : 0; return v > 0 ? (size_t)v : (size_t)0; }();
    const size_t slice_min = slice_env ? slice_env : slice_default;
    const size_t nt = std::max<size_t>(1, std::min<size_t>({(size_t)std::thread::hardware_concurrency(), (size_t)16, size / slice_min}));
    std::vector<size_t> cut(nt + 1), lines(nt, 0);
    for (size_t t = 0; t <= nt; ++t) cut[t] = size / nt * t;
    cut[nt] = size;
    {
        std::vector<std::thread> th;
        for (size_t t = 0; t < nt; ++t)
            th.emplace_back([&, t] {
                size_t n = 0;
                for (const char *p = data + cut[t], *e = data + cut[t + 1]; p < e;) {
                    const char *nl = (const char *)memchr(p, '\n', (size_t)(e - p));
                    if (!nl) break;
                    ++n; p = nl + 1;
                }
                lines[t] = n;
            });
        for (auto &x : th) x.join();
    }
    // first byte of the first record that STARTS in slice t (line number divisible by four)
    std::vector<size_t> start(nt + 1, size);
    start[0] = 0;
    size_t before = 0;
    for (size_t t = 1; t < nt; ++t) {
        before += lines[t - 1];
        size_t need = (4 - before % 4) % 4, pos = cut[t];   // newlines to pass before a record boundary
        if (need == 0 && pos > 0 && data[pos - 1] != '\n') need = 4;   // the slice starts inside the record's header line
        while (need && pos < size) {
            const char *nl = (const char *)memchr(data + pos, '\n', size - pos);
            if (!nl) { pos = size; break; }
            pos = (size_t)(nl - data) + 1; --need;
        }
        start[t] = need ? size : pos;
    }
    for (size_t t = nt - 1; t >= 1; --t) start[t] = std::min(start[t], start[t + 1]);   // slices without a boundary are empty
    std::atomic<bool> ok{true};
    struct Locked : ReadSink {                      // the caller's sink is not thread safe
        ReadSink &inner; std::mutex &mu;
        Locked(ReadSink &i, std::mutex &m) : inner(i), mu(m) {}
        bool operator()(const char *d, size_t l) override { std::lock_guard<std::mutex> lk(mu); return inner(d, l); }
        bool take(std::string &&b) override { std::lock_guard<std::mutex> lk(mu); return inner.take(std::move(b)); }
        bool recycle(std::string &o) override { std::lock_guard<std::mutex> lk(mu); return inner.recycle(o); }
    };
    std::mutex local_mu;
    std::vector<std::thread> th;
    for (size_t t = 0; t < nt; ++t)
        th.emplace_back([&, t] {
            if (start[t] >= start[t + 1]) return;
            Locked locked(sink, sink_mu ? *sink_mu : local_mu);
            FastqState st(batch_bytes, sink.thread_safe() ? sink : static_cast<ReadSink &>(locked));
            if (!st.feed(data + start[t], data + start[t + 1]) || !st.finish()) ok = false;
        });
    for (auto &x : th) x.join();
    return ok;
}
}  // namespace

bool stream_reads(const std::string &path, size_t batch_bytes, ReadSink &sink, std::string &err) {
    // plain file: map it and parse slices on all cores
    {
        const int fd = open(path.c_str(), O_RDONLY);
        if (fd < 0) { err = "cannot open " + path; return false; }
        struct stat st;
        unsigned char magic[2] = {0, 0};
        static const off_t par_min = [] { const char *e = getenv("KQ_CLI_PAR_MIN"); const long v = e ? atol(e) : 0; return v > 0 ? (off_t)v : (off_t)128 << 20; }();   // test knobs
        const bool have = fstat(fd, &st) == 0 && st.st_size >= par_min && pread(fd, magic, 2, 0) == 2;
        if (have && !(magic[0] == 0x1f && magic[1] == 0x8b)) {
            if (magic[0] != '@') { close(fd); return true; }           // loadKmers only takes FASTQ (fastx.h:95-118)
            void *m = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
            if (m != MAP_FAILED) {
                madvise(m, (size_t)st.st_size, MADV_SEQUENTIAL);
                const bool ok = stream_plain_parallel((const char *)m, (size_t)st.st_size, batch_bytes, sink, nullptr);
                munmap(m, (size_t)st.st_size); close(fd);
                if (!ok) err = "reading " + path + " was interrupted";
                return ok;
            }
        }
        close(fd);
    }
    {
        Mapped m(path);
        if (wants_parallel_gz(m)) {
            // Pieces of the decompressed file arrive in order; each is cut at its last record boundary (newline
            // count divisible by four -- the piece starts at a boundary by induction) and parsed in slices.
            struct Pieces : pgz::Sink {
                size_t batch_bytes; ReadSink &sink; int threads; bool first = true, fastq = true;
                Pieces(size_t bb, ReadSink &s, int t) : batch_bytes(bb), sink(s), threads(t) {}
                size_t split(const char *d, size_t n, bool) override {
                    if (!fastq) return n;
                    const size_t nt = std::max<size_t>(1, std::min<size_t>((size_t)threads, n >> 20));
                    std::vector<size_t> lines(nt, 0);
                    std::vector<std::thread> th;
                    for (size_t t = 0; t < nt; ++t)
                        th.emplace_back([&, t] {
                            size_t c = 0;
                            for (const char *p = d + n / nt * t, *e = t + 1 == nt ? d + n : d + n / nt * (t + 1); p < e;) {
                                const char *nl = (const char *)memchr(p, '\n', (size_t)(e - p));
                                if (!nl) break;
                                ++c; p = nl + 1;
                            }
                            lines[t] = c;
                        });
                    for (auto &x : th) x.join();
                    size_t total = 0;
                    for (size_t c : lines) total += c;
                    if (total < 4) return 0;
                    size_t pos = n;                      // walk back over the partial line and total % 4 whole lines
                    for (size_t back = total % 4 + 1; back; --back) {
                        const void *nl = memrchr(d, '\n', pos);
                        pos = (size_t)((const char *)nl - d);
                    }
                    return pos + 1;
                }
                bool consume(const char *d, size_t n, bool) override {
                    if (first && n) { first = false; fastq = d[0] == '@'; }    // loadKmers only takes FASTQ (fastx.h:95-118)
                    if (!fastq || !n) return true;
                    return stream_plain_parallel(d, n, batch_bytes, sink, nullptr, (size_t)16 << 20);
                }
            } pieces(batch_bytes, sink, feed_threads());
            std::string e;
            if (!pgz::inflate_parallel(m.data, m.size, feed_threads(), pieces, e, nullptr, pgz_chunk())) { err = e + " in " + path; return false; }
            return true;
        }
    }
    GzReader r(path);
    if (!r.f) { err = "cannot open " + path; return false; }
    std::vector<char> buf(16u << 20);
    FastqState st(batch_bytes, sink);
    bool first = true;
    for (;;) {
        const long n = r.read(buf.data(), buf.size());
        if (n < 0) { err = "read error in " + path; return false; }
        if (n == 0) break;
        if (first) {
            first = false;
            if (buf[0] != '@') return true;         // loadKmers only takes FASTQ (fastx.h:95-118)
        }
        if (!st.feed(buf.data(), buf.data() + n)) return false;
    }
    return st.finish();
}

namespace {

// traverseInSequence, gfalibs/src/gfa.cpp:90-304: newlines dropped, runs of N/n/X/x become gaps, anything
// else stays inside a segment.
struct ParsedRecord { Path path; std::vector<std::string> segments; };   // component.segment is local to the record

void parse_record(ParsedRecord &r, const std::string &header, const char *s, size_t n) {
    static const std::vector<uint8_t> gap = [] { std::vector<uint8_t> t(256, 0); t['N'] = t['n'] = t['X'] = t['x'] = 1; return t; }();
    r.path.header = header;
    std::string seq;
    seq.reserve(n);
    for (const char *p = s, *end = s + n; p < end;) {
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        const char *stop = nl ? nl : end;
        seq.append(p, (size_t)(stop - p));
        p = stop + 1;
    }
    const uint8_t *q = (const uint8_t *)seq.data();
    const size_t len = seq.size();
    size_t i = 0;
    while (i < len) {
        size_t j = i + 1;
        if (gap[q[i]]) {
            while (j < len && gap[q[j]]) ++j;
            r.path.components.push_back({false, i, j - i, 0});
        } else {
            while (j < len && !gap[q[j]]) ++j;
            r.path.components.push_back({true, i, j - i, r.segments.size()});
            if (i == 0 && j == len) r.segments.push_back(std::move(seq));      // the usual case: one segment, no copy
            else r.segments.emplace_back(seq, i, j - i);
        }
        i = j;
    }
}

void append_record(Genome &g, ParsedRecord &&r) {
    const size_t base = g.segments.size();
    for (Component &c : r.path.components)
        if (c.is_segment) c.segment += base;
    for (std::string &seg : r.segments) g.segments.push_back(std::move(seg));
    g.paths.push_back(std::move(r.path));
}

void add_record(Genome &g, const std::string &header, const char *s, size_t n) {
    ParsedRecord r;
    parse_record(r, header, s, n);
    append_record(g, std::move(r));
}

std::string first_word(const char *s, size_t n) {
    size_t e = 0;
    while (e < n && s[e] != ' ') ++e;
    return std::string(s, e);
}

std::vector<std::string> split(const std::string &s, char sep) {
    std::vector<std::string> out;
    size_t a = 0;
    for (;;) {
        const size_t b = s.find(sep, a);
        out.push_back(s.substr(a, b == std::string::npos ? std::string::npos : b - a));
        if (b == std::string::npos) break;
        a = b + 1;
    }
    return out;
}

// GFA1 as the validate path uses it: S lines carry sequences, P lines name paths over '+' segments;
// without P lines every segment is its own path.
bool load_gfa(const std::string &data, Genome &g, std::string &err) {
    std::unordered_map<std::string, std::string> seqs;
    std::vector<std::string> order;
    std::vector<std::vector<std::string>> plines;
    size_t a = 0;
    while (a < data.size()) {
        size_t b = data.find('\n', a);
        if (b == std::string::npos) b = data.size();
        const std::string ln = data.substr(a, b - a);
        a = b + 1;
        if (ln.empty()) continue;
        std::vector<std::string> f = split(ln, '\t');
        if (f[0] == "S" && f.size() >= 3) {
            if (!seqs.count(f[1])) order.push_back(f[1]);
            seqs[f[1]] = f[2];
        } else if (f[0] == "P" && f.size() >= 3) {
            plines.push_back(std::move(f));
        }
    }
    if (plines.empty()) {
        for (const std::string &name : order) add_record(g, name, seqs[name].data(), seqs[name].size());
        return true;
    }
    for (const auto &f : plines) {
        Path path;
        path.header = f[1];
        uint64_t pos = 0;
        for (const std::string &item : split(f[2], ',')) {
            if (item.empty()) continue;
            const char sign = item.back();
            const std::string name = item.substr(0, item.size() - 1);
            auto it = seqs.find(name);
            if (it == seqs.end()) { err = "GFA path " + f[1] + " names unknown segment " + name; return false; }
            if (sign != '+') { err = "GFA path " + f[1] + ": only '+' oriented segments are supported"; return false; }
            path.components.push_back({true, pos, it->second.size(), g.segments.size()});
            g.segments.push_back(it->second);
            pos += it->second.size();
        }
        g.paths.push_back(std::move(path));
    }
    return true;
}

} // namespace

bool load_genome(const std::string &path, Genome &g, std::string &err) {
    std::string data;
    if (!read_file(path, data, err)) return false;
    if (data.empty()) return true;
    if (data[0] == '>') {
        // header = first line up to the first blank ('>' may occur inside the rest of the header line);
        // sequence = everything up to the next '>' (src/input.cpp:206-228).  Records are found with memchr and
        // parsed on all cores (a chromosome-scale assembly is a handful of records of 10^8 bases each).
        struct Rec { std::string header; size_t body, len; };
        std::vector<Rec> recs;
        size_t pos = 1;
        while (pos < data.size()) {
            const char *nlp = (const char *)memchr(data.data() + pos, '\n', data.size() - pos);
            const size_t nl = nlp ? (size_t)(nlp - data.data()) : data.size();
            const char *gt = nl < data.size() ? (const char *)memchr(data.data() + nl, '>', data.size() - nl) : nullptr;
            const size_t next = gt ? (size_t)(gt - data.data()) : data.size();
            const size_t body = nl < data.size() ? nl + 1 : data.size();
            recs.push_back({first_word(data.data() + pos, nl - pos), body, next - body});
            pos = next + 1;
        }
        std::vector<ParsedRecord> parsed(recs.size());
        std::atomic<size_t> next_rec{0};
        auto work = [&] { for (size_t i; (i = next_rec++) < recs.size();) parse_record(parsed[i], recs[i].header, data.data() + recs[i].body, recs[i].len); };
        std::vector<std::thread> th;
        for (size_t t = 1; t < std::min<size_t>((size_t)feed_threads(), recs.size()); ++t) th.emplace_back(work);
        work();
        for (auto &x : th) x.join();
        for (ParsedRecord &r : parsed) append_record(g, std::move(r));
        return true;
    }
    if (data[0] == '@') {                           // src/input.cpp:230-258: strict 4-line records
        size_t pos = 0;
        auto take_line = [&](size_t &a, size_t &b) {
            if (pos >= data.size()) return false;
            a = pos;
            b = data.find('\n', pos);
            if (b == std::string::npos) b = data.size();
            pos = b + 1;
            return true;
        };
        size_t a, b;
        while (take_line(a, b)) {
            if (b == a) break;
            const std::string header = first_word(data.data() + a + 1, b - a - 1);
            size_t sa = 0, sb = 0, x, y;
            const bool has_seq = take_line(sa, sb);
            take_line(x, y);
            take_line(x, y);
            add_record(g, header, data.data() + sa, has_seq ? sb - sa : 0);
        }
        return true;
    }
    return load_gfa(data, g, err);
}

} // namespace cli

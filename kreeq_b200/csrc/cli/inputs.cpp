// File inputs of the command line: FASTQ read batching and the -f genome (FASTA / FASTQ / GFA1).
#include "kreeq_cli.h"

#include <zlib.h>

#include <cstring>
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

bool read_file(const std::string &path, std::string &out, std::string &err) {
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
// fixed-size chunks so a file never has to be resident.
bool stream_reads(const std::string &path, size_t batch_bytes, ReadSink &sink, std::string &err) {
    GzReader r(path);
    if (!r.f) { err = "cannot open " + path; return false; }
    std::vector<char> buf(16u << 20);
    std::string batch;
    batch.reserve(batch_bytes + (1u << 20));
    int line = 0;                                   // 0 header, 1 sequence, 2 plus, 3 quality
    bool first = true;
    for (;;) {
        const long n = r.read(buf.data(), buf.size());
        if (n < 0) { err = "read error in " + path; return false; }
        if (n == 0) break;
        if (first) {
            first = false;
            if (buf[0] != '@') return true;         // loadKmers only takes FASTQ (fastx.h:95-118)
        }
        const char *p = buf.data(), *end = p + n;
        while (p < end) {
            const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
            const char *stop = nl ? nl : end;
            if (line == 1) batch.append(p, (size_t)(stop - p));
            if (!nl) break;                         // line continues in the next chunk
            if (line == 3) {
                batch.push_back('N');
                if (batch.size() >= batch_bytes) {
                    if (!sink(batch.data(), batch.size())) return false;
                    batch.clear();
                }
            }
            line = (line + 1) & 3;
            p = nl + 1;
        }
    }
    if (!batch.empty() && !sink(batch.data(), batch.size())) return false;
    return true;
}

namespace {

bool is_gap(char c) { return c == 'N' || c == 'n' || c == 'X' || c == 'x'; }

// traverseInSequence, gfalibs/src/gfa.cpp:90-304: newlines dropped, runs of N/n/X/x become gaps, anything
// else stays inside a segment.
void add_record(Genome &g, const std::string &header, const char *s, size_t n) {
    Path path;
    path.header = header;
    std::string seq;
    seq.reserve(n);
    for (size_t i = 0; i < n; ++i)
        if (s[i] != '\n') seq.push_back(s[i]);
    size_t i = 0;
    while (i < seq.size()) {
        size_t j = i;
        if (is_gap(seq[i])) {
            while (j < seq.size() && is_gap(seq[j])) ++j;
            path.components.push_back({false, i, j - i, 0});
        } else {
            while (j < seq.size() && !is_gap(seq[j])) ++j;
            path.components.push_back({true, i, j - i, g.segments.size()});
            g.segments.emplace_back(seq, i, j - i);
        }
        i = j;
    }
    g.paths.push_back(std::move(path));
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
        // sequence = everything up to the next '>' (src/input.cpp:206-228)
        size_t pos = 1;
        while (pos < data.size()) {
            size_t nl = data.find('\n', pos);
            if (nl == std::string::npos) nl = data.size();
            const std::string header = first_word(data.data() + pos, nl - pos);
            size_t next = nl < data.size() ? data.find('>', nl) : std::string::npos;
            if (next == std::string::npos) next = data.size();
            const size_t body = nl < data.size() ? nl + 1 : data.size();
            add_record(g, header, data.data() + body, next - body);
            pos = next + 1;
        }
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

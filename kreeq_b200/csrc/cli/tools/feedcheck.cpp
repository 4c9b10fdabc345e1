// Test helper for the host feed path (no GPU needed): runs the command line's own input layer on a file.
//   kreeq-feedcheck inflate <file.gz> [threads] [chunk_bytes]   decompressed bytes to stdout, stats to stderr
//   kreeq-feedcheck reads <file>                                 what stream_reads hands to kq_count_push:
//        prints  batches bytes separators digest   (digest = order-independent sum of a hash of every
//        maximal N-free stretch, so it does not depend on how reads were grouped into batches)
//   kreeq-feedcheck genome <file>                                paths / segments / bases / digest of load_genome
#include "../kreeq_cli.h"
#include "../pgzip.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

static uint64_t stretch_hash(const char *p, size_t n) {
    uint64_t h = 0xcbf29ce484222325ull;
    for (size_t i = 0; i < n; ++i) { h ^= (unsigned char)p[i]; h *= 0x100000001b3ull; }
    return h * 0x9e3779b97f4a7c15ull + n;
}

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: kreeq-feedcheck inflate|reads|genome <file> ...\n"); return 2; }
    const std::string mode = argv[1], path = argv[2];
    const auto t0 = std::chrono::steady_clock::now();
    auto secs = [&] { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
    if (mode == "inflate") {
        const int fd = open(path.c_str(), O_RDONLY);
        struct stat st;
        if (fd < 0 || fstat(fd, &st) != 0) { perror("open"); return 1; }
        void *m = st.st_size ? mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0) : nullptr;
        if (st.st_size && m == MAP_FAILED) { perror("mmap"); return 1; }
        const bool quiet = getenv("KQ_FEEDCHECK_QUIET") != nullptr;
        struct Out : pgz::Sink {
            bool quiet; size_t n = 0, pieces = 0;
            bool consume(const char *d, size_t len, bool) override { n += len; ++pieces; return quiet || fwrite(d, 1, len, stdout) == len; }
        } out;
        out.quiet = quiet;
        std::string err; pgz::Stats s;
        const int threads = argc > 3 ? atoi(argv[3]) : cli::feed_threads();
        const size_t chunk = argc > 4 ? (size_t)atol(argv[4]) : 0;
        const bool ok = pgz::inflate_parallel((const unsigned char *)m, (size_t)st.st_size, threads, out, err, &s, chunk);
        fflush(stdout);
        fprintf(stderr, "{\"ok\": %s, \"threads\": %d, \"in_bytes\": %zu, \"out_bytes\": %zu, \"pieces\": %zu, \"chunks\": %zu, \"chained\": %zu, \"fallbacks\": %zu, \"members\": %zu, \"seconds\": %.4f, \"stages\": [%.3f, %.3f, %.3f, %.3f, %.3f], \"error\": \"%s\"}\n",
                ok ? "true" : "false", threads, (size_t)st.st_size, out.n, out.pieces, s.chunks, s.chained, s.fallbacks, s.members, secs(), s.t_find, s.t_decode, s.t_chain, s.t_resolve, s.t_hand, err.c_str());
        return ok ? 0 : 1;
    }
    if (mode == "reads") {
        struct Digest : cli::ReadSink {
            std::mutex mu; uint64_t batches = 0, bytes = 0, seps = 0, digest = 0;
            bool cheap = getenv("KQ_FEEDCHECK_CHEAP") != nullptr;      // throughput runs: count bytes only
            bool operator()(const char *d, size_t len) override {
                uint64_t s = 0, h = 0; size_t a = 0;
                if (cheap) { std::lock_guard<std::mutex> lk(mu); ++batches; bytes += len; return true; }
                for (size_t i = 0; i <= len; ++i)
                    if (i == len || d[i] == 'N') { if (i > a) h += stretch_hash(d + a, i - a); if (i < len) ++s; a = i + 1; }
                std::lock_guard<std::mutex> lk(mu);
                ++batches; bytes += len; seps += s; digest += h;
                return true;
            }
            bool thread_safe() const override { return getenv("KQ_FEEDCHECK_SERIAL") == nullptr; }
            // like the command line's queue: batches are taken over and handed back for reuse once consumed
            bool pool = getenv("KQ_FEEDCHECK_NOPOOL") == nullptr;
            std::vector<std::string> spare;
            bool take(std::string &&b) override {
                const bool r = (*this)(b.data(), b.size());
                if (pool) { std::string mine = std::move(b); std::lock_guard<std::mutex> lk(mu); if (spare.size() < 24) spare.push_back(std::move(mine)); }
                else { std::string gone = std::move(b); }
                return r;
            }
            bool recycle(std::string &out) override {
                std::lock_guard<std::mutex> lk(mu);
                if (spare.empty()) return false;
                out = std::move(spare.back()); spare.pop_back();
                return true;
            }
        } sink;
        std::string err;
        const size_t batch = argc > 3 ? (size_t)atol(argv[3]) : (size_t)128 << 20;
        if (!cli::stream_reads(path, batch, sink, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
        printf("%llu %llu %llu %016llx\n", (unsigned long long)sink.batches, (unsigned long long)sink.bytes, (unsigned long long)sink.seps, (unsigned long long)sink.digest);
        fprintf(stderr, "seconds %.4f\n", secs());
        return 0;
    }
    if (mode == "genome") {
        cli::Genome g; std::string err;
        if (!cli::load_genome(path, g, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
        uint64_t bases = 0, h = 0;
        for (const std::string &s : g.segments) { bases += s.size(); h = h * 31 + stretch_hash(s.data(), s.size()); }
        for (const cli::Path &p : g.paths) h = h * 31 + stretch_hash(p.header.data(), p.header.size()) + p.components.size();
        printf("%zu %zu %llu %016llx\n", g.paths.size(), g.segments.size(), (unsigned long long)bases, (unsigned long long)h);
        fprintf(stderr, "seconds %.4f\n", secs());
        return 0;
    }
    if (mode == "rewrite") {
        // kreeq-feedcheck rewrite <asm.fa> <in.bkwig> <out.kwig|.bed|.csvtable|.bkwig>: the (cov, F, B) rows of a
        // .bkwig laid over the genome's segments and written again by the command line's writers
        if (argc < 5) { fprintf(stderr, "rewrite needs <asm> <in.bkwig> <out>\n"); return 2; }
        cli::Genome g; std::string err;
        if (!cli::load_genome(path, g, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
        FILE *f = fopen(argv[3], "rb");
        if (!f) { perror("bkwig"); return 1; }
        uint8_t k = 0; uint32_t npaths = 0;
        bool ok = fread(&k, 1, 1, f) == 1 && fread(&npaths, 4, 1, f) == 1;
        for (uint32_t i = 0; ok && i < npaths; ++i) {
            uint16_t hl = 0; uint32_t ns = 0;
            ok = fread(&hl, 2, 1, f) == 1 && fseek(f, hl, SEEK_CUR) == 0 && fread(&ns, 4, 1, f) == 1 && fseek(f, 17L * ns, SEEK_CUR) == 0;
        }
        g.bases.resize(g.segments.size());
        for (const cli::Path &p : g.paths)
            for (const cli::Component &c : p.components) {
                if (!c.is_segment) continue;
                cli::BaseRows &b = g.bases[c.segment];
                b.resize(c.len);
                for (kq_base &e : b) {
                    uint32_t v[3];
                    ok = ok && fread(v, 4, 3, f) == 3;
                    e.cov = v[0]; e.fw = v[1]; e.bw = v[2]; e.isFw = 1;
                }
            }
        ok = ok && fgetc(f) == EOF;
        fclose(f);
        if (!ok) { fprintf(stderr, "bkwig does not match the genome\n"); return 1; }
        const std::string out = argv[4], ext = cli::file_ext(out);
        ok = ext == "kwig" ? cli::write_kwig(out, k, g) : ext == "bkwig" ? cli::write_bkwig(out, k, g) : cli::write_table(out, ext, k, g);
        return ok ? 0 : 1;
    }
    if (mode == "writebench") {                       // kreeq-feedcheck writebench <out.kwig|.bed|.bkwig> <bases> [k]: writer throughput
        const uint64_t n = argc > 3 ? (uint64_t)atoll(argv[3]) : 10000000;
        const int k = argc > 4 ? atoi(argv[4]) : 21;
        cli::Genome g;
        const uint64_t per = (n + 9) / 10;
        uint64_t x = 88172645463325252ull;
        for (int r = 0; r < 10; ++r) {
            cli::Path p; p.header = "chr" + std::to_string(r + 1);
            p.components.push_back({true, 0, per, g.segments.size()});
            g.segments.emplace_back();               // the writers only read the DBGbase rows
            cli::BaseRows b(per);
            for (auto &e : b) {
                x ^= x << 13; x ^= x >> 7; x ^= x << 17;
                e.cov = 20 + (uint32_t)(x % 21); e.fw = (uint32_t)((x >> 8) % 30); e.bw = (uint32_t)((x >> 16) % 30); e.isFw = (uint8_t)((x >> 24) & 1);
                if ((x >> 40) % 53 == 0) {            // now and then values of every width up to 2^32 - 1
                    const int w = (int)((x >> 50) % 33);
                    e.cov = w ? (uint32_t)(0xffffffffull >> (32 - w)) - (uint32_t)((x >> 30) & 1) : 0; e.fw = e.cov / 7; e.bw = e.cov / 1000;
                }
            }
            g.bases.push_back(std::move(b));
            g.paths.push_back(std::move(p));
        }
        const double t1 = secs();
        const std::string ext = cli::file_ext(path);
        const bool ok = ext == "kwig" ? cli::write_kwig(path, k, g) : ext == "bkwig" ? cli::write_bkwig(path, k, g) : cli::write_table(path, ext, k, g);
        printf("{\"ok\": %s, \"bases\": %llu, \"seconds\": %.4f}\n", ok ? "true" : "false", (unsigned long long)(per * 10), secs() - t1);
        return ok ? 0 : 1;
    }
    fprintf(stderr, "unknown mode %s\n", mode.c_str());
    return 2;
}

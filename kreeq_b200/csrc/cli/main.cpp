// kreeq validate / kreeq union — the reference's command line (src/main.cpp:50-330) driving the GPU
// library.  Flag names, defaults, stdout text and exit codes follow the reference; what differs is listed
// in DESIGN.md ("command line").
#include "kreeq_cli.h"

#include <getopt.h>
#include <sys/stat.h>

#include <cctype>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <deque>
#include <fstream>
#include <iostream>
#include <mutex>
#include <thread>

using namespace cli;

namespace {

int verbose_flag = 0, cmd_flag = 0;
std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();

void verbose(const std::string &msg) {            // Log::verbose, gfalibs/include/log.h
    if (!verbose_flag) return;
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    fprintf(stderr, "%s (done in %.3f s).\n", msg.c_str(), s);
}

bool is_number(const char *s) {
    if (!*s) return false;
    for (; *s; ++s)
        if (!isdigit((unsigned char)*s)) return false;
    return true;
}
bool is_int(const char *s) {                      // isInt: optional sign + digits
    if (*s == '-' || *s == '+') ++s;
    return is_number(s);
}
void require_file(const char *path) {             // ifFileExists, gfalibs/include/functions.h
    struct stat st;
    if (stat(path, &st) != 0) {
        std::cout << "Error - file does not exist: " << path << std::endl;   // on stdout, like the reference
        exit(1);
    }
}

[[noreturn]] void die(kq_multi *ctx, int rc) {
    fprintf(stderr, "%s\n", kq_multi_last_error(ctx));
    (void)rc;
    exit(EXIT_FAILURE);
}
#define CHECK(ctx, call) do { const int rc_ = (call); if (rc_ != KQ_OK) die(ctx, rc_); } while (0)

void print_help_validate() {
    printf("kreeq [command]\n");
    printf("\nOptions:\n");
    printf("\t-c --coverage-cutoff coverage cutoff.\n");
    printf("\t-d --database kreeq database to load.\n");
    printf("\t-f --input-sequence sequence input file (fasta,gfa1/2).\n");
    printf("\t-r --input-reads read input files (fastq).\n");
    printf("\t-k --kmer-length length of kmers.\n");
    printf("\t-o --out-format supported extensions:\n");
    printf("\t\t .kreeq dumps hashmaps to file for reuse.\n");
    printf("\t\t .bed .csvtable .kwig .bkwig per-base coverage and edge support.\n");
    printf("\t\t .hist k-mer coverage histogram.\n");
    printf("\t-t --tmp-prefix accepted for compatibility (no temporary files are written).\n");
    printf("\t-m --max-memory accepted for compatibility.\n");
    printf("\t-j --threads <n> accepted for compatibility.\n");
    printf("\t--gpu <n> first CUDA device to use (default: 0).\n");
    printf("\t--gpus <n> number of GPUs the k-mer table is sharded over (1, 2, 4, 8; 0 = all visible; default: 1).\n");
    printf("\t--expected-distinct <n> size hint for the k-mer table.\n");
    printf("\t--cmd print $0 to stdout.\n");
}

void print_help_union() {
    printf("kreeq union [options]\n");
    printf("\nOptions:\n");
    printf("\t-d --databases DBG databases to merge.\n");
    printf("\t-j --threads <n> accepted for compatibility.\n");
    printf("\t-o --out-format generates various kinds of outputs (currently supported: .kreeq).\n");
    printf("\t--gpu <n> first CUDA device to use (default: 0).\n");
    printf("\t--gpus <n> number of GPUs the merge is sharded over (default: 1).\n");
    printf("\t--cmd print $0 to stdout.\n");
}

// -r and union's -d take every following argument that is not a flag or a number (main.cpp:176-181).
void take_list(int argc, char **argv, std::vector<std::string> &dst) {
    optind--;
    for (; optind < argc && *argv[optind] != '-' && !is_int(argv[optind]); optind++) {
        require_file(argv[optind]);
        dst.push_back(argv[optind]);
    }
}

int parse(int argc, char **argv, Options &o) {
    const std::string mode = argv[1];
    if (mode == "validate") o.mode = 0;
    else if (mode == "union") o.mode = 1;
    else if (mode == "subgraph") {
        fprintf(stderr, "mode subgraph is not part of the GPU path (see DESIGN.md). Terminating\n");
        return EXIT_FAILURE;
    } else {
        fprintf(stderr, "mode %s does not exist. Terminating\n", argv[1]);
        return EXIT_FAILURE;
    }
    static struct option long_options[] = {
        {"coverage-cutoff", required_argument, 0, 'c'}, {"database", required_argument, 0, 'd'},
        {"databases", required_argument, 0, 'd'},       {"input-positions", required_argument, 0, 'p'},
        {"input-sequence", required_argument, 0, 'f'},  {"kmer-length", required_argument, 0, 'k'},
        {"search-depth", required_argument, 0, 0},      {"max-span", required_argument, 0, 0},
        {"out-format", required_argument, 0, 'o'},      {"input-reads", required_argument, 0, 'r'},
        {"tmp-prefix", required_argument, 0, 't'},      {"max-memory", required_argument, 0, 'm'},
        {"threads", required_argument, 0, 'j'},         {"gpu", required_argument, 0, 0},
        {"gpus", required_argument, 0, 0},              {"devices", required_argument, 0, 0},
        {"expected-distinct", required_argument, 0, 0}, {"verbose", no_argument, &verbose_flag, 1},
        {"cmd", no_argument, &cmd_flag, 1},             {"version", no_argument, 0, 'v'},
        {"help", no_argument, 0, 'h'},                  {0, 0, 0, 0}};
    const char *shorts = o.mode == 0 ? "-:c:d:f:k:o:p:r:t:m:j:v:h" : "-:d:j:o:h";
    for (;;) {
        int idx = 0;
        const int c = getopt_long(argc, argv, shorts, long_options, &idx);
        if (c == -1) break;
        switch (c) {
        case ':':
            fprintf(stderr, "option -%c is missing a required argument\n", optopt);
            return EXIT_FAILURE;
        case 0:
            if (!strcmp(long_options[idx].name, "gpu")) o.device = atoi(optarg);
            if (!strcmp(long_options[idx].name, "gpus")) {
                if (!is_number(optarg)) { fprintf(stderr, "input '%s' to option --gpus must be a number\n", optarg); return EXIT_FAILURE; }
                o.gpus = atoi(optarg);
            }
            if (!strcmp(long_options[idx].name, "devices")) {
                o.devices.clear();
                for (const char *p = optarg; *p;) { o.devices.push_back(atoi(p)); while (*p && *p != ',') ++p; if (*p == ',') ++p; }
            }
            if (!strcmp(long_options[idx].name, "expected-distinct")) o.expected = strtoull(optarg, nullptr, 10);
            break;
        case 'c':
            if (!is_number(optarg)) { fprintf(stderr, "input '%s' to option -%c must be a number\n", optarg, optopt); return EXIT_FAILURE; }
            o.cov_cutoff = (uint32_t)atoi(optarg);
            break;
        case 'd':
            if (o.mode == 1) take_list(argc, argv, o.dbs);
            else { require_file(optarg); o.dbs.push_back(optarg); }
            break;
        case 'f': require_file(optarg); o.sequence = optarg; break;
        case 'k':
            if (!is_number(optarg)) { fprintf(stderr, "input '%s' to option -%c must be a number\n", optarg, optopt); return EXIT_FAILURE; }
            o.k = atoi(optarg); o.k_given = true;
            break;
        case 'o': o.out = optarg; break;
        case 'p':
            fprintf(stderr, "-p (BED include list) is not supported by the GPU path.\n");
            return EXIT_FAILURE;
        case 'r': take_list(argc, argv, o.reads); break;
        case 't': o.prefix = optarg; break;
        case 'j': case 'm': break;
        case 'v':
            printf("kreeq (B200) v0.1, reference-compatible with kreeq v0.0.2\n");
            exit(0);
        case 'h':
            if (o.mode == 0) print_help_validate(); else print_help_union();
            exit(0);
        default: break;                            // positional arguments are ignored, as in the reference
        }
    }
    if (o.mode == 1 && o.dbs.size() < 2) {
        fprintf(stderr, "At least two databases required (-d).\n");
        return EXIT_FAILURE;
    }
    o.verbose = verbose_flag; o.cmd = cmd_flag;
    return -1;
}

void make_dir(const std::string &dir) { mkdir(dir.c_str(), 0777); }

// the k-mer table: one handle over --gpus N devices (shard g on device --gpu + g, or the ordinals of --devices)
int create(kq_multi **ctx, int k, const Options &o) {
    std::vector<int> dev = o.devices;
    int n = o.gpus;
    if (!dev.empty()) n = (int)dev.size();
    else if (n > 0) for (int g = 0; g < n; ++g) dev.push_back(o.device + g);
    return kq_multi_create(ctx, k, n, dev.empty() ? nullptr : dev.data(), o.expected);
}

void print_stats(kq_multi *ctx) {                   // DBG::DBstats, src/graph-builder.cpp:284-295
    kq_stats s;
    CHECK(ctx, kq_multi_stats_get(ctx, &s));
    std::cout << "DBG Summary statistics:\n"
              << "Total kmers: " << s.total << "\n"
              << "Unique kmers: " << s.unique << "\n"
              << "Distinct kmers: " << s.distinct << "\n"
              << "Missing kmers: " << s.missing << "\n"
              << "Total edges: " << s.edges << "\n";
}

double error_rate(uint64_t missing, uint64_t total, int k) {   // src/kreeq.cpp:36-40
    return 1 - pow(1 - (double)missing / total, (double)1 / k);
}

void print_qv(const kq_qv &qv, int k) {           // src/kreeq.cpp:78-106
    std::cout << "Missing" << "\t" << "Total" << "\t" << "QV" << "\t" << "Error" << "\t" << "k" << "\t" << "Method" << std::endl;
    const double me = error_rate(qv.missing, qv.kcount, k), mq = -10 * log10(me);
    std::cout << qv.missing << "\t" << qv.kcount << "\t" << mq << "\t" << me << "\t" << std::to_string(k) << "\t" << "Merqury" << std::endl;
    const double ke = error_rate(qv.missing + qv.edge_missing, qv.kcount, k), kv = -10 * log10(ke);
    std::cout << qv.missing + qv.edge_missing << "\t" << qv.kcount << "\t" << kv << "\t" << ke << "\t" << std::to_string(k) << "\t" << "Kreeq" << std::endl;
}

// Read files are parsed (and gunzipped) by one producer thread each, at most four at a time; batches go
// through a bounded queue to the caller's thread, which is the only one that talks to the library.  Counting
// commutes, so the order in which batches of different files arrive does not change the database.
struct BatchQueue {
    std::mutex mu; std::condition_variable can_put, can_get;
    std::deque<std::string> q; size_t producers = 0; std::string error;
    static constexpr size_t kDepth = 4;
    void put(std::string &&b) {
        std::unique_lock<std::mutex> lk(mu);
        can_put.wait(lk, [&] { return q.size() < kDepth; });
        q.push_back(std::move(b)); can_get.notify_one();
    }
    // consumed batches go back to the parsers: their pages are mapped already, a fresh 128 MB string is not
    std::vector<std::string> spare;
    void give_back(std::string &&b) {
        std::lock_guard<std::mutex> lk(mu);
        if (spare.size() < 24) spare.push_back(std::move(b));
    }
    bool take_spare(std::string &out) {
        std::lock_guard<std::mutex> lk(mu);
        if (spare.empty()) return false;
        out = std::move(spare.back()); spare.pop_back();
        return true;
    }
    bool get(std::string &b) {
        std::unique_lock<std::mutex> lk(mu);
        can_get.wait(lk, [&] { return !q.empty() || producers == 0; });
        if (q.empty()) return false;
        b = std::move(q.front()); q.pop_front(); can_put.notify_one();
        return true;
    }
};
struct QueueSink : ReadSink {
    BatchQueue &bq;
    explicit QueueSink(BatchQueue &b) : bq(b) {}
    bool operator()(const char *data, size_t len) override { bq.put(std::string(data, len)); return true; }
    bool take(std::string &&batch) override { bq.put(std::move(batch)); return true; }
    bool recycle(std::string &out) override { return bq.take_spare(out); }
    bool thread_safe() const override { return true; }     // the queue has its own lock
};

bool count_files(kq_multi *ctx, const std::vector<std::string> &files, std::string &err) {
    BatchQueue bq;
    std::mutex next_mu; size_t next = 0;
    const size_t nthreads = std::min<size_t>(files.size(), 4);
    bq.producers = nthreads;
    g_feed_share = (int)std::max<size_t>(1, nthreads);
    std::vector<std::thread> threads;
    for (size_t t = 0; t < nthreads; ++t)
        threads.emplace_back([&] {
            for (;;) {
                size_t i;
                { std::lock_guard<std::mutex> lk(next_mu); i = next++; }
                if (i >= files.size()) break;
                QueueSink sink(bq); std::string e;
                if (!stream_reads(files[i], 128u << 20, sink, e)) { std::lock_guard<std::mutex> lk(bq.mu); if (bq.error.empty()) bq.error = e; }
            }
            { std::lock_guard<std::mutex> lk(bq.mu); --bq.producers; }
            bq.can_get.notify_all();
        });
    std::string batch; int rc = KQ_OK;
    while (bq.get(batch)) {
        if (rc == KQ_OK) rc = kq_multi_count_push(ctx, batch.data(), batch.size());      // keep draining after an error
        bq.give_back(std::move(batch));
        batch = std::string();
    }
    for (auto &t : threads) t.join();
    g_feed_share = 1;
    if (rc != KQ_OK) die(ctx, rc);
    err = bq.error;
    return err.empty();
}

void write_index(const std::string &dir, int k) {  // kreeq-output.cpp:88-94
    std::ofstream ofs(dir + "/.index");
    ofs << k << "\n" << KQ_MAP_COUNT << std::endl;
}

// DBG::report, src/kreeq-output.cpp:34-136, for the extensions the validate/union path produces.
void report(kq_multi *ctx, int k, const Options &o, Genome *genome) {
    const std::string ext = o.out.empty() ? "stdout" : file_ext("." + o.out);
    const bool dotted = o.out.find('.') != std::string::npos || o.out.empty();
    if (dotted || ext == "kreeq") print_stats(ctx);
    verbose("Writing ouput: " + o.out);
    if (ext == "gfa" || ext == "gfa2" || ext == "gfa.gz" || ext == "gfa2.gz" || ext == "vcf") {
        fprintf(stderr, "Output .%s (sequence correction) is not part of the GPU path (see DESIGN.md).\n", ext.c_str());
        exit(EXIT_FAILURE);
    }
    if (o.mode == 0 && genome) {                  // validateSequences, src/kreeq.cpp:47-108
        verbose("Validating sequence");
        kq_qv qv = {0, 0, 0};
        genome->bases.resize(genome->segments.size());
        for (size_t s = 0; s < genome->segments.size(); ++s) {
            const std::string &seg = genome->segments[s];
            genome->bases[s].resize(seg.size());           // every row is written by the library
            CHECK(ctx, kq_multi_validate_segment(ctx, seg.data(), seg.size(), o.cov_cutoff, genome->bases[s].data(), &qv));
        }
        if (dotted) print_qv(qv, k);
    }
    bool good = true;
    if (ext == "bed" || ext == "csvtable") good = !genome || write_table(o.out, ext, k, *genome);
    else if (ext == "kwig") good = !genome || write_kwig(o.out, k, *genome);
    else if (ext == "bkwig") good = !genome || write_bkwig(o.out, k, *genome);
    else if (ext == "hist") {                      // Kmap::printHist, gfalibs/include/kmer.h:906-914
        size_t len = 0;
        CHECK(ctx, kq_multi_hist(ctx, nullptr, nullptr, &len));
        std::vector<uint64_t> cov(len), cnt(len);
        CHECK(ctx, kq_multi_hist(ctx, cov.data(), cnt.data(), &len));
        std::ofstream ofs(o.out);
        for (size_t i = 0; i < len; ++i) ofs << cov[i] << "\t" << cnt[i] << "\n";
        good = ofs.good();
    } else if (ext == "kreeq") {
        write_index(o.out, k);
    }
    if (!good) {
        fprintf(stderr, "Cannot write output file %s\n", o.out.c_str());
        exit(EXIT_FAILURE);
    }
}

int run_validate(Options &o) {                    // Input::read case 0, src/input.cpp:87-117
    const bool keep_db = o.out.find(".kreeq") != std::string::npos;
    kq_multi *ctx = nullptr;
    int k_used = o.k;
    if (!o.reads.empty()) {
        if (o.k < 1 || o.k > 32) { fprintf(stderr, "Invalid kmer length.\n"); return EXIT_FAILURE; }
        k_used = o.k;
        if (create(&ctx, o.k, o) != KQ_OK) die(nullptr, 0);
        verbose("Loading input reads.");
        std::string err;
        if (!count_files(ctx, o.reads, err)) { fprintf(stderr, "%s\n", err.c_str()); return EXIT_FAILURE; }
        CHECK(ctx, kq_multi_count_finish(ctx));
        verbose("Reads loaded.");
        if (keep_db) { make_dir(o.out); CHECK(ctx, kq_multi_db_save(ctx, o.out.c_str())); }
    } else {
        if (o.dbs.empty()) { fprintf(stderr, "Cannot load DBG input. Exiting.\n"); return EXIT_FAILURE; }   // loadGraph, src/input.cpp:56-74
        const int kdb = kq_db_read_k(o.dbs[0].c_str());
        if (kdb < 1) { fprintf(stderr, "Cannot load DBG input. Exiting.\n"); return EXIT_FAILURE; }
        if (o.k_given && o.k != kdb)
            fprintf(stderr, "Warning: -k %d ignored, database was built with k=%d.\n", o.k, kdb);
        else if (!o.k_given && kdb != 21)   // the reference constructs its DBG before it reads the database's k (src/input.cpp:88 vs :100)
            fprintf(stderr, "Note: database k=%d is used; the reference would validate with its default k=21 unless -k %d is given.\n", kdb, kdb);
        k_used = kdb;
        if (create(&ctx, kdb, o) != KQ_OK) die(nullptr, 0);
        CHECK(ctx, kq_multi_db_load(ctx, o.dbs[0].c_str()));
        verbose("DBG loaded.");
        if (keep_db && o.out != o.dbs[0]) { make_dir(o.out); CHECK(ctx, kq_multi_db_save(ctx, o.out.c_str())); }
    }
    Genome genome;
    bool have_genome = false;
    if (!o.sequence.empty()) {
        verbose("Loading input sequences");
        std::string err;
        if (!load_genome(o.sequence, genome, err)) { fprintf(stderr, "Stream not successful: %s.", o.sequence.c_str()); return EXIT_FAILURE; }
        verbose("Sequences loaded");
        have_genome = true;
    }
    report(ctx, k_used, o, have_genome ? &genome : nullptr);
    kq_multi_destroy(ctx);
    return EXIT_SUCCESS;
}

int run_union(Options &o) {                       // Input::read case 1, src/input.cpp:118-152
    verbose("Merging input databases.");
    int k = 0;
    for (const std::string &db : o.dbs) {
        const int kd = kq_db_read_k(db.c_str());
        if (kd < 0) { fprintf(stderr, "Cannot load DBG input. Exiting.\n"); return EXIT_FAILURE; }
        if (k == 0) k = kd;
        if (k != kd) { fprintf(stderr, "Cannot merge databases with different kmer length.\n"); return EXIT_FAILURE; }
    }
    if (k == 0 || k > 32) { fprintf(stderr, "Invalid kmer length.\n"); return EXIT_FAILURE; }
    kq_multi *ctx = nullptr;
    if (create(&ctx, k, o) != KQ_OK) die(nullptr, 0);
    verbose("DBG object generated. Merging.");
    std::vector<const char *> dirs;
    for (const std::string &db : o.dbs) dirs.push_back(db.c_str());
    CHECK(ctx, kq_multi_union(ctx, dirs.data(), (int)dirs.size(), 1));
    if (o.out.find(".kreeq") != std::string::npos) { make_dir(o.out); CHECK(ctx, kq_multi_db_save(ctx, o.out.c_str())); }
    report(ctx, k, o, nullptr);
    kq_multi_destroy(ctx);
    return EXIT_SUCCESS;
}

} // namespace

int main(int argc, char **argv) {
    if (argc == 1 || argc == 2) {
        printf("kreeq [mode] -h\nfor additional help.\n\nModes:\nvalidate\nunion\n");   // src/main.cpp:40-47, without subgraph
        return EXIT_SUCCESS;
    }
    Options o;
    const int rc = parse(argc, argv, o);
    if (rc >= 0) return rc;
    if (o.cmd) {
        for (int i = 0; i < argc; ++i) printf("%s ", argv[i]);
        printf("\n");
    }
    return o.mode == 0 ? run_validate(o) : run_union(o);
}

// kreeq command line on top of libkreeq_b200.so: same modes, flags, stdout and file formats as the
// reference's `kreeq validate` / `kreeq union` (src/main.cpp:50-330, src/input.cpp:76-186,
// src/kreeq-output.cpp:34-399).  Everything that touches k-mers goes through the C-ABI; this side only
// parses files, splits records into segments and formats the results.
#ifndef KREEQ_CLI_H
#define KREEQ_CLI_H

#include <atomic>
#include <cstdint>
#include <memory>
#include <string>
#include <utility>
#include <vector>

extern "C" {
#include "kreeq_b200.h"
}

namespace cli {

struct Options {                       // UserInputKreeq, include/input.h:11-33 + gfalibs/include/struct.h:20-40
    int mode = 0;                      // 0 validate, 1 union
    std::vector<std::string> reads;    // -r
    std::vector<std::string> dbs;      // -d
    std::string sequence;              // -f
    std::string out;                   // -o
    std::string prefix = ".";          // -t
    int k = 21;                        // -k
    bool k_given = false;
    uint32_t cov_cutoff = 0;           // -c
    int verbose = 0, cmd = 0;
    int device = 0;                    // --gpu (ours)
    int gpus = 1;                      // --gpus (ours): shards of the key space, 0 = all visible GPUs
    std::vector<int> devices;          // --devices a,b,.. (ours): explicit ordinal per shard (may repeat)
    uint64_t expected = 0;             // --expected-distinct (ours): table size hint
};

// One component of a path: a segment (maximal run without N/n/X/x) or a gap (gfalibs/src/gfa.cpp:170-203).
struct Component {
    bool is_segment;
    uint64_t start;                    // offset inside the record
    uint64_t len;
    size_t segment;                    // index into Genome::segments when is_segment
};
struct Path {
    std::string header;
    std::vector<Component> components;
};
// std::vector whose resize() leaves new elements uninitialised: the DBGbase arrays (16 bytes per base) are
// overwritten completely by kq_validate_segment, so zero-filling them first only costs a pass over memory.
template <typename T> struct UninitAlloc : std::allocator<T> {
    template <typename U> struct rebind { using other = UninitAlloc<U>; };
    UninitAlloc() = default;
    template <typename U> UninitAlloc(const UninitAlloc<U> &) {}
    template <typename U> void construct(U *p) noexcept { ::new (static_cast<void *>(p)) U; }
    template <typename U, typename... A> void construct(U *p, A &&...a) { ::new (static_cast<void *>(p)) U(std::forward<A>(a)...); }
};
using BaseRows = std::vector<kq_base, UninitAlloc<kq_base>>;

struct Genome {
    std::vector<Path> paths;           // file order (sortPathsByOriginal)
    std::vector<std::string> segments;
    std::vector<BaseRows> bases;       // DBGbase arrays, one per segment
};

// Whole-file reader that understands gzip and plain files alike (gfalibs/src/stream-obj.cpp:98-158).
bool read_file(const std::string &path, std::string &out, std::string &err);

// FASTQ read batching with getKmers' exact end-of-file behaviour (gfalibs/include/functions.h:69-122):
// sequence lines, each followed by 'N' iff the record's quality line ended with a newline.  Calls
// sink(batch) with batches of about `batch_bytes`; FASTA inputs ('>') yield nothing, as in loadKmers.
struct ReadSink {
    virtual bool operator()(const char *data, size_t len) = 0;
    // a finished batch the producer no longer needs: sinks that keep batches override this to avoid a copy
    virtual bool take(std::string &&batch) { return (*this)(batch.data(), batch.size()); }
    // a used buffer the sink can spare (keeps its capacity and its already-faulted pages); false if none
    virtual bool recycle(std::string &out) { (void)out; return false; }
    // true if several parser threads may call the sink at the same time (else calls are serialised)
    virtual bool thread_safe() const { return false; }
    virtual ~ReadSink() {}
};
bool stream_reads(const std::string &path, size_t batch_bytes, ReadSink &sink, std::string &err);

// Host threads one input file may use for decoding / slicing: all cores divided by the number of files being
// read at the same time (g_feed_share, set by the caller).
extern std::atomic<int> g_feed_share;
int feed_threads();

bool load_genome(const std::string &path, Genome &g, std::string &err);   // src/input.cpp:188-308

// writers, byte-identical with src/kreeq-output.cpp:138-399
bool write_table(const std::string &file, const std::string &ext, int k, const Genome &g);   // .bed .csvtable
bool write_kwig(const std::string &file, int k, const Genome &g);
bool write_bkwig(const std::string &file, int k, const Genome &g);

std::string file_ext(const std::string &name);   // getFileExt, gfalibs/src/functions.cpp

} // namespace cli
#endif

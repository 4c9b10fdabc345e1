// Parallel gzip decoding for the host feed path.  The reference reads .gz inputs through a
// single zlib stream per file (gfalibs/src/stream-obj.cpp:98-158, one thread, `sbumpc` per character); at
// ~0.3 GB/s per file that is two orders of magnitude below what the count kernels take.  A gzip member cannot be
// split at arbitrary bytes -- deflate blocks start at arbitrary BIT offsets and every block may copy from the
// previous 32 KiB -- so the file is cut into chunks that are decoded speculatively:
//   1. a worker looks for the first position in its chunk that parses as the start of a dynamic-Huffman block
//      (or a gzip member header) and decodes from there into 16-bit symbols: literals, or MARKERS naming a byte
//      of the 32 KiB window it could not see;
//   2. it stops at the first block boundary that is another chunk's starting point, so a chain of chunks that
//      starts at a known-good position proves every later starting point on it (false positives fall off);
//   3. windows are propagated along the chain (32 KiB per chunk, sequential), then all chunks replace their
//      markers and compute CRC-32s in parallel; per-member CRC and length are checked like gzread does.
// The output arrives in order as contiguous pieces.  Own inflate implementation (RFC 1951); zlib only for crc32.
#ifndef KREEQ_PGZIP_H
#define KREEQ_PGZIP_H

#include <cstddef>
#include <cstdint>
#include <string>

namespace pgz {

struct Sink {
    // How many leading bytes of this piece to take now; the rest is carried over in front of the next piece
    // (lets a FASTQ parser cut at record boundaries).  `last` pieces must be taken whole.
    virtual size_t split(const char *data, size_t n, bool last) { (void)data; (void)last; return n; }
    // Called from a helper thread, pieces in order, never concurrently.
    virtual bool consume(const char *data, size_t n, bool last) = 0;
    virtual ~Sink() {}
};

struct Stats {
    size_t chunks = 0;        // chunk decodes started
    size_t chained = 0;       // ... that ended up on the chain (their work was used)
    size_t fallbacks = 0;     // sequential re-decodes (speculation failed at a true boundary)
    size_t members = 0;       // gzip members seen
    size_t out_bytes = 0;
    double t_find = 0, t_decode = 0, t_chain = 0, t_resolve = 0, t_hand = 0;   // seconds per stage (decoder thread)
};

bool is_gzip(const unsigned char *p, size_t n);

// Decodes the whole (possibly multi-member) gzip file in `gz` with `threads` workers.  chunk_bytes = 0 picks
// the default compressed chunk size.  Returns false with `err` set on corrupt input or when the sink refuses.
bool inflate_parallel(const unsigned char *gz, size_t size, int threads, Sink &sink, std::string &err,
                      Stats *stats = nullptr, size_t chunk_bytes = 0);

}  // namespace pgz
#endif

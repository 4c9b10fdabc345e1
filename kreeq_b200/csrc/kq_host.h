// kq_host.h -- host-side pieces shared by the C-ABI implementation and the CLI:
// a small thread pool, the 2-bit packer and the .kreeq (phmap dump) codec.
#pragma once
#include <condition_variable>
#include <cstddef>
#include <cstdint>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace kq {

// ---- thread pool: parallel_for over task indices ---------------------------------------------
class Pool {
public:
    explicit Pool(int nthreads);
    ~Pool();
    int size() const { return (int)workers_.size() + 1; }
    // runs fn(i) for i in [0,n) on the workers and the calling thread; returns when all are done.  Calls from
    // several threads are serialised (fn must not call parallel_for on the same pool)
    void parallel_for(size_t n, const std::function<void(size_t)> &fn);
private:
    void worker();
    std::vector<std::thread> workers_;
    std::mutex mu_, call_mu_;
    std::condition_variable cv_, done_cv_;
    const std::function<void(size_t)> *fn_ = nullptr;
    size_t next_ = 0, n_ = 0, running_ = 0;
    uint64_t epoch_ = 0;
    bool stop_ = false;
};
Pool &global_pool();          // sized min(hardware_concurrency, KQ_THREADS env, 32)

// ---- 2-bit packing (include/kreeq_b200.h "packed-sequence stage API") --------------------------
// words of EACH array for a sequence of len bases, including the padding the kernels may touch
size_t packed_words(size_t len);
// pack ascii[0,len) -> bases/invalid (both packed_words(len) words).  Positions >= len are
// invalid with base 0.  ctoi semantics of gfalibs/include/kmer.h:97-114 (ACGTacgt valid).
void pack_sequence(const char *ascii, size_t len, uint64_t *bases, uint64_t *invalid, Pool *pool);

// ---- .kreeq codec (phmap_dump.h:52-124) ---------------------------------------------
struct Rec8Entry  { uint64_t key; uint8_t fw[4], bw[4], cov; };
struct Rec32Entry { uint64_t key; uint32_t fw[4], bw[4], cov; };

uint64_t phmap_hash(uint64_t key);                 // phmap_mix<8>(std::hash<uint64_t>)
inline unsigned phmap_submap(uint64_t h) { return (unsigned)(((h >> 8) ^ (h >> 16) ^ (h >> 24)) & 255); }

// write one parallel_flat_hash_map dump holding `entries` (any order); returns "" or an error
std::string write_map8(const std::string &path, std::vector<Rec8Entry> &entries);
std::string write_map32(const std::string &path, std::vector<Rec32Entry> &entries);
// read one dump, appending to out
std::string read_map8(const std::string &path, std::vector<Rec8Entry> &out);
std::string read_map32(const std::string &path, std::vector<Rec32Entry> &out);
// <dir>/.index: "<k>\n128\n" (src/kreeq-output.cpp:88-94)
std::string write_index(const std::string &dir, int k);
int read_index_k(const std::string &dir);          // <0 on error

}  // namespace kq

// kq_host.cpp -- thread pool and the host 2-bit packer.
#include "kq_host.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace kq {

// ---- pool ------------------------------------------------------------------------------------------

Pool::Pool(int nthreads) {
    for (int i = 1; i < nthreads; ++i) workers_.emplace_back([this] { worker(); });
}

Pool::~Pool() {
    { std::lock_guard<std::mutex> l(mu_); stop_ = true; }
    cv_.notify_all();
    for (auto &t : workers_) t.join();
}

void Pool::worker() {
    uint64_t seen = 0;
    std::unique_lock<std::mutex> l(mu_);
    for (;;) {
        cv_.wait(l, [&] { return stop_ || (epoch_ != seen && next_ < n_); });
        if (stop_) return;
        seen = epoch_;
        ++running_;
        while (next_ < n_) {
            const size_t i = next_++;
            l.unlock(); (*fn_)(i); l.lock();
        }
        if (--running_ == 0) done_cv_.notify_all();
    }
}

void Pool::parallel_for(size_t n, const std::function<void(size_t)> &fn) {
    if (n == 0) return;
    if (workers_.empty() || n == 1) { for (size_t i = 0; i < n; ++i) fn(i); return; }
    std::lock_guard<std::mutex> one_call(call_mu_);                   // concurrent callers (one thread per GPU) take turns
    std::unique_lock<std::mutex> l(mu_);
    fn_ = &fn; n_ = n; next_ = 0; ++epoch_;
    cv_.notify_all();
    ++running_;
    while (next_ < n_) {
        const size_t i = next_++;
        l.unlock(); fn(i); l.lock();
    }
    --running_;
    done_cv_.wait(l, [&] { return running_ == 0; });
    fn_ = nullptr; n_ = 0; next_ = 0;
}

Pool &global_pool() {
    static Pool *p = [] {
        int n = (int)std::thread::hardware_concurrency();
        if (const char *e = std::getenv("KQ_THREADS")) n = std::atoi(e);
        n = std::max(1, std::min(n, 32));
        return new Pool(n);
    }();
    return *p;
}

// ---- packer -------------------------------------------------------------------------------------------

size_t packed_words(size_t len) { return (len + 127) / 32 + 2; }

namespace {

struct Lut { uint8_t v[256]; Lut() { std::memset(v, 4, 256); v['A'] = v['a'] = 0; v['C'] = v['c'] = 1; v['G'] = v['g'] = 2; v['T'] = v['t'] = 3; } };
const Lut kLut;

// one 64-base block [i0, i0+64) clipped to len -> two base words + one invalid word
inline void pack_block_scalar(const unsigned char *a, size_t len, size_t i0, uint64_t &b0, uint64_t &b1, uint64_t &inv) {
    b0 = b1 = 0; inv = 0;
    const size_t n = i0 >= len ? 0 : std::min<size_t>(64, len - i0);
    for (size_t j = 0; j < n; ++j) {
        const uint8_t c = kLut.v[a[i0 + j]];
        const uint64_t code = c & 3, bad = c >> 2;
        if (j < 32) b0 |= (code & (bad - 1)) << (2 * j); else b1 |= (code & (bad - 1)) << (2 * (j - 32));
        inv |= bad << j;
    }
    if (n < 64) inv |= n ? (~0ULL << n) : ~0ULL;
}

#if defined(__x86_64__)
// 32 ASCII bytes -> 64 bits of 2-bit codes + 32 invalid bits.
// (c >> 1) & 3 maps A,C,G,T (either case) to 0,1,3,2; x ^ (x >> 1) swaps the last two.
__attribute__((target("avx2,bmi2")))
inline void pack32_avx2(const unsigned char *p, uint64_t &bases, uint32_t &inv) {
    const __m256i v = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(p));
    const __m256i up = _mm256_and_si256(v, _mm256_set1_epi8((char)0xDF));
    const __m256i ok = _mm256_or_si256(
        _mm256_or_si256(_mm256_cmpeq_epi8(up, _mm256_set1_epi8('A')), _mm256_cmpeq_epi8(up, _mm256_set1_epi8('C'))),
        _mm256_or_si256(_mm256_cmpeq_epi8(up, _mm256_set1_epi8('G')), _mm256_cmpeq_epi8(up, _mm256_set1_epi8('T'))));
    inv = ~(uint32_t)_mm256_movemask_epi8(ok);
    __m256i x = _mm256_and_si256(_mm256_srli_epi16(v, 1), _mm256_set1_epi8(3));
    x = _mm256_xor_si256(x, _mm256_and_si256(_mm256_srli_epi16(x, 1), _mm256_set1_epi8(1)));
    x = _mm256_and_si256(x, ok);                       // invalid bytes carry code 0
    alignas(32) uint64_t q[4];
    _mm256_store_si256(reinterpret_cast<__m256i *>(q), x);
    const uint64_t m = 0x0303030303030303ULL;
    bases = _pext_u64(q[0], m) | (_pext_u64(q[1], m) << 16) | (_pext_u64(q[2], m) << 32) | (_pext_u64(q[3], m) << 48);
}

__attribute__((target("avx2,bmi2")))
void pack_blocks_avx2(const unsigned char *a, size_t len, size_t blk0, size_t blk1, uint64_t *bases, uint64_t *invalid) {
    for (size_t b = blk0; b < blk1; ++b) {
        const size_t i0 = b * 64;
        if (i0 + 64 <= len) {
            uint32_t v0, v1;
            pack32_avx2(a + i0, bases[2 * b], v0);
            pack32_avx2(a + i0 + 32, bases[2 * b + 1], v1);
            invalid[b] = (uint64_t)v0 | ((uint64_t)v1 << 32);
        } else {
            pack_block_scalar(a, len, i0, bases[2 * b], bases[2 * b + 1], invalid[b]);
        }
    }
}
#endif

void pack_blocks_scalar(const unsigned char *a, size_t len, size_t blk0, size_t blk1, uint64_t *bases, uint64_t *invalid) {
    for (size_t b = blk0; b < blk1; ++b) pack_block_scalar(a, len, b * 64, bases[2 * b], bases[2 * b + 1], invalid[b]);
}

}  // namespace

void pack_sequence(const char *ascii, size_t len, uint64_t *bases, uint64_t *invalid, Pool *pool) {
    const size_t words = packed_words(len);
    const size_t nblk = words / 2;                      // 64-base blocks covering the base words
    const unsigned char *a = reinterpret_cast<const unsigned char *>(ascii);
#if defined(__x86_64__)
    static const bool fast = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("bmi2");
#else
    const bool fast = false;
#endif
    auto run = [&](size_t b0, size_t b1) {
#if defined(__x86_64__)
        if (fast) { pack_blocks_avx2(a, len, b0, b1, bases, invalid); return; }
#endif
        pack_blocks_scalar(a, len, b0, b1, bases, invalid);
    };
    const size_t grain = 1 << 14;                       // 1 Mi bases per task
    if (!pool || nblk <= grain) run(0, nblk);
    else {
        const size_t tasks = (nblk + grain - 1) / grain;
        pool->parallel_for(tasks, [&](size_t t) { run(t * grain, std::min(nblk, (t + 1) * grain)); });
    }
    if (words & 1) bases[words - 1] = 0;
    for (size_t w = nblk; w < words; ++w) invalid[w] = ~0ULL;   // unused upper half of the invalid array
}

}  // namespace kq

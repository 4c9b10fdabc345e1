// kq_device.cuh -- device-side data layout and helpers shared by the kernels.
//
// HBM layout (DESIGN.md "Data layout"):
//   main table T : cap x Slot (32 B, 32-B aligned == one DRAM sector), open addressing,
//                  linear probing (4 slots per 128-B line).  tag = key+1, 0 == empty, so a
//                  cudaMemset(0) clears the table (key+1 never wraps: the canonical key of a
//                  k<=32-mer is never 2^64-1).
//   side table S : cap x SideSlot (88 B), 64-bit counters.  Holds (a) the part of a record that
//                  does not fit the 16-bit lanes of a Slot (count path, KQ_SIDE_REMAINDER) and
//                  (b) literal maps32 entries of loaded / merged databases (KQ_SIDE_MAP32).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace kq {

constexpr uint32_t KQ_LIMIT = 65535u;          // instances a Slot counts itself (16-bit lanes stay exact)
constexpr uint32_t KQ_LARGEST = 4294967295u;   // LARGEST, include/kreeq.h:68

constexpr uint32_t F_HAS_SIDE    = 1u;         // logical record = slot + side remainder
constexpr uint32_t F_PLACEHOLDER = 2u;         // 8-bit map entry is a cov==255 placeholder; record lives in maps32

constexpr uint32_t KQ_SIDE_REMAINDER = 1u;
constexpr uint32_t KQ_SIDE_MAP32     = 2u;

struct __align__(32) Slot {
    unsigned long long tag;    // key + 1, 0 = empty
    uint32_t cov;
    uint32_t flags;
    unsigned long long fw;     // 4 x u16, lane w at bits [16w, 16w+16)
    unsigned long long bw;
};
static_assert(sizeof(Slot) == 32, "Slot must be one 32-byte sector");

struct __align__(8) SideSlot {
    unsigned long long tag;    // key + 1, 0 = empty
    uint32_t kind;             // KQ_SIDE_* bits
    uint32_t pad;
    unsigned long long cov;
    unsigned long long fw[4];
    unsigned long long bw[4];
};
static_assert(sizeof(SideSlot) == 88, "SideSlot layout");

struct Rec32 { uint32_t fw[4], bw[4], cov; };                 // == kq_rec32
struct Base  { uint32_t fw, bw, cov; uint8_t isFw, pad[3]; }; // == kq_base

// what every kernel needs to know about the resident database
struct TableView {
    Slot *T; unsigned long long cap;
    SideSlot *S; unsigned long long scap;
    unsigned long long *occ;       // [0] slots claimed in T, [1] slots claimed in S
    int *err;                      // sticky error word (1 T full, 2 S full, 4 inconsistent DB, 8 route overflow)
    int k;
    int shard_shift;               // log2(128 / nranks)
    int shard_rank;                // -1 = own everything
};

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
    x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
    x ^= x >> 33; return x;
}

#ifdef __CUDACC__

__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ bool owns(const TableView &tv, uint64_t key) {
    return tv.shard_rank < 0 || (int)((key & 127) >> tv.shard_shift) == tv.shard_rank;
}

// 64 bits of a little-endian bit stream starting at absolute bit position `bit`
__device__ __forceinline__ uint64_t load_bits(const uint64_t *__restrict__ W, uint64_t bit) {
    const uint64_t w = bit >> 6; const uint32_t s = (uint32_t)bit & 63u;
    const uint64_t lo = __ldg(W + w), hi = __ldg(W + w + 1);
    return s ? (lo >> s) | (hi << (64 - s)) : lo;
}
__device__ __forceinline__ uint32_t base_at(const uint64_t *__restrict__ B, uint64_t i) {
    return (uint32_t)(__ldg(B + (i >> 5)) >> (2 * (i & 31))) & 3u;
}
__device__ __forceinline__ uint32_t invalid_at(const uint64_t *__restrict__ M, uint64_t i) {
    return (uint32_t)(__ldg(M + (i >> 6)) >> (i & 63)) & 1u;
}

// reverse the order of the 32 2-bit groups of x
__device__ __forceinline__ uint64_t reverse_pairs(uint64_t x) {
    x = __brevll(x);
    return ((x & 0xAAAAAAAAAAAAAAAAULL) >> 1) | ((x & 0x5555555555555555ULL) << 1);
}

// One k-mer start position of a packed sequence: canonical key, orientation and neighbours.
// Same arithmetic as Kmap::hash (gfalibs/include/kmer.h:781-794): with LSB-first 2-bit packing
// the forward value IS the bit field [2p, 2p+2k); the reverse complement is the pair-reversed
// complement.
struct Kmer { uint64_t key; bool valid, isFw; uint32_t nxt, prv; /* 0..3, 4 = none */ };

__device__ __forceinline__ Kmer kmer_at(const uint64_t *__restrict__ B, const uint64_t *__restrict__ M,
                                        uint64_t p, int k) {
    Kmer r;
    const uint64_t win = load_bits(B, 2 * p);
    const uint64_t inv = load_bits(M, p);
    const uint64_t kmask = k == 32 ? ~0ULL : ((1ULL << (2 * k)) - 1);
    r.valid = (inv & ((1ULL << k) - 1)) == 0;           // k <= 32 < 64
    const uint64_t fw = win & kmask;
    const uint64_t rv = reverse_pairs(~win) >> (64 - 2 * k);
    r.isFw = fw < rv;                                    // tie => not forward, kmer.h:791
    r.key = r.isFw ? fw : rv;
    r.nxt = ((inv >> k) & 1) ? 4u : base_at(B, p + k);
    r.prv = (p == 0 || invalid_at(M, p - 1)) ? 4u : base_at(B, p - 1);
    return r;
}

// edge byte of DBG::hashSequences, src/graph-builder.cpp:98-112 (edgeBit::assign sets bit 7-e)
__device__ __forceinline__ uint32_t edge_byte(const Kmer &km) {
    uint32_t e = 0;
    if (km.isFw) { if (km.nxt <= 3) e |= 0x80u >> km.nxt;       if (km.prv <= 3) e |= 0x08u >> km.prv; }
    else         { if (km.prv <= 3) e |= 0x80u >> (3 - km.prv); if (km.nxt <= 3) e |= 0x08u >> (3 - km.nxt); }
    return e;
}

// ---- main table ---------------------------------------------------------------------------

__device__ __forceinline__ Slot *find_slot(const TableView &tv, uint64_t key) {
    const unsigned long long tag = key + 1;
    unsigned long long i = __umul64hi(mix64(key), tv.cap);
    for (unsigned long long n = 0; n < tv.cap; ++n) {
        const unsigned long long cur = ld_relaxed(&tv.T[i].tag);
        if (cur == tag) return &tv.T[i];
        if (cur == 0) return nullptr;
        if (++i == tv.cap) i = 0;
    }
    return nullptr;
}

// find the slot of key, claiming an empty one if needed; *is_new reports a claim
__device__ __forceinline__ Slot *find_or_claim(const TableView &tv, uint64_t key, bool *is_new) {
    const unsigned long long tag = key + 1;
    unsigned long long i = __umul64hi(mix64(key), tv.cap);
    *is_new = false;
    for (unsigned long long n = 0; n < tv.cap; ++n) {
        unsigned long long cur = ld_relaxed(&tv.T[i].tag);
        if (cur == 0) {
            cur = atomicCAS(&tv.T[i].tag, 0ULL, tag);
            if (cur == 0) { *is_new = true; return &tv.T[i]; }
        }
        if (cur == tag) return &tv.T[i];
        if (++i == tv.cap) i = 0;
    }
    atomicOr(tv.err, 1);
    return nullptr;
}

// ---- side table -----------------------------------------------------------------------------

__device__ __forceinline__ SideSlot *side_find(const TableView &tv, uint64_t key) {
    const unsigned long long tag = key + 1;
    unsigned long long i = __umul64hi(mix64(key ^ 0x9E3779B97F4A7C15ULL), tv.scap);
    for (unsigned long long n = 0; n < tv.scap; ++n) {
        const unsigned long long cur = ld_relaxed(&tv.S[i].tag);
        if (cur == tag) return &tv.S[i];
        if (cur == 0) return nullptr;
        if (++i == tv.scap) i = 0;
    }
    return nullptr;
}

__device__ __forceinline__ SideSlot *side_find_or_claim(const TableView &tv, uint64_t key) {
    const unsigned long long tag = key + 1;
    unsigned long long i = __umul64hi(mix64(key ^ 0x9E3779B97F4A7C15ULL), tv.scap);
    for (unsigned long long n = 0; n < tv.scap; ++n) {
        unsigned long long cur = ld_relaxed(&tv.S[i].tag);
        if (cur == 0) {
            cur = atomicCAS(&tv.S[i].tag, 0ULL, tag);
            if (cur == 0) { atomicAdd(&tv.occ[1], 1ULL); return &tv.S[i]; }
        }
        if (cur == tag) return &tv.S[i];
        if (++i == tv.scap) i = 0;
    }
    atomicOr(tv.err, 2);
    return nullptr;
}

__device__ __forceinline__ uint32_t clamp32(unsigned long long v) {
    return v > KQ_LARGEST ? KQ_LARGEST : (uint32_t)v;
}

// lookup idiom of src/kreeq.cpp:150-166: returns 1 found, 0 absent, -1 placeholder without maps32 entry
__device__ __forceinline__ int lookup(const TableView &tv, uint64_t key, Rec32 &r) {
    const Slot *s = find_slot(tv, key);
    if (!s) return 0;
    // one 32-byte sector read
    const uint4 lo = __ldcg(reinterpret_cast<const uint4 *>(s));
    const uint4 hi = __ldcg(reinterpret_cast<const uint4 *>(s) + 1);
    const uint32_t cov = lo.z, flags = lo.w;
    const unsigned long long fw = ((unsigned long long)hi.y << 32) | hi.x, bw = ((unsigned long long)hi.w << 32) | hi.z;
    if (flags & F_PLACEHOLDER) {
        const SideSlot *q = side_find(tv, key);
        if (!q || !(q->kind & KQ_SIDE_MAP32)) return -1;
#pragma unroll
        for (int w = 0; w < 4; ++w) { r.fw[w] = clamp32(q->fw[w]); r.bw[w] = clamp32(q->bw[w]); }
        r.cov = clamp32(q->cov);
        return 1;
    }
    unsigned long long c = cov, f[4], b[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) { f[w] = (fw >> (16 * w)) & 0xFFFF; b[w] = (bw >> (16 * w)) & 0xFFFF; }
    if (flags & F_HAS_SIDE) {
        const SideSlot *q = side_find(tv, key);
        if (q) {
            c += q->cov;
#pragma unroll
            for (int w = 0; w < 4; ++w) { f[w] += q->fw[w]; b[w] += q->bw[w]; }
        }
    }
#pragma unroll
    for (int w = 0; w < 4; ++w) { r.fw[w] = clamp32(f[w]); r.bw[w] = clamp32(b[w]); }
    r.cov = clamp32(c);
    return 1;
}

// add `c` instances carrying per-lane edge increments (each lane increment <= c) to key's record.
// Exactness argument: the cov atomic is issued first and decides how many of the c
// instances the Slot takes (at most KQ_LIMIT over its lifetime), so no 16-bit lane can exceed
// KQ_LIMIT; the rest goes to 64-bit counters in the side table.
__device__ __forceinline__ void accumulate(const TableView &tv, Slot *s, uint64_t key, uint32_t c,
                                           const uint32_t fwi[4], const uint32_t bwi[4]) {
    const uint32_t old = atomicAdd(&s->cov, c);
    uint32_t m = old >= KQ_LIMIT ? 0u : min(c, KQ_LIMIT - old);      // instances the slot keeps
    unsigned long long fm = 0, bm = 0; uint32_t fs[4], bs[4]; bool spill = m < c;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
        const uint32_t a = min(fwi[w], m), b = min(bwi[w], m);
        fm |= (unsigned long long)a << (16 * w); bm |= (unsigned long long)b << (16 * w);
        fs[w] = fwi[w] - a; bs[w] = bwi[w] - b;
    }
    if (fm) atomicAdd(&s->fw, fm);
    if (bm) atomicAdd(&s->bw, bm);
    if (spill) {
        atomicSub(&s->cov, c - m);
        SideSlot *q = side_find_or_claim(tv, key);
        if (!q) return;
        atomicOr(&q->kind, KQ_SIDE_REMAINDER);
        atomicAdd(&q->cov, (unsigned long long)(c - m));
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            if (fs[w]) atomicAdd(&q->fw[w], (unsigned long long)fs[w]);
            if (bs[w]) atomicAdd(&q->bw[w], (unsigned long long)bs[w]);
        }
        __threadfence();
        atomicOr(&s->flags, F_HAS_SIDE);
    }
}

#endif  // __CUDACC__
}  // namespace kq

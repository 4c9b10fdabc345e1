// kq_device.cuh -- device-side data layout and helpers shared by the kernels.
//
// HBM layout (DESIGN.md "Data layout"):
//   main table T : nslices x 2^slice_bits Slots of 16 B.  A key hashes to ONE slice and is linearly
//                  probed inside it only, so that the count pipeline can bring one slice (<= 32 MiB)
//                  into L2 and hit it with every tuple that belongs there (random L2-resident traffic is
//                  ~8x cheaper than random HBM traffic on B200, profiles/r01_gups_*.log).
//                  Slot = { tag = key + 1 (0 = empty), cnt = [placeholder:1][cov:7][bw3..bw0, fw3..fw0 : 8 x 7 bit] }.
//                  An empty slot is all zero, so cudaMemset(0) clears the table; key + 1 never wraps (the
//                  canonical key of a k <= 32-mer is never 2^64-1).  Claim + first count is one 128-bit CAS,
//                  every later instance one 64-bit CAS on cnt (exact: a 7-bit lane can never overflow
//                  because the CAS computes the new word from the current one).
//   side table R : what does not fit the 7-bit lanes ("remainder", 64-bit counters); logical record =
//                  slot lanes + remainder.  Only consulted when a lane is saturated (== 127).
//   side table M : literal maps32 entries (DBGkmer32) of loaded / merged databases, looked up for slots
//                  whose placeholder bit is set (the reference's cov == 255 entries).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace kq {

constexpr uint32_t KQ_LANE_MAX = 127u;         // a slot lane counts to 127, the rest lives in R
constexpr uint32_t KQ_REM_FLOOR = 64u;         // counted slots that own a remainder never drop below this coverage
constexpr uint32_t KQ_LARGEST = 4294967295u;   // LARGEST, include/kreeq.h:68
constexpr int KQ_MAX_SLICE_BITS = 21;          // a table of up to 2^21 slots (32 MiB, L2 resident) is ONE slice
constexpr int KQ_PIECE_BITS = 12;              // larger tables: slices of 2^12 slots = 64 KiB, one CTA's shared memory (measured best; 2^9: one warp each)
constexpr int KQ_MAX_COARSE = 4096;            // bins of the level-1 scatter at most

constexpr unsigned long long CNT_PLACEHOLDER = 1ull << 63;
constexpr unsigned long long CNT_COV_ONE = 1ull << 56;

struct __align__(16) Slot {
    unsigned long long tag;    // key + 1, 0 = empty
    unsigned long long cnt;    // packed counters, see above
};
static_assert(sizeof(Slot) == 16, "Slot is 16 bytes");

struct __align__(8) SideSlot {
    unsigned long long tag;    // key + 1, 0 = empty
    unsigned long long cov;
    unsigned long long fw[4];
    unsigned long long bw[4];
};
static_assert(sizeof(SideSlot) == 80, "SideSlot layout");

struct Rec32 { uint32_t fw[4], bw[4], cov; };                 // == kq_rec32
struct Base  { uint32_t fw, bw, cov; uint8_t isFw, pad[3]; }; // == kq_base

// what every kernel needs to know about the resident database
struct TableView {
    Slot *T; unsigned long long nslices; int slice_bits;
    int cshift;                    // count pipeline: 2^cshift consecutive slices form one bin of the level-1 scatter
    SideSlot *R; unsigned long long rcap;      // remainders
    SideSlot *M; unsigned long long mcap;      // literal maps32 entries
    unsigned long long *occ;       // [0] slots claimed in T, [1] in R, [2] in M
    int *err;                      // sticky error word (1 T full, 2 side full, 4 inconsistent DB, 8 route overflow)
    int k;
    int shard_shift;               // log2(128 / nranks)
    int shard_rank;                // -1 = own everything
};

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
    x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
    x ^= x >> 33; return x;
}

// home slot of a key inside its slice: one multiplication, bits 38.. of the product (they depend on every bit of a
// key of up to 60 bits and are independent of the slice hash, which uses another multiplier)
__host__ __device__ __forceinline__ uint64_t home_hash(uint64_t key) { return (key * 0xD6E8FEB86659FD93ULL) >> 38; }

__host__ __device__ __forceinline__ uint32_t cnt_cov(unsigned long long c) { return (uint32_t)(c >> 56) & 127u; }
__host__ __device__ __forceinline__ uint32_t cnt_lane(unsigned long long c, int i) { return (uint32_t)(c >> (7 * i)) & 127u; }
// true if a remainder record may exist: the count path moves 63 of a full slot's 127 to R at a time and so
// never leaves such a slot below KQ_REM_FLOOR; load / union fill fields to 127 and put the rest in R
__host__ __device__ __forceinline__ bool cnt_saturated(unsigned long long c) {
    c &= ~CNT_PLACEHOLDER;
    if (cnt_cov(c) >= KQ_REM_FLOOR) return true;
    unsigned long long t = c & (c >> 1); t &= t >> 2; t &= (t >> 3) & (c >> 6);    // bit 7i set iff bits 7i..7i+6 all set
    return (t & 0x0102040810204081ULL) != 0;
}

#ifdef __CUDACC__

__device__ __forceinline__ Slot ld_slot(const Slot *p) {
    Slot s;
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(s.tag), "=l"(s.cnt) : "l"(p) : "memory");
    return s;
}
__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// 128-bit compare-and-swap of a whole slot; returns the previous content
__device__ __forceinline__ Slot cas_slot(Slot *p, const Slot &expect, const Slot &desired) {
    Slot o;
    asm volatile("{\n .reg .b128 c, s, d;\n mov.b128 c, {%2, %3};\n mov.b128 s, {%4, %5};\n"
                 " atom.relaxed.gpu.global.cas.b128 d, [%6], c, s;\n mov.b128 {%0, %1}, d;\n}"
                 : "=l"(o.tag), "=l"(o.cnt) : "l"(expect.tag), "l"(expect.cnt), "l"(desired.tag), "l"(desired.cnt), "l"(p) : "memory");
    return o;
}

__device__ __forceinline__ bool owns(const TableView &tv, uint64_t key) {
    return tv.shard_rank < 0 || (int)((key & 127) >> tv.shard_shift) == tv.shard_rank;
}

// Slice of a key: multiplicative hash, top 32 bits scaled to nslices (a handful of instructions: the
// histogram and scatter kernels evaluate it for every k-mer instance).  The home slot inside the slice
// comes from home_hash(key), which only the kernels that touch the table compute.
__host__ __device__ __forceinline__ unsigned long long slice_of_key(uint64_t key, unsigned long long nslices) {
    return ((unsigned long long)(uint32_t)((key * 0x9E3779B97F4A7C15ULL) >> 32) * nslices) >> 32;
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

__device__ __forceinline__ Kmer kmer_from_window(uint64_t win, uint64_t inv, uint32_t nxt, uint32_t prv, int k) {
    Kmer r;
    const uint64_t kmask = k == 32 ? ~0ULL : ((1ULL << (2 * k)) - 1);
    r.valid = (inv & ((1ULL << k) - 1)) == 0;           // k <= 32 < 64
    const uint64_t fw = win & kmask;
    const uint64_t rv = reverse_pairs(~win) >> (64 - 2 * k);
    r.isFw = fw < rv;                                    // tie => not forward, kmer.h:791
    r.key = r.isFw ? fw : rv;
    r.nxt = nxt; r.prv = prv;
    return r;
}

__device__ __forceinline__ Kmer kmer_at(const uint64_t *__restrict__ B, const uint64_t *__restrict__ M,
                                        uint64_t p, int k) {
    const uint64_t win = load_bits(B, 2 * p);
    const uint64_t inv = load_bits(M, p);
    const uint32_t nxt = ((inv >> k) & 1) ? 4u : base_at(B, p + k);
    const uint32_t prv = (p == 0 || invalid_at(M, p - 1)) ? 4u : base_at(B, p - 1);
    return kmer_from_window(win, inv, nxt, prv, k);
}

// edge byte of DBG::hashSequences, src/graph-builder.cpp:98-112 (edgeBit::assign sets bit 7-e)
__device__ __forceinline__ uint32_t edge_byte(const Kmer &km) {
    uint32_t e = 0;
    if (km.isFw) { if (km.nxt <= 3) e |= 0x80u >> km.nxt;       if (km.prv <= 3) e |= 0x08u >> km.prv; }
    else         { if (km.prv <= 3) e |= 0x80u >> (3 - km.prv); if (km.nxt <= 3) e |= 0x08u >> (3 - km.nxt); }
    return e;
}

// cnt increment of ONE k-mer instance with edge byte e: cov + 1, lane i + bit (7 - i) of e
__device__ __forceinline__ unsigned long long inc_of_edge(uint32_t e) {
    unsigned long long inc = CNT_COV_ONE;
#pragma unroll
    for (int i = 0; i < 8; ++i) inc |= (unsigned long long)((e >> (7 - i)) & 1u) << (7 * i);
    return inc;
}

// ---- side tables ------------------------------------------------------------------------------------

__device__ __forceinline__ SideSlot *side_find(SideSlot *S, unsigned long long scap, uint64_t key) {
    const unsigned long long tag = key + 1;
    unsigned long long i = __umul64hi(mix64(key ^ 0x9E3779B97F4A7C15ULL), scap);
    for (unsigned long long n = 0; n < scap; ++n) {
        const unsigned long long cur = ld_relaxed(&S[i].tag);
        if (cur == tag) return &S[i];
        if (cur == 0) return nullptr;
        if (++i == scap) i = 0;
    }
    return nullptr;
}

__device__ __forceinline__ SideSlot *side_find_or_claim(const TableView &tv, SideSlot *S, unsigned long long scap,
                                                        int which, uint64_t key) {
    const unsigned long long tag = key + 1;
    unsigned long long i = __umul64hi(mix64(key ^ 0x9E3779B97F4A7C15ULL), scap);
    for (unsigned long long n = 0; n < scap; ++n) {
        unsigned long long cur = ld_relaxed(&S[i].tag);
        if (cur == 0) {
            cur = atomicCAS(&S[i].tag, 0ULL, tag);
            if (cur == 0) { atomicAdd(&tv.occ[which], 1ULL); return &S[i]; }
        }
        if (cur == tag) return &S[i];
        if (++i == scap) i = 0;
    }
    atomicOr(tv.err, 2);
    return nullptr;
}

__device__ __forceinline__ uint32_t clamp32(unsigned long long v) {
    return v > KQ_LARGEST ? KQ_LARGEST : (uint32_t)v;
}

// add (cov, lanes) to the remainder record of key (atomic; any thread, any time)
__device__ __forceinline__ void remainder_add(const TableView &tv, uint64_t key, unsigned long long cov,
                                              const uint32_t fw[4], const uint32_t bw[4]) {
    SideSlot *q = side_find_or_claim(tv, tv.R, tv.rcap, 1, key);
    if (!q) return;
    if (cov) atomicAdd(&q->cov, cov);
#pragma unroll
    for (int w = 0; w < 4; ++w) {
        if (fw[w]) atomicAdd(&q->fw[w], (unsigned long long)fw[w]);
        if (bw[w]) atomicAdd(&q->bw[w], (unsigned long long)bw[w]);
    }
}

// ---- main table -------------------------------------------------------------------------------------

// slot of key, or nullptr; *out receives a consistent snapshot of it
__device__ __forceinline__ Slot *find_slot(const TableView &tv, uint64_t key, Slot *out) {
    const unsigned long long tag = key + 1;
    const uint64_t h = home_hash(key);
    Slot *base = tv.T + (slice_of_key(key, tv.nslices) << tv.slice_bits);
    const unsigned long long mask = (1ull << tv.slice_bits) - 1;
    unsigned long long j = h & mask;
    for (unsigned long long n = 0; n <= mask; ++n) {
        const Slot s = ld_slot(base + j);
        if (s.tag == tag) { *out = s; return base + j; }
        if (s.tag == 0) return nullptr;
        j = (j + 1) & mask;
    }
    return nullptr;
}

// Count ONE instance (edge byte e) of key.  Returns 1 if the key is new, 0 if it existed, -1 if the
// key's slice is full (nothing was counted; the caller keeps the tuple for after a grow).
// Reference semantics: DBG::processBuffers src/graph-builder.cpp:160-206 in closed form (exact sums;
// the 8-bit / 32-bit split of the reference is decided when the database is written).
__device__ __forceinline__ int insert_instance(const TableView &tv, uint64_t key, uint32_t e) {
    const unsigned long long tag = key + 1, inc = inc_of_edge(e);
    const uint64_t h = home_hash(key);
    Slot *base = tv.T + (slice_of_key(key, tv.nslices) << tv.slice_bits);
    const unsigned long long mask = (1ull << tv.slice_bits) - 1;
    unsigned long long j = h & mask;
    // a run of 4096 occupied slots means the slice is (locally) full: give the tuple back instead of walking a
    // full slice slot by slot; the caller grows the table and counts it again
    const unsigned long long limit = mask < 4095 ? mask : 4095;
    for (unsigned long long n = 0; n <= limit; ++n) {
        Slot *p = base + j;
        Slot s = ld_slot(p);
        if (s.tag == 0) {                                           // claim and count in one 128-bit CAS
            const Slot prev = cas_slot(p, Slot{0, 0}, Slot{tag, inc});
            if (prev.tag == 0) return 1;
            s = prev;                                               // somebody else claimed it first
        }
        if (s.tag == tag) {
            unsigned long long cur = s.cnt;
            for (;;) {
                if (cnt_cov(cur) >= KQ_LANE_MAX) {
                    // slot is full: move 63 of the coverage and up to 63 of every lane to the remainder in
                    // the same CAS that counts this instance -- R is touched once per 63 instances of a
                    // high-copy k-mer, and the slot stays >= KQ_REM_FLOOR so readers know to look there
                    uint32_t fw[4], bw[4];
                    unsigned long long take = 63ull << 56;
#pragma unroll
                    for (int w = 0; w < 4; ++w) {
                        fw[w] = min(cnt_lane(cur, w), 63u); bw[w] = min(cnt_lane(cur, 4 + w), 63u);
                        take |= (unsigned long long)fw[w] << (7 * w) | (unsigned long long)bw[w] << (7 * (4 + w));
                    }
                    const unsigned long long old = atomicCAS(&p->cnt, cur, cur - take + inc);
                    if (old == cur) { remainder_add(tv, key, 63, fw, bw); return 0; }
                    cur = old;
                    continue;
                }
                // lanes <= cov < 127 for counted records, so cur + inc cannot carry between lanes
                const unsigned long long old = atomicCAS(&p->cnt, cur, cur + inc);
                if (old == cur) return 0;
                cur = old;
            }
        }
        j = (j + 1) & mask;
    }
    return -1;
}

// find the slot of key, claiming an empty one (cnt 0) if needed; *is_new reports a claim
__device__ __forceinline__ Slot *find_or_claim(const TableView &tv, uint64_t key, bool *is_new) {
    const unsigned long long tag = key + 1;
    const uint64_t h = home_hash(key);
    Slot *base = tv.T + (slice_of_key(key, tv.nslices) << tv.slice_bits);
    const unsigned long long mask = (1ull << tv.slice_bits) - 1;
    unsigned long long j = h & mask;
    *is_new = false;
    for (unsigned long long n = 0; n <= mask; ++n) {
        unsigned long long cur = ld_relaxed(&base[j].tag);
        if (cur == 0) {
            cur = atomicCAS(&base[j].tag, 0ULL, tag);
            if (cur == 0) { *is_new = true; return base + j; }
        }
        if (cur == tag) return base + j;
        j = (j + 1) & mask;
    }
    atomicOr(tv.err, 1);
    return nullptr;
}

// logical record of a non-placeholder slot snapshot: lanes + remainder
__device__ __forceinline__ void slot_logical(const TableView &tv, uint64_t key, unsigned long long cnt, Rec32 &r) {
    unsigned long long c = cnt_cov(cnt), f[4], b[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) { f[w] = cnt_lane(cnt, w); b[w] = cnt_lane(cnt, 4 + w); }
    if (cnt_saturated(cnt)) {
        const SideSlot *q = side_find(tv.R, tv.rcap, key);
        if (q) {
            c += q->cov;
#pragma unroll
            for (int w = 0; w < 4; ++w) { f[w] += q->fw[w]; b[w] += q->bw[w]; }
        }
    }
#pragma unroll
    for (int w = 0; w < 4; ++w) { r.fw[w] = clamp32(f[w]); r.bw[w] = clamp32(b[w]); }
    r.cov = clamp32(c);
}

// lookup idiom of src/kreeq.cpp:150-166: returns 1 found, 0 absent, -1 placeholder without maps32 entry
__device__ __forceinline__ int lookup(const TableView &tv, uint64_t key, Rec32 &r) {
    Slot s;
    if (!find_slot(tv, key, &s)) return 0;
    if (s.cnt & CNT_PLACEHOLDER) {
        const SideSlot *q = side_find(tv.M, tv.mcap, key);
        if (!q) return -1;
#pragma unroll
        for (int w = 0; w < 4; ++w) { r.fw[w] = clamp32(q->fw[w]); r.bw[w] = clamp32(q->bw[w]); }
        r.cov = clamp32(q->cov);
        return 1;
    }
    slot_logical(tv, key, s.cnt, r);
    return 1;
}

// add an arbitrary logical record to the (non-placeholder) slot p of key: lanes take what fits, the
// remainder table the rest.  Used by load / union where keys are unique within a launch.
__device__ __forceinline__ void slot_add_record(const TableView &tv, Slot *p, uint64_t key, const Rec32 &a) {
    unsigned long long cur = ld_relaxed(&p->cnt);
    for (;;) {
        unsigned long long nw = cur & CNT_PLACEHOLDER;
        uint32_t rf[4], rb[4]; unsigned long long rc;
        {
            const uint32_t have = cnt_cov(cur), take = min(a.cov, KQ_LANE_MAX - have);
            nw |= (unsigned long long)(have + take) << 56; rc = a.cov - take;
        }
        bool rem = rc != 0;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const uint32_t hf = cnt_lane(cur, w), tf = min(a.fw[w], KQ_LANE_MAX - hf);
            const uint32_t hb = cnt_lane(cur, 4 + w), tb = min(a.bw[w], KQ_LANE_MAX - hb);
            nw |= (unsigned long long)(hf + tf) << (7 * w); nw |= (unsigned long long)(hb + tb) << (7 * (4 + w));
            rf[w] = a.fw[w] - tf; rb[w] = a.bw[w] - tb; rem |= rf[w] || rb[w];
        }
        const unsigned long long old = atomicCAS(&p->cnt, cur, nw);
        if (old == cur) { if (rem) remainder_add(tv, key, rc, rf, rb); return; }
        cur = old;
    }
}

#endif  // __CUDACC__
}  // namespace kq

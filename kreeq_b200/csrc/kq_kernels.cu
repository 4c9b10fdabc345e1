// kq_kernels.cu -- hand-written sm_100a kernels of the kreeq k-mer hot path: count pipeline,
// validate, point lookups.  (Table maintenance kernels are in kq_table_kernels.cu.)
//
// Nothing here is a dense contraction; no tensor cores.  What bounds these kernels on B200 is the
// number of random table touches the memory system can serve (profiles/r01_gups_*.log): ~36 G loads/s
// and ~20 G atomics/s when the touched footprint is far larger than L2, but ~290 G loads/s and
// ~45 G (load + CAS)/s when it is <= 64 MB.  The count path therefore never touches the table at
// random in HBM: tuples are first scattered into slice order with coalesced 256-byte flushes
// (slice = 32 MiB of table), then inserted slice after slice while that slice sits in L2.
//
// Reference semantics each kernel follows (paths relative to the reference tree):
//   extraction (all count kernels) : DBG::hashSequences   src/graph-builder.cpp:62-113
//   insert                         : DBG::processBuffers  src/graph-builder.cpp:160-206 (closed form)
//   validate                       : DBG::evaluateSegment src/kreeq.cpp:143-219
#include "kq_kernels.h"

#include <algorithm>
#include <cstdlib>

namespace kq {

constexpr int TPB = 256;          // threads per CTA of the grid-stride kernels
constexpr int CTAS_PER_SM = 6;
constexpr int RUN = 8;            // consecutive k-mer start positions one thread extracts per round

// Tuning knobs: process-wide, read ONCE from the KQ_* environment variables (first use) and changeable through
// kq_set_tuning; no kernel launcher and no per-batch function reads the environment.
Tuning &tuning() {
    static Tuning t = [] {
        auto num = [](const char *name, long dflt) { const char *e = std::getenv(name); return e && *e ? std::atol(e) : dflt; };
        Tuning v;
        v.scatter_cap = (int)num("KQ_SCATTER_CAP", 0);
        v.rt_variant = (int)num("KQ_RT_VARIANT", 0);
        v.sc_variant = (int)num("KQ_SC_VARIANT", 0);
        { const char *e = std::getenv("KQ_SCATTER"); v.scatter_direct = e && e[0] == 'd'; }
        v.insert_ctas = (int)num("KQ_INSERT_CTAS", 8); if (v.insert_ctas < 1 || v.insert_ctas > 8) v.insert_ctas = 8;
        v.validate_ctas = (int)num("KQ_VALIDATE_CTAS", 8); if (v.validate_ctas < 1 || v.validate_ctas > 8) v.validate_ctas = 8;
        v.single_slice_bits = (int)num("KQ_SINGLE_SLICE_BITS", KQ_MAX_SLICE_BITS);
        if (v.single_slice_bits < 8 || v.single_slice_bits > 26) v.single_slice_bits = KQ_MAX_SLICE_BITS;
        v.piece_bits = (int)num("KQ_PIECE_BITS", KQ_PIECE_BITS); if (v.piece_bits < 6 || v.piece_bits > 13) v.piece_bits = KQ_PIECE_BITS;
        v.part = (unsigned long long)num("KQ_FINE_PART", 1 << 18); if (v.part < 64) v.part = 64;
        v.smem_insert = (int)num("KQ_SMEM_INSERT", 1);
        v.direct_pct = (int)num("KQ_DIRECT_PCT", 25);
        { const char *e = std::getenv("KQ_DB_HOST"); v.db_host = e && e[0] == '1'; }
        v.fs_variant = (int)num("KQ_FS_VARIANT", 0);
        v.pipeline = (int)num("KQ_PIPELINE", 3); if (v.pipeline < 2 || v.pipeline > 4) v.pipeline = 3;
        v.cursor_stride = (int)num("KQ_CURSOR_STRIDE", 8); if (v.cursor_stride < 1 || v.cursor_stride > 32) v.cursor_stride = 8;
        v.warp_insert = (int)num("KQ_WARP_INSERT", 1);
        v.route_smem = (int)num("KQ_ROUTE_SMEM", 0);
        v.cta_lock = (int)num("KQ_CTA_LOCK", 0);
        v.one_pass = (int)num("KQ_ONE_PASS", 0); if (v.one_pass < 0 || v.one_pass > 2) v.one_pass = 0;
        v.db_ordered = (int)num("KQ_DB_ORDERED", 1);
        v.coarse_bias = (int)num("KQ_COARSE_BIAS", 0); if (v.coarse_bias < 0 || v.coarse_bias > 4) v.coarse_bias = 0;
        v.s3_threads = (int)num("KQ_S3_THREADS", 256); v.s3_ctas = (int)num("KQ_S3_CTAS", 3); if (v.s3_ctas < 1 || v.s3_ctas > 8) v.s3_ctas = 3;
        return v;
    }();
    return t;
}

int sm_count() {
    static int n = 0;
    if (!n) {
        int dev = 0; cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

static inline unsigned grid_for(uint64_t n, int per_thread = 1) {
    const uint64_t full = (uint64_t)sm_count() * CTAS_PER_SM;
    const uint64_t need = (n + (uint64_t)TPB * per_thread - 1) / ((uint64_t)TPB * per_thread);
    return (unsigned)(need < full ? (need ? need : 1) : full);
}

// block-wide sum of per-thread counters (< 2^32 per thread), one global atomic per CTA
__device__ __forceinline__ void block_add(unsigned long long v, unsigned long long *dst, unsigned long long *sh) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);   // 64-bit: a thread may carry more than 2^32
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(sh, v);
    __syncthreads();
    if (threadIdx.x == 0 && *sh) atomicAdd(dst, *sh);
}

// ---- extraction ---------------------------------------------------------------------------------------

// 64 bits at bit offset o (< 128) of the 192-bit little-endian value {w0, w1, w2}
__device__ __forceinline__ uint64_t bits3(uint64_t w0, uint64_t w1, uint64_t w2, uint32_t o) {
    const uint64_t lo = o < 64 ? w0 : w1, hi = o < 64 ? w1 : w2;
    const uint32_t s = o & 63u;
    return s ? (lo >> s) | (hi << (64 - s)) : lo;
}

// k-mers starting at positions p0 .. p0+cnt-1 (cnt <= RUN) of a packed sequence: 5 word loads for the
// whole run instead of 6 per position.  Returns the mask of valid windows.
template <int R>
__device__ __forceinline__ uint32_t extract_run(const uint64_t *__restrict__ B, const uint64_t *__restrict__ M,
                                                uint64_t p0, uint32_t cnt, int k, uint64_t key[R], uint32_t e[R]) {
    static_assert(R <= 16, "three base words / two invalid words cover at most 16 positions + k + 1 bases");
    const uint64_t bw = p0 >> 5; const uint32_t bs = (uint32_t)(2 * p0) & 63u;
    const uint64_t b0 = __ldg(B + bw), b1 = __ldg(B + bw + 1), b2 = __ldg(B + bw + 2);
    const uint64_t iw = p0 >> 6; const uint32_t is = (uint32_t)p0 & 63u;
    const uint64_t i0 = __ldg(M + iw), i1 = __ldg(M + iw + 1);
    uint32_t prv = (p0 == 0 || invalid_at(M, p0 - 1)) ? 4u : base_at(B, p0 - 1);
    const uint64_t kmask = k == 32 ? ~0ULL : ((1ULL << (2 * k)) - 1), vmask = (1ULL << k) - 1;
    // forward and reverse-complement values of the first window (Kmap::hash, kmer.h:781-794), then rolled:
    // fw drops its lowest digit and gains the new base on top, rv gains the complement at the bottom
    const uint64_t win0 = bits3(b0, b1, b2, bs);
    uint64_t fw = win0 & kmask, rv = reverse_pairs(~win0) >> (64 - 2 * k);
    uint32_t mask = 0;
#pragma unroll
    for (int i = 0; i < R; ++i) {
        const uint32_t oi = is + i;                                  // <= 63 + R - 1; k + 1 more bits are needed
        const uint64_t inv = oi < 64 ? ((i0 >> oi) | (oi ? i1 << (64 - oi) : 0)) : (i1 >> (oi - 64));
        const uint32_t on = bs + 2 * i + 2 * k;                      // bit offset of the base after the window
        const uint64_t wn = on < 64 ? b0 : on < 128 ? b1 : b2;
        const uint32_t nb = (uint32_t)(wn >> (on & 63u)) & 3u;       // its code (0 if it is not ACGT)
        const uint32_t nxt = ((inv >> k) & 1) ? 4u : nb;
        if ((uint32_t)i < cnt && (inv & vmask) == 0) {
            Kmer km; km.isFw = fw < rv; km.key = km.isFw ? fw : rv; km.nxt = nxt; km.prv = prv;   // tie => not forward
            key[i] = km.key; e[i] = edge_byte(km); mask |= 1u << i;
        }
        prv = (inv & 1) ? 4u : (uint32_t)fw & 3u;
        fw = (fw >> 2) | ((uint64_t)nb << (2 * k - 2));
        rv = ((rv << 2) | (3u - nb)) & kmask;
    }
    return mask;
}

template <int TW> __device__ __forceinline__ void tuple_unpack(const uint64_t *t, uint64_t i, uint64_t &key, uint32_t &e) {
    if (TW == 1) { const uint64_t w = __ldcs(t + i); key = w >> 8; e = (uint32_t)w & 255u; }
    else { const ulonglong2 w = __ldcs(reinterpret_cast<const ulonglong2 *>(t) + i); key = w.x; e = (uint32_t)w.y & 255u; }
}
template <int TW> __device__ __forceinline__ void tuple_store(uint64_t *t, uint64_t i, uint64_t key, uint32_t e) {
    if (TW == 1) t[i] = (key << 8) | e;
    else reinterpret_cast<ulonglong2 *>(t)[i] = make_ulonglong2(key, e);
}

// One round of a CTA over its tile: every thread produces up to RUN (key, edge) pairs.
// Sequence source: tile = nthreads * RUN consecutive positions, thread t takes RUN consecutive ones.
// Tuple source: tile = nthreads * RUN tuples, element i of thread t is tile + i * nthreads + t (coalesced).
template <int TW, int R = RUN>
__device__ __forceinline__ uint32_t produce(const Source &src, const TableView &tv, uint64_t tile, uint64_t total,
                                            uint64_t key[R], uint32_t e[R]) {
    if (src.tuples == nullptr) {
        const uint64_t first = tile + (uint64_t)threadIdx.x * R;
        if (first >= total) return 0;
        const uint32_t cnt = (uint32_t)min((uint64_t)R, total - first);
        uint32_t m = extract_run<R>(src.bases, src.invalid, src.pos0 + first, cnt, tv.k, key, e);
        if (tv.shard_rank >= 0) {
#pragma unroll
            for (int i = 0; i < R; ++i) if ((m >> i & 1u) && !owns(tv, key[i])) m &= ~(1u << i);
        }
        return m;
    }
    uint32_t m = 0;
#pragma unroll
    for (int i = 0; i < R; ++i) {
        const uint64_t j = tile + (uint64_t)i * blockDim.x + threadIdx.x;
        if (j < total) { tuple_unpack<TW>(src.tuples, j, key[i], e[i]); m |= 1u << i; }
    }
    return m;
}

__device__ __forceinline__ uint64_t source_total(const Source &src) {
    return src.tuples ? (uint64_t)*src.ntuples : src.npos;
}

// ---- count: fused extract + insert (single-slice tables, small inputs) ------------------------------

template <int TW>
__global__ void __launch_bounds__(TPB, 4)
count_direct_kernel(TableView tv, Source src, unsigned long long *instances,
                    uint64_t *ovf, unsigned long long ovf_cap, unsigned long long *ovf_n) {
    __shared__ unsigned long long sh[2];
    if (threadIdx.x < 2) sh[threadIdx.x] = 0;
    __syncthreads();
    unsigned long long inst = 0, fresh = 0;
    const uint64_t total = source_total(src), per = (uint64_t)TPB * RUN;
    for (uint64_t tile = (uint64_t)blockIdx.x * per; tile < total; tile += (uint64_t)gridDim.x * per) {
        uint64_t key[RUN]; uint32_t e[RUN];
        const uint32_t m = produce<TW>(src, tv, tile, total, key, e);
#pragma unroll
        for (int i = 0; i < RUN; ++i) {
            if (!(m >> i & 1u)) continue;
            const int r = insert_instance(tv, key[i], e[i]);
            ++inst;
            if (r < 0) {                                              // table full: keep the tuple for after a grow
                const unsigned long long j = atomicAdd(ovf_n, 1ULL);
                if (j < ovf_cap) tuple_store<TW>(ovf, j, key[i], e[i]);
            } else fresh += r;
        }
    }
    block_add(inst, instances, &sh[0]);
    block_add(fresh, &tv.occ[0], &sh[1]);
}

// ---- count: one-pass scatter into per-slice regions (big tables) -----------------------------------------------

// Slice s of the table owns the tuples [s * cap, s * cap + min(cursor[s], cap)) of `out`: a k-mer instance takes its
// place with ONE returning L2 atomic on its slice's cursor and one 8 / 16-byte store.  No histogram pass and no
// shared-memory bins: the cursors (4 bytes x stride per slice, a few MB) and the frontier line of every region
// (nslices x 128 B, tens of MB) stay in L2, which merges the stores of a line before it goes to HBM.
// cap is 5/4 of the mean plus slack; the instances of a slice that is over-subscribed (high-copy k-mers) are counted
// at once with the L2 CAS path, so nothing is ever dropped and no capacity has to be guessed right.
template <int TW>
__global__ void __launch_bounds__(TPB, 4)
scatter_fine_kernel(TableView tv, Source src, uint64_t *__restrict__ out, unsigned cap, unsigned *cursor,
                    int stride, unsigned long long *instances, uint64_t *ovf, unsigned long long ovf_cap, unsigned long long *ovf_n) {
    __shared__ unsigned long long sh[2];
    if (threadIdx.x < 2) sh[threadIdx.x] = 0;
    __syncthreads();
    unsigned long long inst = 0, fresh = 0;
    const uint64_t total = source_total(src), per = (uint64_t)TPB * RUN;
    for (uint64_t tile = (uint64_t)blockIdx.x * per; tile < total; tile += (uint64_t)gridDim.x * per) {
        uint64_t key[RUN]; uint32_t e[RUN]; unsigned at[RUN], sl[RUN];
        uint32_t m = produce<TW>(src, tv, tile, total, key, e);
        // all RUN atomics of the thread are issued back to back (predicated, one result register each): the L2 round
        // trips overlap instead of adding up
#pragma unroll
        for (int i = 0; i < RUN; ++i) {
            if ((m >> i & 1u) && e[i] == 255u) m &= ~(1u << i);             // padding of the routed layout
            const unsigned live = m >> i & 1u;
            sl[i] = live ? (unsigned)slice_of_key(key[i], tv.nslices) : 0u;
            asm volatile("{\n .reg .pred q;\n setp.ne.u32 q, %2, 0;\n mov.u32 %0, 0;\n"
                         " @q atom.relaxed.gpu.global.add.u32 %0, [%1], 1;\n}"
                         : "=r"(at[i]) : "l"(cursor + (size_t)sl[i] * stride), "r"(live) : "memory");
        }
#pragma unroll
        for (int i = 0; i < RUN; ++i) {
            if (!(m >> i & 1u)) continue;
            ++inst;
            if (at[i] < cap) { tuple_store<TW>(out, (uint64_t)sl[i] * cap + at[i], key[i], e[i]); continue; }
            const int r = insert_instance(tv, key[i], e[i]);                // over-subscribed slice: count it now
            if (r < 0) {
                const unsigned long long j = atomicAdd(ovf_n, 1ULL);
                if (j < ovf_cap) tuple_store<TW>(ovf, j, key[i], e[i]);
            } else fresh += r;
        }
    }
    if (instances) block_add(inst, instances, &sh[0]);
    block_add(fresh, &tv.occ[0], &sh[1]);
}

// ---- count, step 1: tuples per slice -----------------------------------------------------------------

// Bin of a key: its slice in the local pipeline; in the routed (multi-GPU) pipeline the owner rank in the
// high bits and 2^nb_bits equal ranges of the slice hash below.  Both are monotone in the slice hash, so
// tuples in bin order are in slice order on the table that receives them, whatever its geometry.
template <bool RT>
__device__ __forceinline__ unsigned bin_of_key(const TableView &tv, const RouteMap &rt, uint64_t key) {
    if (!RT) return (unsigned)(slice_of_key(key, tv.nslices) >> tv.cshift);
    const unsigned owner = (unsigned)(key & 127u) >> rt.shift;
    const uint64_t h = (uint32_t)((key * 0x9E3779B97F4A7C15ULL) >> 32);
    return (owner << rt.nb_bits) | (unsigned)(h >> (32 - rt.nb_bits));
}
template <bool RT>
__device__ __forceinline__ unsigned bin_count(const TableView &tv, const RouteMap &rt) {
    return RT ? (unsigned)rt.nranks << rt.nb_bits : (unsigned)(((tv.nslices - 1) >> tv.cshift) + 1);
}

template <int TW, bool RT>
__global__ void __launch_bounds__(TPB, 4)
slice_hist_kernel(TableView tv, Source src, unsigned long long *hist, RouteMap rt) {
    extern __shared__ unsigned int h[];
    const unsigned P = bin_count<RT>(tv, rt);
    for (unsigned i = threadIdx.x; i < P; i += TPB) h[i] = 0;
    __syncthreads();
    const uint64_t total = source_total(src), per = (uint64_t)TPB * RUN;
    for (uint64_t tile = (uint64_t)blockIdx.x * per; tile < total; tile += (uint64_t)gridDim.x * per) {
        uint64_t key[RUN]; uint32_t e[RUN];
        const uint32_t m = produce<TW>(src, tv, tile, total, key, e);
#pragma unroll
        for (int i = 0; i < RUN; ++i)
            if (m >> i & 1u) atomicAdd(&h[bin_of_key<RT>(tv, rt, key[i])], 1u);
    }
    __syncthreads();
    for (unsigned i = threadIdx.x; i < P; i += TPB)
        if (h[i]) atomicAdd(&hist[i], (unsigned long long)h[i]);
}

// ---- count, step 2: exclusive scan of the slice histogram (one CTA) ------------------------------------

__global__ void __launch_bounds__(1024, 1)
slice_offsets_kernel(unsigned long long nslices, const unsigned long long *hist, unsigned long long *cursor,
                     unsigned long long *start, unsigned long long *total, unsigned long long *instances) {
    __shared__ unsigned long long part[1024];
    const unsigned long long per = (nslices + 1023) / 1024;
    const unsigned long long b = threadIdx.x * per, e = min(nslices, b + per);
    unsigned long long s = 0;
    for (unsigned long long i = b; i < e; ++i) s += hist[i];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long run = 0;
        for (int i = 0; i < 1024; ++i) { const unsigned long long v = part[i]; part[i] = run; run += v; }
        *total = run; if (instances) *instances += run;
    }
    __syncthreads();
    s = part[threadIdx.x];
    for (unsigned long long i = b; i < e; ++i) { cursor[i] = s; if (start) start[i] = s; s += hist[i]; }
}

// Routed pipeline: H[s][b] = tuples rank s sends into global bin b = (owner d, j).  Destination d lays its
// receive slab out bin by bin and, inside a bin, sender by sender; every (sender, bin) region is padded to
// a multiple of `cap` tuples so that all full-bin flushes are aligned cap * 8 * TW byte stores (one NVLink
// packet, whole sectors).  CTA d computes, for THIS rank (me): front[b] = start of its region in bin b of d,
// back[b] = its end (partial flushes are taken from the back so the front stays aligned), the extent of
// d's slab (recv[d]) and the number of real tuples in it (recv[nranks + d]); it also writes the sentinel
// tuples (all ones: no k-mer instance has edge byte 0xFF) into the padding behind its own regions.
__global__ void __launch_bounds__(1024, 1)
route_offsets_kernel(RouteMap rt, int me, int cap, int tw, unsigned long long slab_cap, const unsigned long long *H,
                     unsigned long long *front, unsigned long long *back, unsigned long long *recv) {
    __shared__ unsigned long long part[1024], real[1024];
    __shared__ bool fits;
    const int nranks = rt.nranks;
    const unsigned nb = 1u << rt.nb_bits, d = blockIdx.x, P = (unsigned)nranks * nb;
    const unsigned per = (nb + 1023) / 1024;
    const unsigned b = threadIdx.x * per, e = min(nb, b + per);
    const unsigned long long c1 = (unsigned long long)cap - 1;
    unsigned long long s = 0, rl = 0;
    for (unsigned j = b; j < e; ++j)
        for (int r = 0; r < nranks; ++r) { const unsigned long long v = H[(size_t)r * P + d * nb + j]; s += (v + c1) / cap * cap; rl += v; }
    part[threadIdx.x] = s; real[threadIdx.x] = rl;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long run = 0, rr = 0;
        for (int i = 0; i < 1024; ++i) { const unsigned long long v = part[i]; part[i] = run; run += v; rr += real[i]; }
        recv[d] = run; recv[nranks + d] = rr;
        fits = run <= slab_cap;                                      // else the host refuses the round: write nothing
    }
    __syncthreads();
    s = part[threadIdx.x];
    for (unsigned j = b; j < e; ++j)
        for (int r = 0; r < nranks; ++r) {
            const unsigned long long v = H[(size_t)r * P + d * nb + j], L = (v + c1) / cap * cap;
            if (r == me) {
                front[d * nb + j] = s; back[d * nb + j] = s + v;
                uint64_t *pad = rt.out[d] + (s + v) * tw;
                if (fits) for (unsigned long long w = 0; w < (L - v) * tw; ++w) pad[w] = ~0ULL;
            }
            s += L;
        }
}

// ---- count, step 3: scatter into slice order -------------------------------------------------------------

// One-pass form (no histogram pass): coarse bin b owns the fixed region [b * rcap, end[b]) of the output.  A chunk of n
// tuples reserved at g that does not fit (a bin with more than its share + 3 %: runs of one key) goes to the overflow
// list instead (position | KQ_REGION_OVF; the list is as large as the batch); what is left of the region in front of
// it is filled with sentinel tuples (edge byte 0xFF), so [start, min(cursor, end)) is all tuples or sentinels.
constexpr unsigned long long KQ_REGION_OVF = 1ULL << 63;
template <int TW>
__device__ __noinline__ unsigned long long region_miss(unsigned long long *stat, uint64_t *out, unsigned long long end, unsigned long long g, unsigned n) {
    if (g < end) {
        for (unsigned long long w = g * TW; w < end * TW; ++w) out[w] = ~0ULL;
        atomicAdd(&stat[1], end - g);                                  // sentinels inside the region
    }
    return atomicAdd(&stat[0], (unsigned long long)n) | KQ_REGION_OVF;
}
template <int TW>
__device__ __forceinline__ unsigned long long region_place(const Regions &rg, uint64_t *out, unsigned b, unsigned long long g, unsigned n) {
    const unsigned long long end = (unsigned long long)(b + 1) * rg.rcap;   // (no load: the reservation's round trip is the critical path)
    return g + n <= end ? g : region_miss<TW>(rg.stat, out, end, g, n);
}

// start / end / cursor of the fixed regions; hist (tuples per bin) and the totals once the scatter is done
__global__ void region_init_kernel(unsigned nbins, unsigned long long rcap, unsigned long long *cursor, unsigned long long *start,
                                   unsigned long long *end, unsigned long long *stat) {
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b == 0) { stat[0] = 0; stat[1] = 0; stat[2] = 0; }
    if (b < nbins) { cursor[b] = start[b] = (unsigned long long)b * rcap; end[b] = (unsigned long long)(b + 1) * rcap; }
}
__global__ void __launch_bounds__(1024, 1)
region_counts_kernel(unsigned nbins, unsigned long long *cursor, const unsigned long long *start, const unsigned long long *end,
                     unsigned long long *hist, unsigned long long *total, const unsigned long long *stat, unsigned long long *instances) {
    // hist[b] = extent of bin b (tuples + sentinels); cursor[b] becomes the exclusive scan of the extents: where the bin's
    // tuples start in the level-2 output (the regions themselves are far apart)
    __shared__ unsigned long long part[1024];
    const unsigned per = (nbins + 1023) / 1024, b0 = threadIdx.x * per, b1 = min(nbins, b0 + per);
    unsigned long long mine = 0;
    for (unsigned b = b0; b < b1; ++b) { const unsigned long long n = min(cursor[b], end[b]) - start[b]; hist[b] = n; mine += n; }
    part[threadIdx.x] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long run = 0;
        for (int i = 0; i < 1024; ++i) { const unsigned long long v = part[i]; part[i] = run; run += v; }
        *total = run;
        if (instances) *instances += run - stat[1] + stat[0];
    }
    __syncthreads();
    unsigned long long s = part[threadIdx.x];
    for (unsigned b = b0; b < b1; ++b) { cursor[b] = s; s += hist[b]; }
}

// ---- count, step 3 (two-pass form) ---------------------------------------------------------------------------------

template <int TW, int R, int NT, bool RT, bool OP = false>
__global__ void __launch_bounds__(NT, 1)
slice_scatter_kernel(TableView tv, Source src, uint64_t *__restrict__ out, unsigned long long *cursor, int cap, RouteMap rt,
                     unsigned long long *back, Regions rg) {
    extern __shared__ uint64_t sm[];
    const unsigned P = bin_count<RT>(tv, rt);
    uint64_t *bins = sm;
    unsigned int *fill = reinterpret_cast<unsigned int *>(bins + (size_t)P * cap * TW);
    for (unsigned i = threadIdx.x; i < P; i += NT) fill[i] = 0;
    __syncthreads();
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = NT / 32;
    const unsigned lpb = min(32u, (unsigned)cap * TW);                 // lanes that copy one bin (cap is 8, 16 or 32)
    const unsigned sub = lane / lpb;
    const uint64_t total = source_total(src), per = (uint64_t)NT * R;
    for (uint64_t tile = (uint64_t)blockIdx.x * per; tile < total; tile += (uint64_t)gridDim.x * per) {
        uint64_t key[R]; uint32_t e[R]; unsigned bin[R], slot[R];
        uint32_t m = produce<TW, R>(src, tv, tile, total, key, e);
        bool filled = false;
#pragma unroll
        for (int i = 0; i < R; ++i) {
            if (!(m >> i & 1u)) continue;
            bin[i] = bin_of_key<RT>(tv, rt, key[i]);
            slot[i] = atomicAdd(&fill[bin[i]], 1u);
            filled |= slot[i] >= (unsigned)cap - 1;
        }
        for (;;) {
#pragma unroll
            for (int i = 0; i < R; ++i) {
                if (!(m >> i & 1u) || slot[i] >= (unsigned)cap) continue;
                tuple_store<TW>(bins, (size_t)bin[i] * cap + slot[i], key[i], e[i]);
                m &= ~(1u << i);
            }
            if (!__syncthreads_or(filled)) break;                    // no bin reached cap in this round
            // lane l of a warp looks at bin b0 + l: all reservations of a warp's 32 bins are issued together
            // (one atomic round trip instead of one per bin), then the warp copies the full bins one by one
            for (unsigned b0 = warp * 32; b0 < P; b0 += nwarps * 32) {
                const unsigned b = b0 + lane;
                const unsigned f = b < P ? fill[b] : 0;
                const bool full = f >= (unsigned)cap;
                unsigned long long g = 0;
                if (full) { g = atomicAdd(&cursor[b], (unsigned long long)cap); fill[b] = f - cap; }
                unsigned todo = __ballot_sync(0xffffffffu, full);
                if (OP) {                                            // (almost never: a chunk that does not fit its bin's region)
                    const unsigned long long end = (unsigned long long)(b + 1) * rg.rcap;
                    const bool miss = full && g + (unsigned)cap > end;
                    if (__any_sync(0xffffffffu, miss) && miss) g = region_miss<TW>(rg.stat, out, end, g, (unsigned)cap);
                }
                // a bin is copied by min(32, words) lanes, so a warp moves 32 / that many full bins at a time
                while (todo) {
                    const unsigned l = lpb == 32 ? (unsigned)__ffs(todo) - 1 : __fns(todo, 0, (int)sub + 1);   // the sub-th bin still to do
#pragma unroll 1
                    for (unsigned i = 0; i < 32 / lpb && todo; ++i) todo &= todo - 1;
                    const unsigned long long gg = __shfl_sync(0xffffffffu, g, l & 31u);
                    if (l < 32) {
                        const uint64_t *srcw = bins + (size_t)(b0 + l) * cap * TW;
                        uint64_t *dst = (RT ? rt.out[(b0 + l) >> rt.nb_bits] : (OP && (gg & KQ_REGION_OVF)) ? rg.ovf : out) + (OP ? gg & ~KQ_REGION_OVF : gg) * TW;
                        for (unsigned w = lane & (lpb - 1); w < (unsigned)cap * TW; w += lpb) __stcs(dst + w, srcw[w]);
                    }
                }
            }
            __syncthreads();
            filled = false;
#pragma unroll
            for (int i = 0; i < R; ++i) {
                if (!(m >> i & 1u)) continue;
                slot[i] -= cap;                                      // the first cap entries of the bin are gone
                filled = true;                                       // something is still waiting: go round again
            }
            if (!__syncthreads_or(filled)) break;
        }
    }
    __syncthreads();
    for (unsigned b0 = warp * 32; b0 < P; b0 += nwarps * 32) {         // what is left in the bins
        const unsigned b = b0 + lane;
        const unsigned f = b < P ? fill[b] : 0;
        unsigned long long g = 0;
        // routed: what is left goes to the END of this rank's region of the bin, the aligned front is for full bins
        if (f) {
            g = RT ? atomicAdd(&back[b], 0ULL - f) - f : atomicAdd(&cursor[b], (unsigned long long)f);
            if (OP) g = region_place<TW>(rg, out, b, g, f);
        }
        unsigned todo = __ballot_sync(0xffffffffu, f != 0);
        while (todo) {
            const int l = __ffs(todo) - 1; todo &= todo - 1;
            const unsigned long long gg = __shfl_sync(0xffffffffu, g, l);
            const unsigned ff = __shfl_sync(0xffffffffu, f, l);
            const uint64_t *srcw = bins + (size_t)(b0 + l) * cap * TW;
            uint64_t *dst = (RT ? rt.out[(b0 + l) >> rt.nb_bits] : (OP && (gg & KQ_REGION_OVF)) ? rg.ovf : out) + (OP ? gg & ~KQ_REGION_OVF : gg) * TW;
            for (unsigned w = lane; w < ff * TW; w += 32) __stcs(dst + w, srcw[w]);
        }
    }
}

// ---- count, step 3 (ring form): scatter into coarse order, single pass ------------------------------------------

// Same job as slice_scatter_kernel<.., false> with the ring bins of the level-2 scatter (kq_count3.cu): a tuple is
// appended with one shared-memory atomic and one store, a bin that holds CAP tuples leaves as one coalesced store at
// a position reserved with one global atomic per chunk; the ring (3/2 CAP) takes what arrives while a chunk waits,
// so a tile needs two barriers and (practically) never a second pass over its tuples.
template <int TW, int R, int NT, int CAP>
__global__ void __launch_bounds__(NT, 1)
coarse_scatter_kernel(TableView tv, Source src, uint64_t *__restrict__ out, unsigned long long *cursor) {
    constexpr unsigned RING = CAP + CAP / 2;
    extern __shared__ uint64_t sm[];
    const unsigned P = bin_count<false>(tv, RouteMap{});
    uint64_t *bins = sm;
    unsigned int *fill = reinterpret_cast<unsigned int *>(bins + (size_t)P * RING * TW);
    unsigned int *done = fill + P;
    for (unsigned i = threadIdx.x; i < P; i += NT) { fill[i] = 0; done[i] = 0; }
    __syncthreads();
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = NT / 32;
    const uint64_t total = source_total(src), per = (uint64_t)NT * R;
    for (uint64_t tile = (uint64_t)blockIdx.x * per; tile < total; tile += (uint64_t)gridDim.x * per) {
        uint64_t key[R]; uint32_t e[R]; unsigned bin[R], pos[R];
        uint32_t m = produce<TW, R>(src, tv, tile, total, key, e);
#pragma unroll
        for (int i = 0; i < R; ++i) {
            if (!(m >> i & 1u)) continue;
            if (e[i] == 255u) { m &= ~(1u << i); continue; }               // padding tuples of a routed source
            bin[i] = (unsigned)(slice_of_key(key[i], tv.nslices) >> tv.cshift);
            pos[i] = atomicAdd(&fill[bin[i]], 1u);
        }
        for (;;) {
            bool full = false;
#pragma unroll
            for (int i = 0; i < R; ++i) {
                if (!(m >> i & 1u)) continue;
                const unsigned d = pos[i] - *reinterpret_cast<volatile unsigned int *>(&done[bin[i]]);
                if (d < RING) { tuple_store<TW>(bins, bin[i] * RING + pos[i] % RING, key[i], e[i]); m &= ~(1u << i); }
                full |= d + 1 >= (unsigned)CAP;
            }
            if (!__syncthreads_or(full)) break;
            for (unsigned b0 = warp * 32; b0 < P; b0 += nwarps * 32) {
                const unsigned b = b0 + lane;
                unsigned d0 = 0, chunks = 0;
                if (b < P) { d0 = done[b]; chunks = min(fill[b] - d0, RING) / CAP; }
                unsigned long long g = 0;
                if (chunks) { g = atomicAdd(&cursor[b], (unsigned long long)chunks * CAP); done[b] = d0 + chunks * CAP; }
                unsigned todo = __ballot_sync(0xffffffffu, chunks != 0);
                while (todo) {
                    const int l = __ffs(todo) - 1; todo &= todo - 1;
                    const unsigned long long gg = __shfl_sync(0xffffffffu, g, l);
                    const unsigned dd = __shfl_sync(0xffffffffu, d0, l);
                    const unsigned nn = __shfl_sync(0xffffffffu, chunks, l) * CAP;
                    const uint64_t *srcw = bins + (size_t)(b0 + l) * RING * TW;
                    uint64_t *dst = out + gg * TW;
                    for (unsigned w = lane; w < nn * TW; w += 32)
                        __stcs(dst + w, srcw[((dd + w / TW) % RING) * TW + w % TW]);
                }
            }
            if (!__syncthreads_or(m != 0)) break;                        // nobody had to wait for a flush
        }
    }
    __syncthreads();
    for (unsigned b0 = warp * 32; b0 < P; b0 += nwarps * 32) {             // what is left in the bins
        const unsigned b = b0 + lane;
        unsigned d0 = 0, f = 0;
        if (b < P) { d0 = done[b]; f = fill[b] - d0; }
        unsigned long long g = 0;
        if (f) g = atomicAdd(&cursor[b], (unsigned long long)f);
        unsigned todo = __ballot_sync(0xffffffffu, f != 0);
        while (todo) {
            const int l = __ffs(todo) - 1; todo &= todo - 1;
            const unsigned long long gg = __shfl_sync(0xffffffffu, g, l);
            const unsigned dd = __shfl_sync(0xffffffffu, d0, l);
            const unsigned nn = __shfl_sync(0xffffffffu, f, l);
            const uint64_t *srcw = bins + (size_t)(b0 + l) * RING * TW;
            uint64_t *dst = out + gg * TW;
            for (unsigned w = lane; w < nn * TW; w += 32)
                __stcs(dst + w, srcw[((dd + w / TW) % RING) * TW + w % TW]);
        }
    }
}

// Variant without shared-memory bins: one global atomic on the slice cursor per tuple and one 8/16-byte
// store; L2 is the write combiner (the ~600 cursor lines and the ~600 frontier lines of the output stay
// resident, every output line is complete before it is evicted).
template <int TW>
__global__ void __launch_bounds__(TPB, 4)
slice_scatter_direct_kernel(TableView tv, Source src, uint64_t *__restrict__ out, unsigned long long *cursor) {
    const uint64_t total = source_total(src), per = (uint64_t)TPB * RUN;
    for (uint64_t tile = (uint64_t)blockIdx.x * per; tile < total; tile += (uint64_t)gridDim.x * per) {
        uint64_t key[RUN]; uint32_t e[RUN]; unsigned long long at[RUN];
        const uint32_t m = produce<TW>(src, tv, tile, total, key, e);
#pragma unroll
        for (int i = 0; i < RUN; ++i)
            if (m >> i & 1u) at[i] = atomicAdd(&cursor[slice_of_key(key[i], tv.nslices) >> tv.cshift], 1ULL);
#pragma unroll
        for (int i = 0; i < RUN; ++i)
            if (m >> i & 1u) tuple_store<TW>(out, at[i], key[i], e[i]);
    }
}

// ---- count, step 4: insert slice-ordered tuples -----------------------------------------------------------

// Consecutive CTAs work on consecutive tuples, i.e. on the same one or two slices at a time, so the
// slot loads and CASes below hit L2.  Tuples are streamed with evict-first loads to keep L2 for the table.
constexpr int INS_TILE = TPB * 8;   // tuples per ticket

// CTAs take tiles of INS_TILE consecutive tuples from a global ticket counter, so at any moment all
// resident CTAs work within ~1.8 M consecutive tuples (less than half a slice) however their speeds
// differ; with a static grid-stride assignment CTAs on the far L2 die fall behind and the set of
// slices in use outgrows L2.
template <int TW, int CPS>
__global__ void __launch_bounds__(TPB, CPS)
insert_tuples_kernel(TableView tv, const uint64_t *__restrict__ tuples, const unsigned long long *n_ptr,
                     uint64_t *ovf, unsigned long long ovf_cap, unsigned long long *ovf_n, unsigned long long *ticket, unsigned long long n_cap) {
    __shared__ unsigned long long sh[2];
    if (threadIdx.x == 0) sh[0] = 0;
    unsigned long long fresh = 0;
    const uint64_t n = min(*n_ptr, n_cap);                            // (a list that overflowed was flagged by its writer)
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) sh[1] = atomicAdd(ticket, 1ULL);
        __syncthreads();
        const uint64_t tile = sh[1] * INS_TILE;
        if (tile >= n) break;
#pragma unroll 2
        for (int r = 0; r < INS_TILE / TPB; ++r) {
            const uint64_t i = tile + (uint64_t)r * TPB + threadIdx.x;
            if (i >= n) break;
            uint64_t key; uint32_t e;
            tuple_unpack<TW>(tuples, i, key, e);
            if (e == 255u) continue;                                  // padding of the routed layout (no instance has all 8 edges)
            const int st = insert_instance(tv, key, e);
            if (st < 0) {                                             // slice full: keep the tuple for after a grow
                const unsigned long long j = atomicAdd(ovf_n, 1ULL);
                if (j < ovf_cap) tuple_store<TW>(ovf, j, key, e); else atomicOr(tv.err, 1);
            } else fresh += st;
        }
    }
    block_add(fresh, &tv.occ[0], &sh[0]);
}

// ---- count: extract with routing (multi-GPU) ------------------------------------------------------------

template <int TW>
__global__ void __launch_bounds__(TPB, 4)
extract_route_kernel(TableView tv, Source src, RouteArgs ra) {
    const uint64_t total = src.npos, per = (uint64_t)TPB * RUN;
    const unsigned lane = threadIdx.x & 31;
    tv.shard_rank = -1;                                               // every window is routed, none dropped
    for (uint64_t tile = (uint64_t)blockIdx.x * per; tile < total; tile += (uint64_t)gridDim.x * per) {
        uint64_t key[RUN]; uint32_t e[RUN];
        const uint32_t m = produce<TW>(src, tv, tile, total, key, e);
#pragma unroll
        for (int i = 0; i < RUN; ++i) {
            const int dest = (m >> i & 1u) ? (int)((key[i] & 127) >> ra.shift) : -1;
            // warp-aggregated append: one atomic per (warp, destination)
            const unsigned peers = __match_any_sync(0xffffffffu, dest);
            if (dest < 0) continue;
            const unsigned leader = __ffs(peers) - 1;
            unsigned long long base = 0;
            if (lane == leader) base = atomicAdd(&ra.counts[dest], (unsigned long long)__popc(peers));
            base = __shfl_sync(peers, base, leader);
            const unsigned long long slot = base + __popc(peers & ((1u << lane) - 1));
            if (slot < ra.cap_per_rank) tuple_store<TW>(ra.tuples, (unsigned long long)dest * ra.cap_per_rank + slot, key[i], e[i]);
            else atomicOr(tv.err, 8);
        }
    }
}

// ---- validate ------------------------------------------------------------------------------------

// DBG::evaluateSegment src/kreeq.cpp:143-219 for positions [c0, c0+n) of one segment.
template <int CPS>
__global__ void __launch_bounds__(TPB, CPS)
validate_kernel(TableView tv, const uint64_t *__restrict__ B, const uint64_t *__restrict__ M,
                uint64_t base0, uint64_t c0, uint64_t n, uint64_t kcount, uint32_t cutoff,
                Base *__restrict__ out, unsigned long long *acc3, const TableView *__restrict__ peers, int npeers) {
    __shared__ unsigned long long sh[3];
    // Multi-GPU: peers[r] describes the shard of rank r (its T / R / M mapped into this process); every
    // position is probed in the table of its key's owner, over NVLink when that is another GPU.
    __shared__ TableView pv[KQ_MAX_RANKS];
    if (threadIdx.x < 3) sh[threadIdx.x] = 0;
    if (peers) {
        const unsigned words = (unsigned)(npeers * sizeof(TableView) / 4);
        for (unsigned i = threadIdx.x; i < words; i += TPB)
            reinterpret_cast<uint32_t *>(pv)[i] = reinterpret_cast<const uint32_t *>(peers)[i];
    }
    __syncthreads();
    unsigned long long missing = 0, counted = 0, edge_missing = 0;
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        const uint64_t c = c0 + i;
        uint4 o = make_uint4(0, 0, 0, 0);                       // fw, bw, cov, isFw
        if (c < kcount) {
            const Kmer km = kmer_at(B, M, c - base0, tv.k);
            // a window holding a non-ACGT byte can never be found; it still counts (kreeq.cpp:216),
            // on shard 0 only so that the per-shard counters add up
            if (peers || (km.valid ? owns(tv, km.key) : tv.shard_rank <= 0)) {
                Rec32 kh; int found = 0;
                if (km.valid) {
                    found = lookup(peers ? pv[(km.key & 127) >> tv.shard_shift] : tv, km.key, kh);
                    if (found < 0) { atomicOr(tv.err, 4); found = 0; }
                }
                if (found) { o.z = kh.cov; o.w = km.isFw ? 1u : 0u; }                // :168-169
                if (o.z == 0 || o.z < cutoff) {                                      // :172-175
                    ++missing;
                } else {
                    bool no_left = false, no_right = false;
                    const bool has_next = c < kcount - 1, has_prev = c > 0;
                    if (km.isFw) {                                                   // :178-192
                        if (has_next) { const uint32_t v = km.nxt <= 3 ? kh.fw[km.nxt] : 0; if (v) o.x = v; else no_right = true; }
                        if (has_prev) { const uint32_t v = km.prv <= 3 ? kh.bw[km.prv] : 0; if (v) o.y = v; else no_left = true; }
                    } else {                                                         // :194-210
                        if (has_prev) { const uint32_t v = km.prv <= 3 ? kh.fw[3 - km.prv] : 0; if (v) o.x = v; else no_left = true; }
                        if (has_next) { const uint32_t v = km.nxt <= 3 ? kh.bw[3 - km.nxt] : 0; if (v) o.y = v; else no_right = true; }
                    }
                    if (no_left && no_right) ++edge_missing;                         // :211-214
                }
                ++counted;
            }
        }
        __stcs(reinterpret_cast<uint4 *>(out) + i, o);          // one coalesced 16-byte store per base
    }
    block_add(missing, &acc3[0], &sh[0]);
    block_add(counted, &acc3[1], &sh[1]);
    block_add(edge_missing, &acc3[2], &sh[2]);
}

// 16-byte DBGbase rows -> {cov, fw, bw, flags} bytes; rows with a value above 254 also go to the wide list with their position
__global__ void __launch_bounds__(TPB, 8)
compact_rows_kernel(const Base *__restrict__ rows, uint64_t n, uint64_t pos0, uint32_t *__restrict__ out4, WideRow *wide,
                    unsigned long long wide_cap, unsigned long long *wide_n) {
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        const uint4 r = __ldcs(reinterpret_cast<const uint4 *>(rows) + i);          // fw, bw, cov, isFw
        const bool big = r.x > 254u || r.y > 254u || r.z > 254u;
        out4[i] = min(r.z, 255u) | min(r.x, 255u) << 8 | min(r.y, 255u) << 16 | ((r.w & 1u) | (big ? 2u : 0u)) << 24;
        if (big) {
            const unsigned long long j = atomicAdd(wide_n, 1ULL);
            if (j < wide_cap) { wide[j].pos = pos0 + i; wide[j].fw = r.x; wide[j].bw = r.y; wide[j].cov = r.z; wide[j].isfw = r.w; }
        }
    }
}

// ---- validate across GPUs: queries to the owners, rows back (SURVEY 8e "Collective") -------------------------------
//
// Phase 1 (origin GPU): every valid window of the chunk becomes a 16-byte query {key, position, neighbours} appended
// to the region this rank owns in the OWNER's inbox (peer-mapped memory: the stores are the all-to-all; cursors are
// local, appends warp-aggregated).  Windows without a key (non-ACGT) and positions past the last window get their
// row here.  Phase 2 (owner GPU): one thread per received query probes the owner's own table in its own HBM and
// stores the finished 16-byte DBGbase row straight into the origin's row slab (peer store).  Only rows and queries
// cross NVLink, in bulk-friendly streams; no GPU ever dereferences another GPU's table.

__device__ __forceinline__ uint4 evaluate_row(const Rec32 &kh, int found, uint32_t cutoff, bool isFw, uint32_t nxt, uint32_t prv,
                                              bool has_next, bool has_prev, unsigned long long &missing, unsigned long long &edge_missing) {
    uint4 o = make_uint4(0, 0, 0, 0);                                                // fw, bw, cov, isFw
    if (found) { o.z = kh.cov; o.w = isFw ? 1u : 0u; }                               // src/kreeq.cpp:168-169
    if (o.z == 0 || o.z < cutoff) { ++missing; return o; }                           // :172-175
    bool no_left = false, no_right = false;
    if (isFw) {                                                                      // :178-192
        if (has_next) { const uint32_t v = nxt <= 3 ? kh.fw[nxt] : 0; if (v) o.x = v; else no_right = true; }
        if (has_prev) { const uint32_t v = prv <= 3 ? kh.bw[prv] : 0; if (v) o.y = v; else no_left = true; }
    } else {                                                                         // :194-210
        if (has_prev) { const uint32_t v = prv <= 3 ? kh.fw[3 - prv] : 0; if (v) o.x = v; else no_left = true; }
        if (has_next) { const uint32_t v = nxt <= 3 ? kh.bw[3 - nxt] : 0; if (v) o.y = v; else no_right = true; }
    }
    if (no_left && no_right) ++edge_missing;                                         // :211-214
    return o;
}

__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
vroute_send_kernel(TableView tv, const uint64_t *__restrict__ B, const uint64_t *__restrict__ M, uint64_t base0, uint64_t c0,
                   uint64_t n, uint64_t kcount, Base *__restrict__ rows, unsigned long long *acc3, VRoute vr, uint64_t row0) {
    __shared__ unsigned long long sh[2];
    if (threadIdx.x < 2) sh[threadIdx.x] = 0;
    __syncthreads();
    unsigned long long missing = 0, counted = 0;
    const unsigned lane = threadIdx.x & 31;
    const uint64_t stride = (uint64_t)gridDim.x * TPB, rounds = (n + stride - 1) / stride;
    for (uint64_t r = 0; r < rounds; ++r) {                          // warp-uniform trip count (match_any below)
        const uint64_t i = r * stride + (uint64_t)blockIdx.x * TPB + threadIdx.x;
        int dest = -1; Kmer km; km.key = 0;
        if (i < n) {
            const uint64_t c = c0 + i;
            if (c < kcount) {
                km = kmer_at(B, M, c - base0, tv.k);
                if (km.valid) dest = (int)((km.key & 127) >> vr.shift);
                else { ++missing; ++counted; }                        // never found; still counts (kreeq.cpp:216)
            }
            if (dest < 0) __stcs(reinterpret_cast<uint4 *>(rows) + row0 + i, make_uint4(0, 0, 0, 0));
        }
        const unsigned peers = __match_any_sync(0xffffffffu, dest);
        if (dest < 0) continue;
        const unsigned leader = __ffs(peers) - 1;
        unsigned long long slot = 0;
        if (lane == leader) slot = atomicAdd(&vr.cursor[dest], (unsigned long long)__popc(peers));
        slot = __shfl_sync(peers, slot, leader) + __popc(peers & ((1u << lane) - 1));
        const uint64_t c = c0 + i;
        const unsigned meta = km.nxt | km.prv << 3 | (km.isFw ? 64u : 0u) | (c < kcount - 1 ? 128u : 0u) | (c > 0 ? 256u : 0u);
        uint4 q; q.x = (unsigned)km.key; q.y = (unsigned)(km.key >> 32); q.z = (unsigned)(row0 + i); q.w = meta;
        reinterpret_cast<uint4 *>(vr.inbox[dest])[(uint64_t)vr.me * vr.cap + slot] = q;
        vr.poslog[(uint64_t)dest * vr.cap + slot] = q.z;               // where the answer in slot `slot` of owner dest's box belongs
    }
    block_add(missing, &acc3[0], &sh[0]);
    block_add(counted, &acc3[1], &sh[1]);
}

// after the send kernel: tell every owner how many queries this rank left in its inbox, reset the cursors
__global__ void vroute_publish_kernel(VRoute vr) {
    const int o = threadIdx.x;
    if (o < vr.nranks) { vr.counts[o][vr.me] = vr.cursor[o]; vr.sent[o] = vr.cursor[o]; vr.cursor[o] = 0; }
}

__global__ void __launch_bounds__(TPB, 8)
vroute_answer_kernel(TableView tv, const uint4 *__restrict__ inbox, const unsigned long long *__restrict__ counts, uint32_t cutoff,
                     unsigned long long *acc3, VRoute vr) {
    __shared__ unsigned long long sh[3], first[KQ_MAX_RANKS + 1];
    if (threadIdx.x < 3) sh[threadIdx.x] = 0;
    if (threadIdx.x == 0) {
        unsigned long long run = 0;
        for (int s = 0; s < vr.nranks; ++s) { first[s] = run; run += counts[s]; }
        first[vr.nranks] = run;
    }
    __syncthreads();
    unsigned long long missing = 0, counted = 0, edge_missing = 0;
    const uint64_t total = first[vr.nranks], stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t j = (uint64_t)blockIdx.x * TPB + threadIdx.x; j < total; j += stride) {
        int s = 0;
        while (j >= first[s + 1]) ++s;
        const uint4 q = __ldcs(inbox + (uint64_t)s * vr.cap + (j - first[s]));
        const uint64_t key = (uint64_t)q.y << 32 | q.x;
        Rec32 kh; int found = lookup(tv, key, kh);
        if (found < 0) { atomicOr(tv.err, 4); found = 0; }
        const uint4 o = evaluate_row(kh, found, cutoff, (q.w & 64u) != 0, q.w & 7u, (q.w >> 3) & 7u, (q.w & 128u) != 0, (q.w & 256u) != 0,
                                     missing, edge_missing);
        ++counted;
        // answers go to this owner's box at the origin in query order: consecutive threads store consecutive 16-byte
        // rows (whole lines over NVLink); the origin puts them in their places (vroute_place_kernel)
        reinterpret_cast<uint4 *>(vr.rows[s])[vr.cap + (uint64_t)vr.me * vr.cap + (j - first[s])] = o;
    }
    block_add(missing, &acc3[0], &sh[0]);
    block_add(counted, &acc3[1], &sh[1]);
    block_add(edge_missing, &acc3[2], &sh[2]);
}

// origin side: answer j of owner o's box belongs to row poslog[o][j] (the boxes are read front to back, the rows of one
// box ascend: neighbouring threads write neighbouring lines)
__global__ void __launch_bounds__(TPB, 8)
vroute_place_kernel(VRoute vr) {
    __shared__ unsigned long long first[KQ_MAX_RANKS + 1];
    if (threadIdx.x == 0) {
        unsigned long long run = 0;
        for (int o = 0; o < vr.nranks; ++o) { first[o] = run; run += vr.sent[o]; }
        first[vr.nranks] = run;
    }
    __syncthreads();
    uint4 *rows = reinterpret_cast<uint4 *>(vr.rows[vr.me]);
    const uint64_t total = first[vr.nranks], stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t j = (uint64_t)blockIdx.x * TPB + threadIdx.x; j < total; j += stride) {
        int o = 0;
        while (j >= first[o + 1]) ++o;
        const uint64_t at = (uint64_t)o * vr.cap + (j - first[o]);
        rows[vr.poslog[at]] = __ldcs(rows + vr.cap + at);
    }
}

__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
lookup_kernel(TableView tv, const uint64_t *__restrict__ keys, uint64_t n, Rec32 *rec, uint8_t *found) {
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        Rec32 r; const int f = lookup(tv, keys[i], r);
        if (f < 0) atomicOr(tv.err, 4);
        found[i] = f > 0;
        if (f > 0) rec[i] = r;
    }
}

// ---- launchers ---------------------------------------------------------------------------------------

size_t scatter_smem_bytes(unsigned long long nslices, int cap, int tw) {
    return (size_t)nslices * cap * tw * 8 + (size_t)nslices * 4;
}

// bin capacity (tuples) of the scatter kernel's shared-memory bins; one CTA per SM owns them
int scatter_pick_cap(unsigned long long nslices, int tw, int max_cap) {
    static int max_smem = 0, sm_smem = 0;
    if (!max_smem) {
        int dev = 0; cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        cudaDeviceGetAttribute(&sm_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
        if (max_smem <= 0) max_smem = 227 * 1024;
        if (sm_smem <= 0) sm_smem = 228 * 1024;
    }
    if (const int cap = tuning().scatter_cap)                          // tuning knob
        if (cap <= max_cap && scatter_smem_bytes(nslices, cap, tw) <= (size_t)max_smem - 1024) return cap;
    // a tile brings NT * R = 8192 tuples: with few bins a small capacity would overflow several times per
    // tile (each time costs a round of barriers), so take the largest power of two that fits
    for (int cap : {512, 256, 128, 64, 32, 16, 8})
        if (cap <= max_cap && scatter_smem_bytes(nslices, cap, tw) <= (size_t)max_smem - 1024) return cap;
    return 0;
}

static inline uint64_t source_bound(const Source &src) { return src.tuples ? src.ntuples_max : src.npos; }

#define KQ_TW_DISPATCH(tw, CALL) do { if ((tw) == 1) { constexpr int TW = 1; CALL; } else { constexpr int TW = 2; CALL; } } while (0)

cudaError_t launch_count_direct(cudaStream_t st, const TableView &tv, const Source &src,
                                unsigned long long *instances, uint64_t *ovf, unsigned long long ovf_cap,
                                unsigned long long *ovf_n, uint64_t *launches) {
    const uint64_t n = source_bound(src);
    if (!n) return cudaSuccess;
    const unsigned grid = std::min<uint64_t>((n + TPB * RUN - 1) / (TPB * RUN), (uint64_t)sm_count() * 4);
    KQ_TW_DISPATCH(tuple_words(tv.k), (count_direct_kernel<TW><<<grid, TPB, 0, st>>>(tv, src, instances, ovf, ovf_cap, ovf_n)));
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_scatter_fine(cudaStream_t st, const TableView &tv, const Source &src, uint64_t *out, unsigned cap,
                                unsigned *cursor, int stride, unsigned long long *instances, uint64_t *ovf,
                                unsigned long long ovf_cap, unsigned long long *ovf_n, uint64_t *launches) {
    const uint64_t n = source_bound(src);
    if (!n) return cudaSuccess;
    const unsigned grid = std::min<uint64_t>((n + TPB * RUN - 1) / (TPB * RUN), (uint64_t)sm_count() * 4);
    KQ_TW_DISPATCH(tuple_words(tv.k), (scatter_fine_kernel<TW><<<grid, TPB, 0, st>>>(tv, src, out, cap, cursor, stride, instances, ovf, ovf_cap, ovf_n)));
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_slice_hist(cudaStream_t st, const TableView &tv, const Source &src, unsigned long long *hist,
                              uint64_t *launches) {
    const uint64_t n = source_bound(src);
    if (!n) return cudaSuccess;
    const unsigned grid = std::min<uint64_t>((n + TPB * RUN - 1) / (TPB * RUN), (uint64_t)sm_count() * 4);
    const size_t smem = (size_t)coarse_bins(tv) * 4;
    KQ_TW_DISPATCH(tuple_words(tv.k), (slice_hist_kernel<TW, false><<<grid, TPB, smem, st>>>(tv, src, hist, RouteMap{})));
    ++*launches;
    return cudaGetLastError();
}

// ---- routed (multi-GPU) pipeline: the same histogram and scatter kernels over (owner, hash range) bins ----

static TableView unsharded(const TableView &tv) { TableView v = tv; v.shard_rank = -1; return v; }   // route every k-mer

cudaError_t launch_route_hist(cudaStream_t st, const TableView &tv, const Source &src, const RouteMap &rt,
                              unsigned long long *hist, uint64_t *launches) {
    const uint64_t n = source_bound(src);
    const unsigned P = (unsigned)rt.nranks << rt.nb_bits;
    cudaError_t e = cudaMemsetAsync(hist, 0, (size_t)P * 8, st);
    if (e != cudaSuccess || !n) return e;
    const unsigned grid = std::min<uint64_t>((n + TPB * RUN - 1) / (TPB * RUN), (uint64_t)sm_count() * 4);
    KQ_TW_DISPATCH(tuple_words(tv.k), (slice_hist_kernel<TW, true><<<grid, TPB, (size_t)P * 4, st>>>(unsharded(tv), src, hist, rt)));
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_route_offsets(cudaStream_t st, const RouteMap &rt, int me, int cap, int tw, unsigned long long slab_cap,
                                 const unsigned long long *H, unsigned long long *front, unsigned long long *back,
                                 unsigned long long *recv, uint64_t *launches) {
    route_offsets_kernel<<<rt.nranks, 1024, 0, st>>>(rt, me, cap, tw, slab_cap, H, front, back, recv);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_route_scatter(cudaStream_t st, const TableView &tv, const Source &src, const RouteMap &rt,
                                 unsigned long long *cursor, unsigned long long *back, int cap, uint64_t *launches) {
    const uint64_t n = source_bound(src);
    if (!n) return cudaSuccess;
    const int tw = tuple_words(tv.k);
    const size_t smem = scatter_smem_bytes((unsigned long long)rt.nranks << rt.nb_bits, cap, tw);
    const TableView v = unsharded(tv);
    cudaError_t e = cudaSuccess;
#define KQ_ROUTE_LAUNCH(TWV, RV, NTV) do {                                                                           \
        e = cudaFuncSetAttribute(slice_scatter_kernel<TWV, RV, NTV, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                             \
        const unsigned grid = std::min<uint64_t>((n + NTV * RV - 1) / (NTV * RV), (uint64_t)sm_count());           \
        slice_scatter_kernel<TWV, RV, NTV, true><<<grid, NTV, smem, st>>>(v, src, nullptr, cursor, cap, rt, back, Regions{0, nullptr, 0, nullptr}); } while (0)
    const int variant = tuning().rt_variant;
    if (tw == 1) {
        if (variant == 1) KQ_ROUTE_LAUNCH(1, 8, 1024);
        else if (variant == 2) KQ_ROUTE_LAUNCH(1, 8, 512);
        else KQ_ROUTE_LAUNCH(1, 16, 512);
    } else {
        KQ_ROUTE_LAUNCH(2, 8, 512);
    }
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_slice_offsets(cudaStream_t st, unsigned long long nslices, const unsigned long long *hist,
                                 unsigned long long *cursor, unsigned long long *start, unsigned long long *total,
                                 unsigned long long *instances, uint64_t *launches) {
    slice_offsets_kernel<<<1, 1024, 0, st>>>(nslices, hist, cursor, start, total, instances);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_region_init(cudaStream_t st, unsigned nbins, unsigned long long rcap, unsigned long long *cursor,
                               unsigned long long *start, unsigned long long *end, unsigned long long *stat, uint64_t *launches) {
    region_init_kernel<<<(nbins + 255) / 256, 256, 0, st>>>(nbins, rcap, cursor, start, end, stat);
    ++*launches;
    return cudaGetLastError();
}
cudaError_t launch_region_counts(cudaStream_t st, unsigned nbins, unsigned long long *cursor, const unsigned long long *start,
                                 const unsigned long long *end, unsigned long long *hist, unsigned long long *total,
                                 const unsigned long long *stat, unsigned long long *instances, uint64_t *launches) {
    region_counts_kernel<<<1, 1024, 0, st>>>(nbins, cursor, start, end, hist, total, stat, instances);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_slice_scatter(cudaStream_t st, const TableView &tv, const Source &src, uint64_t *out,
                                 unsigned long long *cursor, int cap, uint64_t *launches, Regions rg) {
    const uint64_t n = source_bound(src);
    if (!n) return cudaSuccess;
    const int tw = tuple_words(tv.k);
    if (tuning().scatter_direct && !rg.rcap) {
        const unsigned g = std::min<uint64_t>((n + TPB * RUN - 1) / (TPB * RUN), (uint64_t)sm_count() * 4);
        KQ_TW_DISPATCH(tw, (slice_scatter_direct_kernel<TW><<<g, TPB, 0, st>>>(tv, src, out, cursor)));
        ++*launches;
        return cudaGetLastError();
    }
    const int variant = tuning().sc_variant;
    if (variant == 5 && !rg.rcap) {
        // ring form (measured slower than the two-pass form at one CTA per SM: 51.8 vs 31.6 ms on C2, the global
        // reservation of every flush round is on the critical path): bins of 3/2 cap tuples + two counters; the largest chunk of 64 / 32 / 16 / 8 tuples that fits
        static int max_smem = 0;
        if (!max_smem) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev); if (max_smem <= 0) max_smem = 227 * 1024; }
        const unsigned P = coarse_bins(tv);
        auto need = [&](unsigned c) { return (size_t)P * (c + c / 2) * tw * 8 + (size_t)P * 8; };
        unsigned rc = 0;
        for (unsigned c : {64u, 32u, 16u, 8u}) if (need(c) <= (size_t)max_smem - 1024) { rc = c; break; }
        if (rc) {
            const size_t smem_r = need(rc);
            cudaError_t er = cudaSuccess;
            const unsigned grid_r = (unsigned)std::min<uint64_t>((n + 512 * 8 - 1) / (512 * 8), (uint64_t)sm_count());
#define KQ_CS_GO(TWV, CAPV) do {                                                                                       \
            er = cudaFuncSetAttribute(coarse_scatter_kernel<TWV, 8, 512, CAPV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_r); \
            if (er != cudaSuccess) return er;                                                                          \
            coarse_scatter_kernel<TWV, 8, 512, CAPV><<<grid_r, 512, smem_r, st>>>(tv, src, out, cursor); } while (0)
#define KQ_CS_CAP(TWV) do { if (rc == 64) KQ_CS_GO(TWV, 64); else if (rc == 32) KQ_CS_GO(TWV, 32); else if (rc == 16) KQ_CS_GO(TWV, 16); else KQ_CS_GO(TWV, 8); } while (0)
            if (tw == 1) KQ_CS_CAP(1); else KQ_CS_CAP(2);
#undef KQ_CS_CAP
#undef KQ_CS_GO
            ++*launches;
            return cudaGetLastError();
        }
    }
    const size_t smem = scatter_smem_bytes(coarse_bins(tv), cap, tw);
    cudaError_t e = cudaSuccess;
#define KQ_SCATTER_LAUNCH(TWV, RV, NTV) do {                                                                        \
        e = cudaFuncSetAttribute(slice_scatter_kernel<TWV, RV, NTV, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                             \
        e = cudaFuncSetAttribute(slice_scatter_kernel<TWV, RV, NTV, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                             \
        const unsigned grid = std::min<uint64_t>((n + NTV * RV - 1) / (NTV * RV), (uint64_t)sm_count());           \
        if (rg.rcap) slice_scatter_kernel<TWV, RV, NTV, false, true><<<grid, NTV, smem, st>>>(tv, src, out, cursor, cap, RouteMap{}, nullptr, rg); \
        else slice_scatter_kernel<TWV, RV, NTV, false><<<grid, NTV, smem, st>>>(tv, src, out, cursor, cap, RouteMap{}, nullptr, rg); } while (0)
    if (tw == 1) {                                                   // the two-pass form (sc_variant 1..4)
        if (variant == 1) KQ_SCATTER_LAUNCH(1, 8, 512);
        else if (variant == 2) KQ_SCATTER_LAUNCH(1, 8, 1024);
        else if (variant == 3) KQ_SCATTER_LAUNCH(1, 4, 1024);
        else KQ_SCATTER_LAUNCH(1, 16, 512);
    } else {
        KQ_SCATTER_LAUNCH(2, 8, 512);
    }
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_insert_tuples(cudaStream_t st, const TableView &tv, const uint64_t *tuples,
                                 const unsigned long long *n, uint64_t n_max, uint64_t *ovf,
                                 unsigned long long ovf_cap, unsigned long long *ovf_n, unsigned long long *ticket,
                                 uint64_t *launches, uint64_t n_cap) {
    if (!n_max) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    // 8 CTAs of 256 threads = every thread slot of the SM (32 registers, a few spilled words): the kernel waits on
    // L2 round trips and its rate follows the warps in flight (3/4/5/6/7/8 CTAs: 125/101/86/76/70/67 ms on C2)
    const int ctas = tuning().insert_ctas;
    const unsigned grid = (unsigned)std::min<uint64_t>((n_max + INS_TILE - 1) / INS_TILE, (uint64_t)sm_count() * ctas);
    if (ctas > CTAS_PER_SM) KQ_TW_DISPATCH(tuple_words(tv.k), (insert_tuples_kernel<TW, 8><<<grid, TPB, 0, st>>>(tv, tuples, n, ovf, ovf_cap, ovf_n, ticket, n_cap ? n_cap : n_max)));
    else KQ_TW_DISPATCH(tuple_words(tv.k), (insert_tuples_kernel<TW, CTAS_PER_SM><<<grid, TPB, 0, st>>>(tv, tuples, n, ovf, ovf_cap, ovf_n, ticket, n_cap ? n_cap : n_max)));
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_extract_route(cudaStream_t st, const TableView &tv, const Source &src, RouteArgs ra,
                                 uint64_t *launches) {
    if (!src.npos) return cudaSuccess;
    const unsigned grid = std::min<uint64_t>((src.npos + TPB * RUN - 1) / (TPB * RUN), (uint64_t)sm_count() * 4);
    KQ_TW_DISPATCH(tuple_words(tv.k), (extract_route_kernel<TW><<<grid, TPB, 0, st>>>(tv, src, ra)));
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_validate(cudaStream_t st, const TableView &tv, const uint64_t *B, const uint64_t *M,
                            uint64_t base0, uint64_t c0, uint64_t n, uint64_t kcount, uint32_t cutoff,
                            Base *out, unsigned long long *acc3, const TableView *peers, int npeers, uint64_t *launches) {
    if (!n) return cudaSuccess;
    const int ctas = tuning().validate_ctas;                         // 8: 0.42 ms, 6: 0.45 ms per 1e7 positions
    const unsigned grid = (unsigned)std::min<uint64_t>((n + TPB - 1) / TPB, (uint64_t)sm_count() * ctas);
    if (ctas > CTAS_PER_SM) validate_kernel<8><<<grid, TPB, 0, st>>>(tv, B, M, base0, c0, n, kcount, cutoff, out, acc3, peers, npeers);
    else validate_kernel<CTAS_PER_SM><<<grid, TPB, 0, st>>>(tv, B, M, base0, c0, n, kcount, cutoff, out, acc3, peers, npeers);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_compact_rows(cudaStream_t st, const Base *rows, uint64_t n, uint64_t pos0, uint32_t *out4, void *wide,
                                unsigned long long wide_cap, unsigned long long *wide_n, uint64_t *launches) {
    if (!n) return cudaSuccess;
    compact_rows_kernel<<<grid_for(n), TPB, 0, st>>>(rows, n, pos0, out4, static_cast<WideRow *>(wide), wide_cap, wide_n);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_vroute_send(cudaStream_t st, const TableView &tv, const uint64_t *B, const uint64_t *M, uint64_t base0,
                               uint64_t c0, uint64_t n, uint64_t kcount, Base *rows, unsigned long long *acc3, const VRoute &vr,
                               uint64_t *launches, uint64_t row0, bool publish) {
    if (n) { vroute_send_kernel<<<grid_for(n), TPB, 0, st>>>(tv, B, M, base0, c0, n, kcount, rows, acc3, vr, row0); ++*launches; }
    if (publish) { vroute_publish_kernel<<<1, 32, 0, st>>>(vr); ++*launches; }   // also for an empty chunk: the owners read the counts
    return cudaGetLastError();
}

cudaError_t launch_vroute_answer(cudaStream_t st, const TableView &tv, const void *inbox, const unsigned long long *counts,
                                 uint64_t max_queries, uint32_t cutoff, unsigned long long *acc3, const VRoute &vr, uint64_t *launches) {
    if (!max_queries) return cudaSuccess;
    const unsigned grid = (unsigned)std::min<uint64_t>((max_queries + TPB - 1) / TPB, (uint64_t)sm_count() * 8);
    vroute_answer_kernel<<<grid, TPB, 0, st>>>(tv, static_cast<const uint4 *>(inbox), counts, cutoff, acc3, vr);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_vroute_place(cudaStream_t st, uint64_t max_queries, const VRoute &vr, uint64_t *launches) {
    if (!max_queries) return cudaSuccess;
    const unsigned grid = (unsigned)std::min<uint64_t>((max_queries + TPB - 1) / TPB, (uint64_t)sm_count() * 8);
    vroute_place_kernel<<<grid, TPB, 0, st>>>(vr);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_lookup(cudaStream_t st, const TableView &tv, const uint64_t *keys, uint64_t n, Rec32 *rec,
                          uint8_t *found, uint64_t *launches) {
    if (!n) return cudaSuccess;
    lookup_kernel<<<grid_for(n), TPB, 0, st>>>(tv, keys, n, rec, found);
    ++*launches;
    return cudaGetLastError();
}

}  // namespace kq

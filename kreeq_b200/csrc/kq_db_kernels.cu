// kq_db_kernels.cu -- the .kreeq file images are built and decoded on the GPU.
//
// A .map.<m>.bin file is a dump of phmap::parallel_flat_hash_map<uint64_t, DBGkmer> (phmap_dump.h:52-151):
//   u64 256, then per submap s:  u64 version (SIZE_MAX - 10), u64 size, u64 capacity, and iff size > 0:
//   int8 ctrl[capacity + 17], slot[capacity] (24 bytes: key, fw[4], bw[4], cov, 7 pad), u64 growth_left.
// Nothing but the first 8 bytes is aligned, so every field is written and read byte-wise.
// Save: records are bucketed by (map, submap) on the device, the host turns the 32 768 bucket counts into
// capacities and byte offsets, and three kernels lay the files out in one device image: headers + control
// bytes, placement (a record claims the first free position of its phmap probe sequence with a byte CAS --
// positions only ever go from empty to full, so the result is a table some sequential insertion order would
// have produced), then slot bytes.  Load is one kernel over the control bytes of an uploaded file image.
#include "kq_kernels.h"

namespace kq {

namespace {

constexpr int TPB = 256;
constexpr unsigned char kEmpty = 0x80, kSentinel = 0xFF;     // phmap.h ctrl_t: kEmpty = -128, kSentinel = -1
constexpr unsigned kWidth = 16;                              // Group::kWidth (SSE2 group)

__device__ __forceinline__ uint64_t phmap_hash_dev(uint64_t key) {                  // phmap_mix<8>: high + low of the 128-bit product
    return __umul64hi(key, 0xde5fb9d2630458e9ULL) + key * 0xde5fb9d2630458e9ULL;
}
__device__ __forceinline__ unsigned phmap_submap_dev(uint64_t h) {
    return (unsigned)(((h >> 8) ^ (h >> 16) ^ (h >> 24)) & 255);
}

__device__ __forceinline__ void put64(unsigned char *p, unsigned long long v) {
#pragma unroll
    for (int i = 0; i < 8; ++i) p[i] = (unsigned char)(v >> (8 * i));
}
__device__ __forceinline__ unsigned long long get64(const unsigned char *p) {
    unsigned long long v = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) v |= (unsigned long long)p[i] << (8 * i);
    return v;
}

// compare-and-swap of ONE byte through the aligned 32-bit word that holds it
__device__ __forceinline__ bool cas_byte(unsigned char *p, unsigned char expect, unsigned char val) {
    unsigned *w = reinterpret_cast<unsigned *>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
    const unsigned sh = (unsigned)(reinterpret_cast<uintptr_t>(p) & 3) * 8;
    unsigned cur = *reinterpret_cast<volatile unsigned *>(w);
    for (;;) {
        if (((cur >> sh) & 0xFFu) != expect) return false;
        const unsigned nw = (cur & ~(0xFFu << sh)) | ((unsigned)val << sh);
        const unsigned old = atomicCAS(w, cur, nw);
        if (old == cur) return true;
        cur = old;
    }
}

// ---- save -------------------------------------------------------------------------------------------

// records per (map, submap) bucket of the maps [lo, hi); records whose coverage does not fit the 8-bit map
// (>= 255 instances, graph-builder.cpp:160-206) are also appended to hc (their 32-bit form)
__global__ void __launch_bounds__(TPB)
db_count_kernel(const XRec *__restrict__ xs, unsigned long long n, int lo, int hi, unsigned long long *counts,
                XRec *hc, unsigned long long hc_cap, unsigned long long *hc_n) {
    const unsigned long long stride = (unsigned long long)gridDim.x * TPB;
    for (unsigned long long i = (unsigned long long)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        const uint64_t key = xs[i].key;
        const int m = (int)(key & 127);
        if (m < lo || m >= hi) continue;
        atomicAdd(&counts[(unsigned)m * 256 + phmap_submap_dev(phmap_hash_dev(key))], 1ULL);
        if (!(xs[i].flags & F_PLACEHOLDER) && xs[i].r.cov > 254) {
            const unsigned long long j = atomicAdd(hc_n, 1ULL);
            if (j < hc_cap) hc[j] = xs[i];
        }
    }
}

// one CTA per bucket: headers, control bytes, growth_left (the image was zeroed: slots and padding are 0)
__global__ void __launch_bounds__(TPB)
db_layout_kernel(unsigned char *img, const DbBucket *__restrict__ buckets, unsigned first, unsigned count) {
    for (unsigned b = blockIdx.x; b < count; b += gridDim.x) {
        const DbBucket d = buckets[first + b];
        if (threadIdx.x == 0) {
            if (((first + b) & 255u) == 0) put64(img + d.hdr_off - 8, 256);          // submap count opens the file
            put64(img + d.hdr_off, 0xFFFFFFFFFFFFFFF5ULL);          // version word: SIZE_MAX - 10, phmap_dump.h:47-48
            put64(img + d.hdr_off + 8, d.n); put64(img + d.hdr_off + 16, d.cap);
            if (d.n) put64(img + d.slot_off + d.cap * 24, d.cap - d.cap / 8 - d.n);
        }
        if (!d.n) continue;
        unsigned char *ctrl = img + d.ctrl_off;
        for (unsigned long long i = threadIdx.x; i < d.cap + kWidth + 1; i += TPB) ctrl[i] = i == d.cap ? kSentinel : kEmpty;
    }
}

// placement: find_first_non_full (phmap.h:2172-2183) with the claim made by a byte CAS; pos_out[i] = position
__global__ void __launch_bounds__(TPB)
db_place_kernel(const XRec *__restrict__ xs, unsigned long long n, int lo, int hi, unsigned char *img,
                const DbBucket *__restrict__ buckets, unsigned long long *pos_out) {
    const unsigned long long stride = (unsigned long long)gridDim.x * TPB;
    for (unsigned long long i = (unsigned long long)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        const uint64_t key = xs[i].key;
        const int m = (int)(key & 127);
        if (m < lo || m >= hi) continue;
        const uint64_t h = phmap_hash_dev(key);
        const DbBucket d = buckets[(unsigned)m * 256 + phmap_submap_dev(h)];
        unsigned char *ctrl = img + d.ctrl_off;
        const unsigned char h2 = (unsigned char)(h & 0x7F);
        unsigned long long off = (h >> 7) & d.cap, step = 0, pos = ~0ULL;
        while (pos == ~0ULL) {
            for (unsigned g = 0; g < kWidth; ++g) {
                const unsigned long long p = off + g;              // p == cap is the sentinel; p > cap mirrors p & cap
                if (p == d.cap) continue;
                const unsigned long long q = p & d.cap;
                if (*reinterpret_cast<volatile unsigned char *>(ctrl + q) == kEmpty && cas_byte(ctrl + q, kEmpty, h2)) { pos = q; break; }
            }
            step += kWidth; off = (off + step) & d.cap;            // probe_seq::next, phmap.h:160-164
        }
        pos_out[i] = pos;
    }
}

// slot bytes and the mirrored control byte (set_ctrl, phmap.h:2264-2276); 8-bit form of the record
__global__ void __launch_bounds__(TPB)
db_fill_kernel(const XRec *__restrict__ xs, unsigned long long n, int lo, int hi, unsigned char *img,
               const DbBucket *__restrict__ buckets, const unsigned long long *__restrict__ pos_in) {
    const unsigned long long stride = (unsigned long long)gridDim.x * TPB;
    for (unsigned long long i = (unsigned long long)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        const XRec x = xs[i];
        const int m = (int)(x.key & 127);
        if (m < lo || m >= hi) continue;
        const uint64_t h = phmap_hash_dev(x.key);
        const DbBucket d = buckets[(unsigned)m * 256 + phmap_submap_dev(h)];
        const unsigned long long pos = pos_in[i];
        unsigned char *ctrl = img + d.ctrl_off;
        ctrl[((pos - kWidth) & d.cap) + 1 + ((kWidth - 1) & d.cap)] = (unsigned char)(h & 0x7F);
        unsigned char *s = img + d.slot_off + pos * 24;
        put64(s, x.key);
        if (!(x.flags & F_PLACEHOLDER) && x.r.cov <= 254) {        // exact 8-bit record
#pragma unroll
            for (int w = 0; w < 4; ++w) { s[8 + w] = (unsigned char)x.r.fw[w]; s[12 + w] = (unsigned char)x.r.bw[w]; }
            s[16] = (unsigned char)x.r.cov;
        } else {
            s[16] = 255;                                            // placeholder: fw / bw unobservable, zeros
        }
    }
}

// ---- load -------------------------------------------------------------------------------------------

// every full control byte of the image's submaps becomes one raw record (cov == 255 -> placeholder)
__global__ void __launch_bounds__(TPB)
db_decode_kernel(const unsigned char *__restrict__ img, const DbBucket *__restrict__ buckets, unsigned nbuckets,
                 XRec *out, unsigned long long cap_out, unsigned long long *n_out) {
    const unsigned lane = threadIdx.x & 31;
    for (unsigned b = blockIdx.x; b < nbuckets; b += gridDim.x) {
        const DbBucket d = buckets[b];
        if (!d.n) continue;
        const unsigned char *ctrl = img + d.ctrl_off;
        const unsigned long long rounds = (d.cap + TPB - 1) / TPB;
        for (unsigned long long r = 0; r < rounds; ++r) {
            const unsigned long long i = r * TPB + threadIdx.x;
            const bool full = i < d.cap && !(ctrl[i] & 0x80);      // IsFull(ctrl) <=> ctrl >= 0
            const unsigned mask = __ballot_sync(0xffffffffu, full);
            if (!mask) continue;
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(n_out, (unsigned long long)__popc(mask));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (!full) continue;
            const unsigned long long j = base + __popc(mask & ((1u << lane) - 1));
            if (j >= cap_out) continue;
            const unsigned char *s = img + d.slot_off + i * 24;
            XRec x; x.key = get64(s);
#pragma unroll
            for (int w = 0; w < 4; ++w) { x.r.fw[w] = s[8 + w]; x.r.bw[w] = s[12 + w]; }
            x.r.cov = s[16];
            x.flags = x.r.cov == 255 ? F_PLACEHOLDER : 0u;
            out[j] = x;
        }
    }
}

// ---- load in table order ---------------------------------------------------------------------------------------
//
// slice_of_key is monotone in the slice hash, so records sorted by the top KQ_DB_ORDER_BITS bits of that hash are in
// slice order on the table that receives them whatever its geometry: put / merge kernels that walk such an array
// touch one narrow band of the table at a time (DRAM pages and TLB entries are reused, the band sits in L2), and
// a union applies database after database over the table in the same sweep order.  Two passes over the uploaded
// image: records per band, then decode straight to the record's place in band order (one L2 atomic per record).
__device__ __forceinline__ unsigned order_band(uint64_t key) {
    return (unsigned)((uint32_t)((key * 0x9E3779B97F4A7C15ULL) >> 32) >> (32 - KQ_DB_ORDER_BITS));
}

__global__ void __launch_bounds__(TPB)
db_band_hist_kernel(const unsigned char *__restrict__ img, const DbBucket *__restrict__ buckets, unsigned nbuckets,
                    unsigned long long *hist) {
    __shared__ unsigned h[1u << KQ_DB_ORDER_BITS];
    for (unsigned i = threadIdx.x; i < (1u << KQ_DB_ORDER_BITS); i += TPB) h[i] = 0;
    __syncthreads();
    for (unsigned b = blockIdx.x; b < nbuckets; b += gridDim.x) {
        const DbBucket d = buckets[b];
        if (!d.n) continue;
        const unsigned char *ctrl = img + d.ctrl_off;
        for (unsigned long long i = threadIdx.x; i < d.cap; i += TPB)
            if (!(ctrl[i] & 0x80)) atomicAdd(&h[order_band(get64(img + d.slot_off + i * 24))], 1u);
    }
    __syncthreads();
    for (unsigned i = threadIdx.x; i < (1u << KQ_DB_ORDER_BITS); i += TPB)
        if (h[i]) atomicAdd(&hist[i], (unsigned long long)h[i]);
}

__global__ void __launch_bounds__(TPB)
db_decode_ordered_kernel(const unsigned char *__restrict__ img, const DbBucket *__restrict__ buckets, unsigned nbuckets,
                         XRec *out, unsigned long long cap_out, unsigned long long *cursor) {
    for (unsigned b = blockIdx.x; b < nbuckets; b += gridDim.x) {
        const DbBucket d = buckets[b];
        if (!d.n) continue;
        const unsigned char *ctrl = img + d.ctrl_off;
        for (unsigned long long i = threadIdx.x; i < d.cap; i += TPB) {
            if (ctrl[i] & 0x80) continue;
            const unsigned char *s = img + d.slot_off + i * 24;
            XRec x; x.key = get64(s);
#pragma unroll
            for (int w = 0; w < 4; ++w) { x.r.fw[w] = s[8 + w]; x.r.bw[w] = s[12 + w]; }
            x.r.cov = s[16];
            x.flags = x.r.cov == 255 ? F_PLACEHOLDER : 0u;
            const unsigned long long j = atomicAdd(&cursor[order_band(x.key)], 1ULL);
            if (j < cap_out) out[j] = x;
        }
    }
}

// records with a field above what a slot lane holds: only these can create remainder / maps32 entries
__global__ void __launch_bounds__(TPB)
db_count_big_kernel(const XRec *__restrict__ xs, unsigned long long n, unsigned long long *out) {
    unsigned long long mine = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * TPB;
    for (unsigned long long i = (unsigned long long)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        const XRec x = xs[i];
        uint32_t mx = x.r.cov;
#pragma unroll
        for (int w = 0; w < 4; ++w) mx = max(mx, max(x.r.fw[w], x.r.bw[w]));
        mine += mx > KQ_LANE_MAX || (x.flags & F_PLACEHOLDER);
    }
    for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);   // 64-bit: a thread may carry more than 2^32
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(out, mine);
}

inline unsigned grid_for(unsigned long long n) {
    const unsigned long long want = (n + TPB - 1) / TPB, most = (unsigned long long)sm_count() * 8;
    return (unsigned)(want < most ? (want ? want : 1) : most);
}

}  // namespace

unsigned long long db_capacity_for(unsigned long long n) {           // smallest 2^j - 1 whose growth budget holds n
    unsigned long long cap = 1;
    while (cap - cap / 8 < n) cap = cap * 2 + 1;
    return cap;
}

cudaError_t launch_db_count(cudaStream_t st, const XRec *xs, unsigned long long n, int lo, int hi,
                            unsigned long long *counts, XRec *hc, unsigned long long hc_cap, unsigned long long *hc_n,
                            uint64_t *launches) {
    if (!n) return cudaSuccess;
    db_count_kernel<<<grid_for(n), TPB, 0, st>>>(xs, n, lo, hi, counts, hc, hc_cap, hc_n);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_db_build(cudaStream_t st, const XRec *xs, unsigned long long n, int lo, int hi, unsigned char *img,
                            const DbBucket *buckets, unsigned long long *pos_scratch, uint64_t *launches) {
    db_layout_kernel<<<(unsigned)(hi - lo) * 256, TPB, 0, st>>>(img, buckets, (unsigned)lo * 256, (unsigned)(hi - lo) * 256);
    ++*launches;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess || !n) return e;
    db_place_kernel<<<grid_for(n), TPB, 0, st>>>(xs, n, lo, hi, img, buckets, pos_scratch);
    db_fill_kernel<<<grid_for(n), TPB, 0, st>>>(xs, n, lo, hi, img, buckets, pos_scratch);
    *launches += 2;
    return cudaGetLastError();
}

cudaError_t launch_db_count_big(cudaStream_t st, const XRec *xs, unsigned long long n, unsigned long long *out, uint64_t *launches) {
    if (!n) return cudaSuccess;
    db_count_big_kernel<<<grid_for(n), TPB, 0, st>>>(xs, n, out);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_db_decode(cudaStream_t st, const unsigned char *img, const DbBucket *buckets, unsigned nbuckets,
                             XRec *out, unsigned long long cap_out, unsigned long long *n_out, uint64_t *launches) {
    if (!nbuckets) return cudaSuccess;
    db_decode_kernel<<<nbuckets < 65535u ? nbuckets : 65535u, TPB, 0, st>>>(img, buckets, nbuckets, out, cap_out, n_out);
    ++*launches;
    return cudaGetLastError();
}

// band histogram of the image's records -> hist[1 << KQ_DB_ORDER_BITS] (zeroed by the caller)
cudaError_t launch_db_band_hist(cudaStream_t st, const unsigned char *img, const DbBucket *buckets, unsigned nbuckets,
                                unsigned long long *hist, uint64_t *launches) {
    if (!nbuckets) return cudaSuccess;
    db_band_hist_kernel<<<std::min<unsigned>(nbuckets, (unsigned)sm_count() * 8), TPB, 0, st>>>(img, buckets, nbuckets, hist);
    ++*launches;
    return cudaGetLastError();
}
// decode into band order; cursor[] = exclusive scan of the band histogram (advanced by the kernel)
cudaError_t launch_db_decode_ordered(cudaStream_t st, const unsigned char *img, const DbBucket *buckets, unsigned nbuckets,
                                     XRec *out, unsigned long long cap_out, unsigned long long *cursor, uint64_t *launches) {
    if (!nbuckets) return cudaSuccess;
    db_decode_ordered_kernel<<<nbuckets < 65535u ? nbuckets : 65535u, TPB, 0, st>>>(img, buckets, nbuckets, out, cap_out, cursor);
    ++*launches;
    return cudaGetLastError();
}

}  // namespace kq

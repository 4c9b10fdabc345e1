// kq_kernels.cu -- hand-written sm_100a kernels of the kreeq k-mer hot path.
//
// All kernels are HBM-bound integer work (random 32-byte sector traffic on the table); no tensor
// cores are involved.  Grids are sized as a multiple of the SM count (148 on B200) times the
// resident CTAs per SM and loop grid-stride, so every SM keeps its full complement of warps in
// flight to cover the DRAM latency of the dependent probe -> atomic chain.
//
// Reference semantics each kernel follows (paths relative to the reference tree):
//   count_packed / extract_route : DBG::hashSequences   src/graph-builder.cpp:62-113
//   insert (accumulate)          : DBG::processBuffers  src/graph-builder.cpp:160-206 (closed form)
//   validate                     : DBG::evaluateSegment src/kreeq.cpp:143-219
//   stats_scan                   : DBG::summary         src/graph-builder.cpp:240-282
//   merge_ref                    : DBG::mergeSubMaps    src/graph-builder.cpp:353-432
#include "kq_kernels.h"

namespace kq {

constexpr int TPB = 256;          // threads per CTA
constexpr int CTAS_PER_SM = 6;    // 1536 resident threads per SM, 40 registers each

int sm_count() {
    static int n = 0;
    if (!n) {
        int dev = 0; cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

static inline unsigned grid_for(uint64_t n) {
    const uint64_t full = (uint64_t)sm_count() * CTAS_PER_SM;
    const uint64_t need = (n + TPB - 1) / TPB;
    return (unsigned)(need < full ? (need ? need : 1) : full);
}

// block-wide sum of per-thread counters, one global atomic per CTA
__device__ __forceinline__ void block_add(unsigned long long v, unsigned long long *dst, unsigned long long *sh) {
    v = __reduce_add_sync(0xffffffffu, (unsigned)v);   // per-thread partial sums stay < 2^32 (see callers)
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(sh, v);
    __syncthreads();
    if (threadIdx.x == 0 && *sh) atomicAdd(dst, *sh);
}

// one k-mer instance -> table (src/graph-builder.cpp:160-206)
__device__ __forceinline__ bool insert_instance(const TableView &tv, uint64_t key, uint32_t e) {
    bool is_new;
    Slot *s = find_or_claim(tv, key, &is_new);
    if (!s) return false;
    uint32_t fwi[4], bwi[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) { fwi[w] = (e >> (7 - w)) & 1u; bwi[w] = (e >> (3 - w)) & 1u; }
    accumulate(tv, s, key, 1u, fwi, bwi);
    return is_new;
}

// ---- count: fused extract + insert ----------------------------------------------------------

__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
count_packed_kernel(TableView tv, const uint64_t *__restrict__ B, const uint64_t *__restrict__ M,
                    uint64_t pos0, uint64_t npos, unsigned long long *instances) {
    __shared__ unsigned long long sh[2];
    if (threadIdx.x < 2) sh[threadIdx.x] = 0;
    __syncthreads();
    unsigned long long inst = 0, fresh = 0;
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < npos; i += stride) {
        const Kmer km = kmer_at(B, M, pos0 + i, tv.k);
        if (!km.valid || !owns(tv, km.key)) continue;
        ++inst;
        fresh += insert_instance(tv, km.key, edge_byte(km));
    }
    block_add(inst, instances, &sh[0]);
    block_add(fresh, &tv.occ[0], &sh[1]);
}

// ---- count: extract with routing (multi-GPU) --------------------------------------------------

__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
extract_route_kernel(TableView tv, const uint64_t *__restrict__ B, const uint64_t *__restrict__ M,
                     uint64_t pos0, uint64_t npos, RouteArgs ra, unsigned long long *instances) {
    __shared__ unsigned long long sh[2];
    if (threadIdx.x < 2) sh[threadIdx.x] = 0;
    __syncthreads();
    unsigned long long inst = 0, fresh = 0;
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    const uint64_t rounds = (npos + stride - 1) / stride;        // warp-uniform trip count
    const unsigned lane = threadIdx.x & 31;
    for (uint64_t r = 0; r < rounds; ++r) {
        const uint64_t i = r * stride + (uint64_t)blockIdx.x * TPB + threadIdx.x;
        int dest = -1; uint64_t key = 0; uint32_t e = 0;
        if (i < npos) {
            const Kmer km = kmer_at(B, M, pos0 + i, tv.k);
            if (km.valid) {
                key = km.key; e = edge_byte(km);
                dest = (int)((key & 127) >> ra.shift);
                if (dest == ra.my_rank) {                        // fused local insert
                    ++inst; fresh += insert_instance(tv, key, e); dest = -1;
                }
            }
        }
        // warp-aggregated append: one atomic per (warp, destination)
        const unsigned peers = __match_any_sync(0xffffffffu, dest);
        if (dest >= 0) {
            const unsigned leader = __ffs(peers) - 1;
            unsigned long long base = 0;
            if (lane == leader) base = atomicAdd(&ra.counts[dest], (unsigned long long)__popc(peers));
            base = __shfl_sync(peers, base, leader);
            const unsigned long long slot = base + __popc(peers & ((1u << lane) - 1));
            if (slot < ra.cap_per_rank) {
                ra.keys[(unsigned long long)dest * ra.cap_per_rank + slot] = key;
                ra.edges[(unsigned long long)dest * ra.cap_per_rank + slot] = (uint8_t)e;
            } else {
                atomicOr(tv.err, 8);
            }
        }
    }
    block_add(inst, instances, &sh[0]);
    block_add(fresh, &tv.occ[0], &sh[1]);
}

// ---- count: insert routed tuples ---------------------------------------------------------------

__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
insert_tuples_kernel(TableView tv, const uint64_t *__restrict__ keys, const uint8_t *__restrict__ edges,
                     uint64_t n, unsigned long long *instances) {
    __shared__ unsigned long long sh[2];
    if (threadIdx.x < 2) sh[threadIdx.x] = 0;
    __syncthreads();
    unsigned long long inst = 0, fresh = 0;
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        ++inst;
        fresh += insert_instance(tv, __ldg(keys + i), __ldg(edges + i));
    }
    block_add(inst, instances, &sh[0]);
    block_add(fresh, &tv.occ[0], &sh[1]);
}

cudaError_t launch_count_packed(cudaStream_t st, const TableView &tv, const uint64_t *B, const uint64_t *M,
                                uint64_t pos0, uint64_t npos, unsigned long long *instances, uint64_t *launches) {
    if (!npos) return cudaSuccess;
    count_packed_kernel<<<grid_for(npos), TPB, 0, st>>>(tv, B, M, pos0, npos, instances);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_extract_route(cudaStream_t st, const TableView &tv, const uint64_t *B, const uint64_t *M,
                                 uint64_t pos0, uint64_t npos, RouteArgs ra, unsigned long long *instances,
                                 uint64_t *launches) {
    if (!npos) return cudaSuccess;
    extract_route_kernel<<<grid_for(npos), TPB, 0, st>>>(tv, B, M, pos0, npos, ra, instances);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_insert_tuples(cudaStream_t st, const TableView &tv, const uint64_t *keys, const uint8_t *edges,
                                 uint64_t n, unsigned long long *instances, uint64_t *launches) {
    if (!n) return cudaSuccess;
    insert_tuples_kernel<<<grid_for(n), TPB, 0, st>>>(tv, keys, edges, n, instances);
    ++*launches;
    return cudaGetLastError();
}

// ---- validate ------------------------------------------------------------------------------------

// DBG::evaluateSegment src/kreeq.cpp:143-219 for positions [c0, c0+n) of one segment.
__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
validate_kernel(TableView tv, const uint64_t *__restrict__ B, const uint64_t *__restrict__ M,
                uint64_t base0, uint64_t c0, uint64_t n, uint64_t kcount, uint32_t cutoff,
                Base *__restrict__ out, unsigned long long *acc3) {
    __shared__ unsigned long long sh[3];
    if (threadIdx.x < 3) sh[threadIdx.x] = 0;
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
            if (km.valid ? owns(tv, km.key) : tv.shard_rank <= 0) {
                Rec32 kh; int found = 0;
                if (km.valid) {
                    found = lookup(tv, km.key, kh);
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
        reinterpret_cast<uint4 *>(out)[i] = o;                  // one coalesced 16-byte store per base
    }
    block_add(missing, &acc3[0], &sh[0]);
    block_add(counted, &acc3[1], &sh[1]);
    block_add(edge_missing, &acc3[2], &sh[2]);
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

cudaError_t launch_validate(cudaStream_t st, const TableView &tv, const uint64_t *B, const uint64_t *M,
                            uint64_t base0, uint64_t c0, uint64_t n, uint64_t kcount, uint32_t cutoff,
                            Base *out, unsigned long long *acc3, uint64_t *launches) {
    if (!n) return cudaSuccess;
    validate_kernel<<<grid_for(n), TPB, 0, st>>>(tv, B, M, base0, c0, n, kcount, cutoff, out, acc3);
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

// ---- stats / export ----------------------------------------------------------------------------------

// logical record of an occupied, non-placeholder slot
__device__ __forceinline__ void slot_logical(const TableView &tv, const Slot &s, Rec32 &r) {
    unsigned long long c = s.cov, f[4], b[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) { f[w] = (s.fw >> (16 * w)) & 0xFFFF; b[w] = (s.bw >> (16 * w)) & 0xFFFF; }
    if (s.flags & F_HAS_SIDE) {
        const SideSlot *q = side_find(tv, s.tag - 1);
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

// DBG::summary src/graph-builder.cpp:240-257 over the 8-bit side of the database
__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
stats_kernel(TableView tv, DevStats *ds, unsigned long long *hist, unsigned long long *big, unsigned long long big_cap) {
    __shared__ unsigned long long sh[4];
    if (threadIdx.x < 4) sh[threadIdx.x] = 0;
    __syncthreads();
    unsigned long long total = 0, uniq = 0, distinct = 0, edges = 0;
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < tv.cap; i += stride) {
        const Slot s = tv.T[i];
        if (s.tag == 0 || (s.flags & F_PLACEHOLDER)) continue;          // :246-247
        Rec32 r; slot_logical(tv, s, r);
        uniq += r.cov == 1;
#pragma unroll
        for (int w = 0; w < 4; ++w) edges += (r.fw[w] > 0 || r.bw[w] > 0);   // precedence quirk :253-254
        ++distinct;
        if (r.cov < (uint32_t)HIST_BINS) { atomicAdd(&hist[r.cov], 1ULL); total += r.cov; }
        else {
            const unsigned long long j = atomicAdd(&ds->big_n, 1ULL);
            if (j < big_cap) big[j] = r.cov;
        }
    }
    // per-thread totals can exceed 32 bits: add them directly, one atomic per warp
    for (int o = 16; o; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
    if ((threadIdx.x & 31) == 0 && total) atomicAdd(&ds->total, total);
    block_add(uniq, &ds->unique, &sh[1]);
    block_add(distinct, &ds->distinct, &sh[2]);
    block_add(edges, &ds->edges, &sh[3]);
}

__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
export_kernel(TableView tv, XRec *out, unsigned long long cap, unsigned long long *n_out) {
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    const unsigned lane = threadIdx.x & 31;
    const uint64_t rounds = (tv.cap + stride - 1) / stride;
    for (uint64_t r = 0; r < rounds; ++r) {
        const uint64_t i = r * stride + (uint64_t)blockIdx.x * TPB + threadIdx.x;
        Slot s; s.tag = 0;
        if (i < tv.cap) s = tv.T[i];
        const bool have = s.tag != 0;
        const unsigned m = __ballot_sync(0xffffffffu, have);
        if (!m) continue;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(n_out, (unsigned long long)__popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (!have) continue;
        const unsigned long long j = base + __popc(m & ((1u << lane) - 1));
        if (j >= cap) continue;
        XRec x; x.key = s.tag - 1; x.flags = s.flags;
        if (s.flags & F_PLACEHOLDER) {
#pragma unroll
            for (int w = 0; w < 4; ++w) { x.r.fw[w] = (uint32_t)(s.fw >> (16 * w)) & 0xFFFF; x.r.bw[w] = (uint32_t)(s.bw >> (16 * w)) & 0xFFFF; }
            x.r.cov = s.cov;
        } else {
            slot_logical(tv, s, x.r);
        }
        out[j] = x;
    }
}

cudaError_t launch_stats(cudaStream_t st, const TableView &tv, DevStats *ds, unsigned long long *hist,
                         unsigned long long *big, unsigned long long big_cap, uint64_t *launches) {
    stats_kernel<<<grid_for(tv.cap), TPB, 0, st>>>(tv, ds, hist, big, big_cap);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_export(cudaStream_t st, const TableView &tv, XRec *out, unsigned long long cap,
                          unsigned long long *n_out, uint64_t *launches) {
    export_kernel<<<grid_for(tv.cap), TPB, 0, st>>>(tv, out, cap, n_out);
    ++*launches;
    return cudaGetLastError();
}

// ---- grow / load / union -----------------------------------------------------------------------------

// move every occupied slot of the old table into tv.T (keys are unique: claim, then plain stores)
__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
rehash_kernel(TableView tv, const Slot *__restrict__ from, unsigned long long from_cap) {
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < from_cap; i += stride) {
        const Slot s = from[i];
        if (s.tag == 0) continue;
        bool is_new; Slot *d = find_or_claim(tv, s.tag - 1, &is_new);
        if (!d) continue;
        d->cov = s.cov; d->flags = s.flags; d->fw = s.fw; d->bw = s.bw;
    }
}

__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
side_rehash_kernel(TableView tv, const SideSlot *__restrict__ from, unsigned long long from_cap) {
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < from_cap; i += stride) {
        const SideSlot s = from[i];
        if (s.tag == 0) continue;
        SideSlot *d = side_find_or_claim(tv, s.tag - 1);
        if (!d) continue;
        d->kind = s.kind; d->cov = s.cov;
#pragma unroll
        for (int w = 0; w < 4; ++w) { d->fw[w] = s.fw[w]; d->bw[w] = s.bw[w]; }
    }
}

__device__ __forceinline__ unsigned long long pack16(const uint32_t v[4]) {
    return (unsigned long long)v[0] | ((unsigned long long)v[1] << 16) | ((unsigned long long)v[2] << 32) |
           ((unsigned long long)v[3] << 48);
}

// raw entries of an 8-bit map (Kmap::loadMap kmer.h:592-602): unique keys, values <= 255
__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
put_kernel(TableView tv, const XRec *__restrict__ in, uint64_t n) {
    __shared__ unsigned long long sh[1];
    if (threadIdx.x == 0) sh[0] = 0;
    __syncthreads();
    unsigned long long fresh = 0;
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        const XRec x = in[i];
        bool is_new; Slot *d = find_or_claim(tv, x.key, &is_new);
        if (!d) continue;
        fresh += is_new;
        d->cov = x.r.cov; d->flags = x.flags; d->fw = pack16(x.r.fw); d->bw = pack16(x.r.bw);
    }
    block_add(fresh, &tv.occ[0], &sh[0]);
}

// exact union: add the logical record x.r of a key (keys unique within one launch)
__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
add_records_kernel(TableView tv, const XRec *__restrict__ in, uint64_t n) {
    __shared__ unsigned long long sh[1];
    if (threadIdx.x == 0) sh[0] = 0;
    __syncthreads();
    unsigned long long fresh = 0;
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        const XRec x = in[i];
        bool is_new; Slot *d = find_or_claim(tv, x.key, &is_new);
        if (!d) continue;
        fresh += is_new;
        // the slot takes what fits its 16-bit lanes, the side table the rest
        const uint32_t room = d->cov >= KQ_LIMIT ? 0u : KQ_LIMIT - d->cov;
        const uint32_t m = min(x.r.cov, room);
        uint32_t fm[4], bm[4]; bool spill = m < x.r.cov;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            fm[w] = min(x.r.fw[w], m); bm[w] = min(x.r.bw[w], m);
            spill |= fm[w] < x.r.fw[w] || bm[w] < x.r.bw[w];
        }
        d->cov += m; d->fw += pack16(fm); d->bw += pack16(bm);
        if (spill) {
            SideSlot *q = side_find_or_claim(tv, x.key);
            if (!q) continue;
            q->kind |= KQ_SIDE_REMAINDER;
            q->cov += x.r.cov - m;
#pragma unroll
            for (int w = 0; w < 4; ++w) { q->fw[w] += x.r.fw[w] - fm[w]; q->bw[w] += x.r.bw[w] - bm[w]; }
            d->flags |= F_HAS_SIDE;
        }
    }
    block_add(fresh, &tv.occ[0], &sh[0]);
}

// DBG::mergeSubMaps src/graph-builder.cpp:353-432: one (key, a) of a later database's 8-bit map
// against the merged table (T = "submap2", S/KQ_SIDE_MAP32 = maps32[m]).  Keys are unique within
// one launch and launches follow the database order, so the result is the reference's.
__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
merge_ref_kernel(TableView tv, const XRec *__restrict__ in, uint64_t n) {
    __shared__ unsigned long long sh[1];
    if (threadIdx.x == 0) sh[0] = 0;
    __syncthreads();
    unsigned long long fresh = 0;
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < n; i += stride) {
        const XRec a = in[i];
        if (a.r.cov == 255) continue;                                        // :367-368
        SideSlot *q = side_find(tv, a.key);
        bool overflow = false;
        Slot *d = nullptr;
        if (q && (q->kind & KQ_SIDE_MAP32)) {                                // :370-372
            overflow = true;
        } else {
            bool is_new; d = find_or_claim(tv, a.key, &is_new);
            if (!d) continue;
            if (is_new) {                                                    // :375-377
                ++fresh;
                d->cov = a.r.cov; d->flags = 0; d->fw = pack16(a.r.fw); d->bw = pack16(a.r.bw);
            } else {
                uint32_t bf[4], bb[4];
#pragma unroll
                for (int w = 0; w < 4; ++w) { bf[w] = (uint32_t)(d->fw >> (16 * w)) & 0xFFFF; bb[w] = (uint32_t)(d->bw >> (16 * w)) & 0xFFFF; }
                if (255 - (int)d->cov <= (int)a.r.cov) overflow = true;      // :382-383
#pragma unroll
                for (int w = 0; w < 4; ++w)                                  // :385-391
                    if (255 - (int)bf[w] <= (int)a.r.fw[w] || 255 - (int)bb[w] <= (int)a.r.bw[w]) overflow = true;
                if (!overflow) {                                             // :393-400
                    d->cov += a.r.cov; d->fw += pack16(a.r.fw); d->bw += pack16(a.r.bw);
                } else {                                                     // :404-412 first spill of this key
                    q = side_find_or_claim(tv, a.key);
                    if (!q) continue;
                    q->kind |= KQ_SIDE_MAP32;
                    q->cov = d->cov;
#pragma unroll
                    for (int w = 0; w < 4; ++w) { q->fw[w] = bf[w]; q->bw[w] = bb[w]; }
                    d->cov = 255; d->flags |= F_PLACEHOLDER;
                }
            }
        }
        if (overflow) {                                                      // :414-427 (clamped on read)
            q->cov += a.r.cov;
#pragma unroll
            for (int w = 0; w < 4; ++w) { q->fw[w] += a.r.fw[w]; q->bw[w] += a.r.bw[w]; }
        }
    }
    block_add(fresh, &tv.occ[0], &sh[0]);
}

cudaError_t launch_rehash(cudaStream_t st, const TableView &tv, const Slot *from, unsigned long long from_cap,
                          uint64_t *launches) {
    rehash_kernel<<<grid_for(from_cap), TPB, 0, st>>>(tv, from, from_cap);
    ++*launches;
    return cudaGetLastError();
}
cudaError_t launch_side_rehash(cudaStream_t st, const TableView &tv, const SideSlot *from,
                               unsigned long long from_cap, uint64_t *launches) {
    side_rehash_kernel<<<grid_for(from_cap), TPB, 0, st>>>(tv, from, from_cap);
    ++*launches;
    return cudaGetLastError();
}
cudaError_t launch_put(cudaStream_t st, const TableView &tv, const XRec *in, uint64_t n, uint64_t *launches) {
    if (!n) return cudaSuccess;
    put_kernel<<<grid_for(n), TPB, 0, st>>>(tv, in, n);
    ++*launches;
    return cudaGetLastError();
}
cudaError_t launch_add_records(cudaStream_t st, const TableView &tv, const XRec *in, uint64_t n, uint64_t *launches) {
    if (!n) return cudaSuccess;
    add_records_kernel<<<grid_for(n), TPB, 0, st>>>(tv, in, n);
    ++*launches;
    return cudaGetLastError();
}
cudaError_t launch_merge_ref(cudaStream_t st, const TableView &tv, const XRec *in, uint64_t n, uint64_t *launches) {
    if (!n) return cudaSuccess;
    merge_ref_kernel<<<grid_for(n), TPB, 0, st>>>(tv, in, n);
    ++*launches;
    return cudaGetLastError();
}

}  // namespace kq

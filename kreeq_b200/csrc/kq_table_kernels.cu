// kq_table_kernels.cu -- table maintenance kernels: statistics, export, grow, load and union.
//
// Reference semantics (paths relative to the reference tree):
//   stats_kernel     : DBG::summary        src/graph-builder.cpp:240-282
//   put_kernel       : Kmap::loadMap       gfalibs/include/kmer.h:592-602 (raw 8-bit map entries)
//   merge_ref_kernel : DBG::mergeSubMaps   src/graph-builder.cpp:353-432
#include "kq_kernels.h"

namespace kq {

constexpr int TPB = 256;
constexpr int CTAS_PER_SM = 6;

static inline unsigned grid_for(uint64_t n) {
    const uint64_t full = (uint64_t)sm_count() * CTAS_PER_SM;
    const uint64_t need = (n + TPB - 1) / TPB;
    return (unsigned)(need < full ? (need ? need : 1) : full);
}

__device__ __forceinline__ void block_add(unsigned long long v, unsigned long long *dst, unsigned long long *sh) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);   // 64-bit: a thread may carry more than 2^32
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(sh, v);
    __syncthreads();
    if (threadIdx.x == 0 && *sh) atomicAdd(dst, *sh);
}

__device__ __forceinline__ unsigned long long table_slots(const TableView &tv) { return tv.nslices << tv.slice_bits; }

// ---- stats / export ----------------------------------------------------------------------------------

// DBG::summary src/graph-builder.cpp:240-257 over the 8-bit side of the database
__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
stats_kernel(TableView tv, DevStats *ds, unsigned long long *hist, unsigned long long *big, unsigned long long big_cap) {
    __shared__ unsigned long long sh[4];
    // coverages below 1024 (nearly every key: error k-mers at 1, genomic ones around the sequencing depth) are
    // counted in shared memory -- one global atomic per key on hist[1] serialises in L2
    constexpr unsigned LOW = 1024;
    __shared__ unsigned low[LOW];
    if (threadIdx.x < 4) sh[threadIdx.x] = 0;
    for (unsigned i = threadIdx.x; i < LOW; i += TPB) low[i] = 0;
    __syncthreads();
    unsigned long long total = 0, uniq = 0, distinct = 0, edges = 0;
    const uint64_t stride = (uint64_t)gridDim.x * TPB, cap = table_slots(tv);
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < cap; i += stride) {
        const Slot s = tv.T[i];
        if (s.tag == 0 || (s.cnt & CNT_PLACEHOLDER)) continue;          // :246-247
        Rec32 r; slot_logical(tv, s.tag - 1, s.cnt, r);
        uniq += r.cov == 1;
#pragma unroll
        for (int w = 0; w < 4; ++w) edges += (r.fw[w] > 0 || r.bw[w] > 0);   // precedence quirk :253-254
        ++distinct;
        if (r.cov < LOW) { atomicAdd(&low[r.cov], 1u); total += r.cov; }
        else if (r.cov < (uint32_t)HIST_BINS) { atomicAdd(&hist[r.cov], 1ULL); total += r.cov; }
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
    for (unsigned i = threadIdx.x; i < LOW; i += TPB)
        if (low[i]) atomicAdd(&hist[i], (unsigned long long)low[i]);
}

__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
export_kernel(TableView tv, XRec *out, unsigned long long cap, unsigned long long *n_out) {
    const uint64_t stride = (uint64_t)gridDim.x * TPB, slots = table_slots(tv);
    const unsigned lane = threadIdx.x & 31;
    const uint64_t rounds = (slots + stride - 1) / stride;
    for (uint64_t r = 0; r < rounds; ++r) {
        const uint64_t i = r * stride + (uint64_t)blockIdx.x * TPB + threadIdx.x;
        Slot s; s.tag = 0; s.cnt = 0;
        if (i < slots) s = tv.T[i];
        const bool have = s.tag != 0;
        const unsigned m = __ballot_sync(0xffffffffu, have);
        if (!m) continue;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(n_out, (unsigned long long)__popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (!have) continue;
        const unsigned long long j = base + __popc(m & ((1u << lane) - 1));
        if (j >= cap) continue;
        XRec x; x.key = s.tag - 1;
        if (s.cnt & CNT_PLACEHOLDER) {
            x.flags = F_PLACEHOLDER;
#pragma unroll
            for (int w = 0; w < 4; ++w) { x.r.fw[w] = 0; x.r.bw[w] = 0; }     // unobservable in the reference
            x.r.cov = 255;
        } else {
            x.flags = 0;
            slot_logical(tv, x.key, s.cnt, x.r);
        }
        out[j] = x;
    }
}

// ---- grow ----------------------------------------------------------------------------------------------

// move every occupied slot of the old table into tv.T (keys are unique: claim, then plain store)
__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
rehash_kernel(TableView tv, const Slot *__restrict__ from, unsigned long long from_cap) {
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < from_cap; i += stride) {
        const Slot s = from[i];
        if (s.tag == 0) continue;
        bool is_new; Slot *d = find_or_claim(tv, s.tag - 1, &is_new);
        if (!d) continue;
        d->cnt = s.cnt;
    }
}

__global__ void __launch_bounds__(TPB, CTAS_PER_SM)
side_rehash_kernel(TableView tv, int which, const SideSlot *__restrict__ from, unsigned long long from_cap) {
    const uint64_t stride = (uint64_t)gridDim.x * TPB;
    SideSlot *S = which == 1 ? tv.R : tv.M; const unsigned long long scap = which == 1 ? tv.rcap : tv.mcap;
    for (uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x; i < from_cap; i += stride) {
        const SideSlot s = from[i];
        if (s.tag == 0) continue;
        SideSlot *d = side_find_or_claim(tv, S, scap, which, s.tag - 1);
        if (!d) continue;
        d->cov = s.cov;
#pragma unroll
        for (int w = 0; w < 4; ++w) { d->fw[w] = s.fw[w]; d->bw[w] = s.bw[w]; }
    }
}

// ---- load / union ----------------------------------------------------------------------------------------

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
        if (x.flags & F_PLACEHOLDER) d->cnt = CNT_PLACEHOLDER;
        else slot_add_record(tv, d, x.key, x.r);
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
        slot_add_record(tv, d, x.key, x.r);
    }
    block_add(fresh, &tv.occ[0], &sh[0]);
}

// DBG::mergeSubMaps src/graph-builder.cpp:353-432: one (key, a) of a later database's 8-bit map
// against the merged table (T = "submap2", M = maps32[m]).  Keys are unique within one launch and
// launches follow the database order, so the result is the reference's.
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
        SideSlot *q = side_find(tv.M, tv.mcap, a.key);
        bool overflow = false;
        if (q) {                                                             // :370-372
            overflow = true;
        } else {
            bool is_new; Slot *d = find_or_claim(tv, a.key, &is_new);
            if (!d) continue;
            if (is_new) {                                                    // :375-377
                ++fresh;
                slot_add_record(tv, d, a.key, a.r);
            } else if (d->cnt & CNT_PLACEHOLDER) {
                // cov == 255 entry of the merged map whose maps32 record is absent: 255 - 255 <= a.cov
                // always overflows (:382-383) and creates the record from the placeholder's bytes,
                // which we do not keep (they are zero when written by us): widen(0s, cov 255)
                q = side_find_or_claim(tv, tv.M, tv.mcap, 2, a.key);
                if (!q) continue;
                q->cov = 255;
                overflow = true;
            } else {
                Rec32 b; slot_logical(tv, a.key, d->cnt, b);                  // the 8-bit entry (all fields <= 254)
                if (255 - (int)b.cov <= (int)a.r.cov) overflow = true;       // :382-383
#pragma unroll
                for (int w = 0; w < 4; ++w)                                  // :385-391
                    if (255 - (int)b.fw[w] <= (int)a.r.fw[w] || 255 - (int)b.bw[w] <= (int)a.r.bw[w]) overflow = true;
                if (!overflow) {                                             // :393-400
                    slot_add_record(tv, d, a.key, a.r);
                } else {                                                     // :404-412 first spill of this key
                    q = side_find_or_claim(tv, tv.M, tv.mcap, 2, a.key);
                    if (!q) continue;
                    q->cov = b.cov;
#pragma unroll
                    for (int w = 0; w < 4; ++w) { q->fw[w] = b.fw[w]; q->bw[w] = b.bw[w]; }
                    d->cnt = CNT_PLACEHOLDER;
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

// ---- launchers --------------------------------------------------------------------------------------------

cudaError_t launch_stats(cudaStream_t st, const TableView &tv, DevStats *ds, unsigned long long *hist,
                         unsigned long long *big, unsigned long long big_cap, uint64_t *launches) {
    stats_kernel<<<grid_for(tv.nslices << tv.slice_bits), TPB, 0, st>>>(tv, ds, hist, big, big_cap);
    ++*launches;
    return cudaGetLastError();
}
cudaError_t launch_export(cudaStream_t st, const TableView &tv, XRec *out, unsigned long long cap,
                          unsigned long long *n_out, uint64_t *launches) {
    export_kernel<<<grid_for(tv.nslices << tv.slice_bits), TPB, 0, st>>>(tv, out, cap, n_out);
    ++*launches;
    return cudaGetLastError();
}
cudaError_t launch_rehash(cudaStream_t st, const TableView &tv, const Slot *from, unsigned long long from_cap,
                          uint64_t *launches) {
    rehash_kernel<<<grid_for(from_cap), TPB, 0, st>>>(tv, from, from_cap);
    ++*launches;
    return cudaGetLastError();
}
cudaError_t launch_side_rehash(cudaStream_t st, const TableView &tv, int which, const SideSlot *from,
                               unsigned long long from_cap, uint64_t *launches) {
    side_rehash_kernel<<<grid_for(from_cap), TPB, 0, st>>>(tv, which, from, from_cap);
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

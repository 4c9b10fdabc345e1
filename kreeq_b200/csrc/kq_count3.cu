// kq_count3.cu -- the table side of the count pipeline: second scatter level and shared-memory insert.
//
// A table larger than 32 MiB is cut into slices of 2^12 slots (64 KiB): a key is probed inside its slice
// only, so ONE CTA can own a slice while it counts into it.  The level-1 scatter (kq_kernels.cu) brings the
// (key, edgeBit) tuples of a batch into COARSE order (2^cshift consecutive slices per bin); here
//   fine_plan_kernel     cuts the coarse bins into tasks of at most `part` tuples,
//   fine_hist_kernel     counts the tuples of every slice (shared-memory histogram per task),
//   fine_scatter_kernel  brings them into slice order through shared-memory ring bins, a full chunk leaves
//                        as ONE coalesced 256-byte store,
//   smem_insert_kernel   per slice: the 64 KiB table piece comes in with ONE bulk async copy
//                        (cp.async.bulk + mbarrier, SASS UBLKCP), the slice's tuples are counted with
//                        shared-memory CAS (exact 7-bit lanes, same closed form as insert_instance), and
//                        the piece goes back with one bulk store.
// The table is streamed through the SMs exactly once per batch and no k-mer instance costs an L2 atomic
// (the L2 CAS kernel insert_tuples_kernel sustains 35 G tuples/s on B200, profiles/r01c_ncu_insert_full.md).
//
// Reference semantics: DBG::processBuffers src/graph-builder.cpp:160-206 in closed form (exact sums; the
// 8-bit / 32-bit split is decided when the database is written).
#include "kq_kernels.h"

#include <algorithm>

namespace kq {

namespace {

template <int TW> __device__ __forceinline__ void t_unpack(const uint64_t *t, uint64_t i, uint64_t &key, uint32_t &e) {
    if (TW == 1) { const uint64_t w = __ldcs(t + i); key = w >> 8; e = (uint32_t)w & 255u; }
    else { const ulonglong2 w = __ldcs(reinterpret_cast<const ulonglong2 *>(t) + i); key = w.x; e = (uint32_t)w.y & 255u; }
}
template <int TW> __device__ __forceinline__ void t_store(uint64_t *t, uint64_t i, uint64_t key, uint32_t e) {
    if (TW == 1) t[i] = (key << 8) | e;
    else reinterpret_cast<ulonglong2 *>(t)[i] = make_ulonglong2(key, e);
}

// raw tuple words (prefetched one round ahead of their use)
template <int TW> __device__ __forceinline__ ulonglong2 t_raw(const uint64_t *t, uint64_t i) {
    if (TW == 1) return make_ulonglong2(__ldcs(t + i), 0ULL);
    return __ldcs(reinterpret_cast<const ulonglong2 *>(t) + i);
}
template <int TW> __device__ __forceinline__ void t_split(const ulonglong2 &w, uint64_t &key, uint32_t &e) {
    if (TW == 1) { key = w.x >> 8; e = (uint32_t)w.x & 255u; }
    else { key = w.x; e = (uint32_t)w.y & 255u; }
}
constexpr unsigned long long T_PAD1 = ~0ULL;       // a tuple word that unpacks to edge byte 0xFF (= "no tuple")

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

}  // namespace

// ---- plan -------------------------------------------------------------------------------------------------

// Coarse bin b of the level-1 output holds the tuples [start[b], start[b] + n[b]) of slices
// [b << cshift, (b + 1) << cshift).  One task = at most `part` consecutive tuples of one coarse bin.
__global__ void __launch_bounds__(1024, 1)
fine_plan_kernel(TableView tv, unsigned ncoarse, const unsigned long long *start, const unsigned long long *n,
                 unsigned long long part, FineTask *tasks, unsigned task_cap, unsigned *ntasks, unsigned *task_first) {
    __shared__ unsigned first[KQ_MAX_COARSE + 1];
    for (unsigned b = threadIdx.x; b < ncoarse; b += 1024) first[b] = (unsigned)((n[b] + part - 1) / part);
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned run = 0;
        for (unsigned b = 0; b < ncoarse; ++b) { const unsigned v = first[b]; first[b] = run; run += v; }
        first[ncoarse] = run;
        *ntasks = run < task_cap ? run : task_cap;                    // task_cap is an upper bound by construction
    }
    __syncthreads();
    for (unsigned b = threadIdx.x; b <= ncoarse; b += 1024) task_first[b] = first[b] < task_cap ? first[b] : task_cap;
    for (unsigned b = threadIdx.x; b < ncoarse; b += 1024) {
        const unsigned long long lo = start[b], hi = lo + n[b];
        const unsigned s_lo = b << tv.cshift;
        const unsigned ns = (unsigned)min((unsigned long long)(1u << tv.cshift), tv.nslices - s_lo);
        unsigned t = first[b];
        for (unsigned long long a = lo; a < hi && t < task_cap; a += part, ++t)
            tasks[t] = FineTask{a, min(hi, a + part), s_lo, ns};
    }
}

// Routed (multi-GPU) pipeline: the receive slab of rank `me` is laid out bin by bin ((owner, hash range) bins,
// route_offsets_kernel), every (sender, bin) region padded to whole flush chunks with sentinel tuples.  Bin j of
// this owner holds the keys whose slice hash lies in [j, j + 1) << (32 - nb_bits), i.e. slices [s_lo, s_hi] of the
// local table whatever its geometry: one task = at most `part` consecutive tuples of one bin's extent.
__global__ void __launch_bounds__(1024, 1)
route_plan_kernel(TableView tv, int nranks, int nb_bits, int me, int cap, const unsigned long long *H,
                  unsigned long long part, FineTask *tasks, unsigned task_cap, unsigned *ntasks, unsigned *task_first) {
    __shared__ unsigned long long ext[KQ_ROUTE_PLAN_BINS + 1];
    __shared__ unsigned first[KQ_ROUTE_PLAN_BINS + 1];
    const unsigned nb = 1u << nb_bits, P = (unsigned)nranks * nb;
    const unsigned long long c1 = (unsigned long long)cap - 1;
    for (unsigned j = threadIdx.x; j < nb; j += 1024) {
        unsigned long long L = 0;
        for (int r = 0; r < nranks; ++r) L += (H[(size_t)r * P + (unsigned)me * nb + j] + c1) / cap * cap;
        ext[j] = L; first[j] = (unsigned)((L + part - 1) / part);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long run = 0; unsigned t = 0;
        for (unsigned j = 0; j < nb; ++j) { const unsigned long long L = ext[j]; ext[j] = run; run += L; const unsigned v = first[j]; first[j] = t; t += v; }
        ext[nb] = run; first[nb] = t;
        *ntasks = t < task_cap ? t : task_cap;
    }
    __syncthreads();
    for (unsigned j = threadIdx.x; j <= nb; j += 1024) task_first[j] = first[j] < task_cap ? first[j] : task_cap;
    for (unsigned j = threadIdx.x; j < nb; j += 1024) {
        const unsigned long long lo = ext[j], hi = ext[j + 1];
        const unsigned long long h_lo = (unsigned long long)j << (32 - nb_bits), h_hi = (((unsigned long long)j + 1) << (32 - nb_bits)) - 1;
        const unsigned s_lo = (unsigned)((h_lo * tv.nslices) >> 32), s_hi = (unsigned)((h_hi * tv.nslices) >> 32);
        unsigned t = first[j];
        for (unsigned long long a = lo; a < hi && t < task_cap; a += part, ++t)
            tasks[t] = FineTask{a, min(hi, a + part), s_lo, s_hi - s_lo + 1};
    }
}

// ---- tuples per slice ---------------------------------------------------------------------------------------

constexpr int FH_NT = 512;

template <int TW>
__global__ void __launch_bounds__(FH_NT, 2)
fine_hist_kernel(TableView tv, const uint64_t *__restrict__ in, const FineTask *__restrict__ tasks,
                 const unsigned *__restrict__ ntasks, unsigned long long *fhist, unsigned *hist2, unsigned ns_max) {
    extern __shared__ unsigned int fh[];
    const unsigned nt = *ntasks;
    for (unsigned t = blockIdx.x; t < nt; t += gridDim.x) {
        const FineTask tk = tasks[t];
        for (unsigned i = threadIdx.x; i < tk.ns; i += FH_NT) fh[i] = 0;
        __syncthreads();
        for (unsigned long long i0 = tk.lo; i0 < tk.hi; i0 += FH_NT * 4) {
            uint64_t key[4]; uint32_t e[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const unsigned long long i = i0 + (unsigned)j * FH_NT + threadIdx.x;
                e[j] = 255u;
                if (i < tk.hi) t_unpack<TW>(in, i, key[j], e[j]);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (e[j] != 255u) atomicAdd(&fh[(unsigned)slice_of_key(key[j], tv.nslices) - tk.s_lo], 1u);
        }
        __syncthreads();
        if (hist2) {                                                     // per task: fine_offsets_kernel turns them into positions
            for (unsigned i = threadIdx.x; i < tk.ns; i += FH_NT) hist2[(size_t)t * ns_max + i] = fh[i];
        } else {
            for (unsigned i = threadIdx.x; i < tk.ns; i += FH_NT)
                if (fh[i]) atomicAdd(&fhist[tk.s_lo + i], (unsigned long long)fh[i]);
        }
        __syncthreads();
    }
}

// One CTA per coarse bin b: hist2[t][f] (tuples of slice s_lo + f in task t) becomes the offset of task t inside the
// slice's region, fcount[s] the slice's total and fstart[s] its first tuple (regions follow each other inside the
// coarse bin's range of the output, which starts where the bin starts in the level-1 output).
__global__ void __launch_bounds__(1024, 1)
fine_offsets_kernel(TableView tv, const unsigned long long *__restrict__ start, const unsigned *__restrict__ task_first,
                    unsigned *hist2, unsigned ns_max, unsigned long long *fstart, unsigned long long *fcount) {
    extern __shared__ unsigned long long tot[];
    const unsigned b = blockIdx.x, s_lo = b << tv.cshift;
    const unsigned ns = (unsigned)min((unsigned long long)(1u << tv.cshift), tv.nslices - s_lo);
    const unsigned t0 = task_first[b], t1 = task_first[b + 1];
    for (unsigned f = threadIdx.x; f < ns; f += 1024) {
        unsigned run = 0;
        for (unsigned t = t0; t < t1; ++t) { unsigned *p = hist2 + (size_t)t * ns_max + f; const unsigned v = *p; *p = run; run += v; }
        tot[f] = run; fcount[s_lo + f] = run;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long run = start[b];
        for (unsigned f = 0; f < ns; ++f) { const unsigned long long v = tot[f]; tot[f] = run; run += v; }
    }
    __syncthreads();
    for (unsigned f = threadIdx.x; f < ns; f += 1024) fstart[s_lo + f] = tot[f];
}

// Routed slab: the bins are hash ranges, and a slice can lie across the border of two of them (or span many, on a small
// table).  One thread per slice s: the tasks of all bins whose hash range meets the slice's, in task order --
// hist2[t][s - s_lo(t)] becomes the offset of task t inside the slice's region, fcount[s] the slice's total
// (slice_offsets_kernel then turns the totals into fstart).
__global__ void __launch_bounds__(256)
route_fine_offsets_kernel(TableView tv, int nb_bits, const FineTask *__restrict__ tasks, const unsigned *__restrict__ task_first,
                          unsigned *hist2, unsigned ns_max, unsigned long long *fcount) {
    const unsigned long long ns = tv.nslices;
    for (unsigned long long s = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; s < ns; s += (unsigned long long)gridDim.x * blockDim.x) {
        // hashes h with (h * ns) >> 32 == s: [ceil(s 2^32 / ns), ceil((s + 1) 2^32 / ns) - 1]
        const unsigned long long h_lo = ((s << 32) + ns - 1) / ns, h_hi = (((s + 1) << 32) + ns - 1) / ns - 1;
        const unsigned j_lo = (unsigned)(h_lo >> (32 - nb_bits)), j_hi = (unsigned)(h_hi >> (32 - nb_bits));
        unsigned long long run = 0;
        for (unsigned t = task_first[j_lo]; t < task_first[j_hi + 1]; ++t) {
            const unsigned s_lo = tasks[t].s_lo;
            if (s < s_lo || s >= (unsigned long long)s_lo + tasks[t].ns) continue;
            unsigned *p = hist2 + (size_t)t * ns_max + (unsigned)(s - s_lo);
            const unsigned v = *p; *p = (unsigned)run; run += v;
        }
        fcount[s] = run;
    }
}

// ---- scatter into slice order ----------------------------------------------------------------------------------

// Shared memory: `ns` ring bins of RING = 3/2 CAP tuples, an append counter and a flushed counter per bin.
// A tuple is appended with one shared-memory atomic and one store; a bin that holds CAP tuples leaves as one
// coalesced CAP * 8 * TW byte store.  The ring takes what arrives while a chunk waits for its flush, so a round of
// the CTA needs two barriers and (practically) never a second pass.
// Where a chunk goes: PRIV (local pipeline) -- every task owns its part of every slice's region (fine_offsets_kernel
// turned the per-task histograms into start positions), the cursors live in shared memory and a flush costs no global
// round trip; !PRIV (routed slab, whose bins overlap at their edges) -- one global atomic on the slice's cursor.
template <int TW, int NT, int R, int CAP, bool PRIV>
__global__ void __launch_bounds__(NT, NT >= 1024 ? 1 : 2)
fine_scatter_kernel(TableView tv, const uint64_t *__restrict__ in, const FineTask *__restrict__ tasks,
                    const unsigned *__restrict__ ntasks, uint64_t *__restrict__ out, unsigned long long *cursor,
                    const unsigned *__restrict__ hist2, const unsigned long long *__restrict__ fstart, unsigned ns_max) {
    constexpr unsigned RING = CAP + CAP / 2;
    extern __shared__ uint64_t sm3[];
    uint64_t *bins = sm3;
    unsigned long long *cur = reinterpret_cast<unsigned long long *>(bins + (size_t)ns_max * RING * TW);   // PRIV only
    unsigned int *fill = reinterpret_cast<unsigned int *>(cur + (PRIV ? ns_max : 0));
    unsigned int *done = fill + ns_max;
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = NT / 32;
    const unsigned nt = *ntasks;
    for (unsigned t = blockIdx.x; t < nt; t += gridDim.x) {
        const FineTask tk = tasks[t];
        for (unsigned i = threadIdx.x; i < tk.ns; i += NT) {
            fill[i] = 0; done[i] = 0;
            if (PRIV) cur[i] = fstart[tk.s_lo + i] + hist2[(size_t)t * ns_max + i];
        }
        __syncthreads();
        ulonglong2 nxt[R];                                               // the next round's tuples, loaded a round ahead
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const unsigned long long j = tk.lo + (unsigned)i * NT + threadIdx.x;
            nxt[i] = j < tk.hi ? t_raw<TW>(in, j) : make_ulonglong2(T_PAD1, T_PAD1);
        }
        for (unsigned long long tile = tk.lo; tile < tk.hi; tile += (unsigned long long)NT * R) {
            uint64_t key[R]; uint32_t e[R]; unsigned bin[R], pos[R];
            uint32_t m = 0;
#pragma unroll
            for (int i = 0; i < R; ++i) {
                t_split<TW>(nxt[i], key[i], e[i]);
                const unsigned long long j = tile + (unsigned long long)NT * R + (unsigned)i * NT + threadIdx.x;
                nxt[i] = j < tk.hi ? t_raw<TW>(in, j) : make_ulonglong2(T_PAD1, T_PAD1);
            }
#pragma unroll
            for (int i = 0; i < R; ++i) {
                if (e[i] == 255u) continue;                              // padding of the routed layout
                bin[i] = (unsigned)slice_of_key(key[i], tv.nslices) - tk.s_lo;
                pos[i] = atomicAdd(&fill[bin[i]], 1u);
                m |= 1u << i;
            }
            for (;;) {
                bool full = false;
#pragma unroll
                for (int i = 0; i < R; ++i) {
                    if (!(m >> i & 1u)) continue;
                    const unsigned d = pos[i] - *reinterpret_cast<volatile unsigned int *>(&done[bin[i]]);
                    if (d < RING) { t_store<TW>(bins, bin[i] * RING + pos[i] % RING, key[i], e[i]); m &= ~(1u << i); }
                    full |= d + 1 >= (unsigned)CAP;
                }
                if (!__syncthreads_or(full)) break;
                // lane l of a warp looks at bin b0 + l, then the warp copies the full chunks of its 32 bins one by one
                for (unsigned b0 = warp * 32; b0 < tk.ns; b0 += nwarps * 32) {
                    const unsigned b = b0 + lane;
                    unsigned d0 = 0, chunks = 0;
                    if (b < tk.ns) { d0 = done[b]; chunks = min(fill[b] - d0, RING) / CAP; }
                    unsigned long long g = 0;
                    if (chunks) {
                        if (PRIV) { g = cur[b]; cur[b] = g + chunks * CAP; }
                        else g = atomicAdd(&cursor[tk.s_lo + b], (unsigned long long)chunks * CAP);
                        done[b] = d0 + chunks * CAP;
                    }
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
                if (!__syncthreads_or(m != 0)) break;                    // nobody had to wait for a flush
            }
        }
        __syncthreads();
        for (unsigned b0 = warp * 32; b0 < tk.ns; b0 += nwarps * 32) {       // what is left in the bins
            const unsigned b = b0 + lane;
            unsigned d0 = 0, f = 0;
            if (b < tk.ns) { d0 = done[b]; f = fill[b] - d0; }
            unsigned long long g = 0;
            if (f) g = PRIV ? cur[b] : atomicAdd(&cursor[tk.s_lo + b], (unsigned long long)f);
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
        __syncthreads();
    }
}

// ---- shared-memory insert -----------------------------------------------------------------------------------------

namespace {

__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
    asm volatile(
        "{\n .reg .pred p;\n"
        "KQ_WAIT_%=:\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra KQ_DONE_%=;\n"
        " bra KQ_WAIT_%=;\n"
        "KQ_DONE_%=:\n}"
        ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk async copy, completion reported to the mbarrier in bytes
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src, unsigned bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// shared -> global bulk async copy (bulk group)
__device__ __forceinline__ void bulk_s2g(void *dst, const void *src_smem, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// a whole range starts its way into L2 (address and size multiples of 16)
__device__ __forceinline__ void bulk_prefetch_l2(const void *src, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
// 64-bit CAS on a shared-memory address (shared window offset)
__device__ __forceinline__ unsigned long long atom_cas_shared(uint32_t addr, unsigned long long cmp, unsigned long long val) {
    unsigned long long old;
    asm volatile("atom.shared.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "r"(addr), "l"(cmp), "l"(val) : "memory");
    return old;
}

// Slow paths of the shared-memory insert, kept out of line so that the hot loop stays small.

// cov + n would pass a slot lane: move 63 of the coverage and up to 63 of every lane to the remainder table in the
// same CAS that counts these instances (the slot stays >= KQ_REM_FLOOR: readers then look into R)
__device__ __noinline__ void smem_spill(unsigned long long *pcnt, const TableView &tv, uint64_t key, unsigned long long inc) {
    const unsigned n = (unsigned)(inc >> 56);
    unsigned long long cur = *reinterpret_cast<volatile unsigned long long *>(pcnt);
    for (;;) {
        if (cnt_cov(cur) + n <= KQ_LANE_MAX) {                             // somebody else made room meanwhile
            const unsigned long long old = atomicCAS(pcnt, cur, cur + inc);
            if (old == cur) return;
            cur = old; continue;
        }
        uint32_t fw[4], bw[4];
        unsigned long long take = 63ull << 56;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            fw[w] = min(cnt_lane(cur, w), 63u); bw[w] = min(cnt_lane(cur, 4 + w), 63u);
            take |= (unsigned long long)fw[w] << (7 * w) | (unsigned long long)bw[w] << (7 * (4 + w));
        }
        const unsigned long long old = atomicCAS(pcnt, cur, cur - take + inc);
        if (old == cur) { remainder_add(tv, key, 63, fw, bw); return; }
        cur = old;
    }
}

// the slice is full: the instances go back to the overflow list as single tuples (instance c of a combined group
// carries edge b iff c < count of b, which preserves every sum)
template <int TW>
__device__ __noinline__ void smem_give_back(uint64_t *ovf, unsigned long long ovf_cap, unsigned long long *ovf_n, uint64_t key, unsigned long long inc) {
    const unsigned cnt = (unsigned)(inc >> 56);
    const unsigned long long q = atomicAdd(ovf_n, (unsigned long long)cnt);
    for (unsigned c = 0; c < cnt; ++c) {
        uint32_t eb = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) if (c < cnt_lane(inc, b)) eb |= 0x80u >> b;
        if (q + c < ovf_cap) t_store<TW>(ovf, q + c, key, eb);
    }
}

// complete insert of one (combined) tuple: probe, claim, count; 1 = new key, 0 = existed / given back
template <int TW>
__device__ __noinline__ unsigned smem_insert_slow(Slot *piece, unsigned mask, unsigned limit, const TableView &tv, uint64_t key,
                                                  unsigned long long inc, uint64_t *ovf, unsigned long long ovf_cap, unsigned long long *ovf_n) {
    const unsigned long long tag = key + 1;
    const unsigned n = (unsigned)(inc >> 56);
    unsigned j = (unsigned)home_hash(key) & mask, fresh = 0;
    for (unsigned probes = 0; probes < limit; ++probes) {
        unsigned long long t = *reinterpret_cast<volatile unsigned long long *>(&piece[j].tag);
        if (t == 0) {
            t = atomicCAS(&piece[j].tag, 0ULL, tag);
            if (t == 0) { t = tag; fresh = 1; }
        }
        if (t == tag) {
            unsigned long long cur = *reinterpret_cast<volatile unsigned long long *>(&piece[j].cnt);
            for (;;) {
                if (cnt_cov(cur) + n > KQ_LANE_MAX) { smem_spill(&piece[j].cnt, tv, key, inc); return fresh; }
                const unsigned long long old = atomicCAS(&piece[j].cnt, cur, cur + inc);   // lanes <= cov: no carry between lanes
                if (old == cur) return fresh;
                cur = old;
            }
        }
        j = (j + 1) & mask;
    }
    smem_give_back<TW>(ovf, ovf_cap, ovf_n, key, inc);
    return 0;
}

// runs of one key in the stream (poly-A, satellites): the lanes of a warp that hold the same key add their
// increments up and the lowest lane counts for the group, so a hot key costs one CAS per warp, not a CAS storm
__device__ __noinline__ unsigned long long warp_combine(uint64_t key, uint32_t e, bool live) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    unsigned long long sum = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        const unsigned has = __ballot_sync(0xffffffffu, live && ((e >> (7 - b)) & 1u));
        sum |= (unsigned long long)__popc(has & peers) << (7 * b);
    }
    sum |= (unsigned long long)__popc(peers) << 56;
    return live && lane == (unsigned)__ffs(peers) - 1 ? sum : 0ULL;
}

}  // namespace

// One CTA per slice at a time (ticket counter; slices without tuples are skipped): bulk-load the slice (UBLKCP),
// count its tuples in shared memory, bulk-store it.  Three CTAs per SM overlap each other's copies.
// A warp owns a contiguous share of the slice's tuples and its lanes POP tuples from it: every lane holds the tuple
// it is counting and the next one (already loaded); an iteration is ONE probe step per lane -- slot load, claim or
// compare, one CAS -- and the lanes that finished take the next unread tuples of the share (ballot + prefix count,
// no atomics).  Lanes never wait for the longest probe sequence of the warp and stay busy until the share ends.
// Four or more lanes losing their CAS in one iteration means a run of one key in the stream: the warp combines.
template <int TW, int NT>
__global__ void __launch_bounds__(NT, 3)
smem_insert_kernel(TableView tv, const uint64_t *__restrict__ tuples, const unsigned long long *__restrict__ fstart,
                   const void *__restrict__ fcount_, unsigned long long cap, int stride, uint64_t *ovf,
                   unsigned long long ovf_cap, unsigned long long *ovf_n, unsigned long long *ticket, HotList hot) {
    const unsigned long long *fcount = static_cast<const unsigned long long *>(fcount_);     // level-2 layout: exact counts
    const unsigned *fcur = static_cast<const unsigned *>(fcount_);                             // one-pass layout: 32-bit cursors
    extern __shared__ __align__(128) unsigned char sm_piece[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ unsigned long long sh_ticket, sh_fresh, sh_n;
    __shared__ unsigned long long inc_lut[256];
    Slot *piece = reinterpret_cast<Slot *>(sm_piece);
    const unsigned slots = 1u << tv.slice_bits, mask = slots - 1, bytes = slots * (unsigned)sizeof(Slot);
    const unsigned limit = min(slots, 1024u);                          // a longer run means the slice is (locally) full
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, lt_mask = (1u << lane) - 1;
    const uint32_t piece_a = smem_u32(piece);
    for (unsigned i = threadIdx.x; i < 256; i += NT) inc_lut[i] = inc_of_edge(i);
    if (threadIdx.x == 0) { mbar_init(&bar, 1); sh_fresh = 0; asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    unsigned parity = 0;
    unsigned fresh = 0;
    for (;;) {
        if (threadIdx.x == 0) {
            // next slice that has tuples (the counts are read by one thread: most slices of a small batch are empty)
            unsigned long long s, cnt = 0;
            for (;;) {
                s = atomicAdd(ticket, 1ULL); if (s >= tv.nslices) break;
                cnt = fstart ? fcount[s] : min((unsigned long long)fcur[s * stride], cap);
                if (!cnt) continue;
                if (hot.buf && cnt > hot.limit) {                       // a run of one key: left to hot_slices_kernel (all SMs)
                    const unsigned long long j = atomicAdd(&hot.buf[0], 1ULL);
                    if (j < hot.cap) { hot.buf[1 + 3 * j] = fstart ? fstart[s] : s * cap; hot.buf[2 + 3 * j] = cnt; hot.buf[3 + 3 * j] = s; continue; }
                }
                break;
            }
            sh_ticket = s; sh_n = cnt;
            if (s < tv.nslices) {
                mbar_expect_tx(&bar, bytes); bulk_g2s(piece, tv.T + (s << tv.slice_bits), bytes, &bar);
                // the slice's tuples start their way into L2 now (one instruction for the whole range)
                const unsigned long long a0 = (unsigned long long)(tuples + (fstart ? fstart[s] : s * cap) * TW);
                const unsigned long long a = a0 & ~15ULL, e = (a0 + sh_n * TW * 8) & ~15ULL;
                if (e > a) bulk_prefetch_l2(reinterpret_cast<const void *>(a), (unsigned)(e - a));
            }
        }
        __syncthreads();
        const unsigned long long s = sh_ticket;
        if (s >= tv.nslices) break;
        const uint64_t *tp = tuples + (fstart ? fstart[s] : s * cap) * TW;   // this slice's tuples
        const unsigned n = (unsigned)sh_n;                                 // (a batch brings fewer than 2^32 tuples)
        // this warp's share [lo, hi): whole groups of 32 tuples
        const unsigned per = ((n + NT / 32 - 1) / (NT / 32) + 31u) & ~31u;
        const unsigned lo = min(n, warp * per), hi = min(n, lo + per);
        unsigned head = lo + 64;                                           // first tuple of the share nobody holds yet
        ulonglong2 cur = lo + lane < hi ? t_raw<TW>(tp, lo + lane) : make_ulonglong2(T_PAD1, T_PAD1);
        bool more = lo + 32 + lane < hi;                                   // nxt holds an element of the stream
        ulonglong2 nxt = t_raw<TW>(tp, min(lo + 32 + lane, n - 1));
        mbar_wait(&bar, parity); parity ^= 1;
        // state of the tuple in hand
        uint64_t ck; uint32_t ce; t_split<TW>(cur, ck, ce);
        unsigned long long ci = inc_lut[ce];
        uint32_t ca = piece_a + (((unsigned)home_hash(ck) & mask) << 4);   // shared address of the slot to look at
        unsigned probes = 0;
        for (;;) {
            const bool live = ce != 255u;
            if (!__any_sync(0xffffffffu, live || more)) break;
            bool done = !live && more;                                     // padding inside the stream: move on
            bool clash = false;
            if (live) {
                const unsigned long long tag = ck + 1;
                unsigned long long t, c;                                   // one 16-byte shared load: {tag, cnt}
                asm volatile("ld.volatile.shared.v2.u64 {%0, %1}, [%2];" : "=l"(t), "=l"(c) : "r"(ca) : "memory");
                if (t == 0) {
                    t = atom_cas_shared(ca, 0ULL, tag);
                    if (t == 0) { t = tag; ++fresh; c = 0; }
                }
                if (t == tag) {
                    if (cnt_cov(c) + 1u > KQ_LANE_MAX) { smem_spill(&piece[(ca - piece_a) >> 4].cnt, tv, ck, ci); done = true; }
                    else if (atom_cas_shared(ca + 8, c, c + ci) == c) done = true;
                    else clash = true;                                     // somebody else counted into the slot: look again
                } else {
                    ca = piece_a + ((ca - piece_a + 16u) & (bytes - 1));
                    if (++probes >= limit) { smem_give_back<TW>(ovf, ovf_cap, ovf_n, ck, ci); done = true; }
                }
            }
            if (__popc(__ballot_sync(0xffffffffu, clash)) >= 4) {
                // neighbours in the stream with one key (poly-A, satellites): combine inside the warp
                const bool todo = live && !done;
                const unsigned long long inc = warp_combine(todo ? ck : ~0ULL - lane, ce, todo);   // distinct dummies: no false match
                if (inc) fresh += smem_insert_slow<TW>(piece, mask, limit, tv, ck, inc, ovf, ovf_cap, ovf_n);
                done |= todo;
            }
            const unsigned dm = __ballot_sync(0xffffffffu, done);
            if (done) {
                cur = more ? nxt : make_ulonglong2(T_PAD1, T_PAD1);
                const unsigned i = head + __popc(dm & lt_mask);
                more = i < hi;
                nxt = t_raw<TW>(tp, min(i, n - 1));                        // (always a valid address; used only if `more`)
                t_split<TW>(cur, ck, ce);
                ci = inc_lut[ce];
                ca = piece_a + (((unsigned)home_hash(ck) & mask) << 4);
                probes = 0;
            }
            head += __popc(dm);
        }
        __syncthreads();                                                  // every update is in the piece
        if (threadIdx.x == 0) {
            fence_async_smem();
            bulk_s2g(tv.T + (s << tv.slice_bits), piece, bytes);
            bulk_commit();
            bulk_wait_read();                                             // the piece may be overwritten again
        }
    }
    if (threadIdx.x == 0) bulk_wait_all();
    // new keys of this CTA -> occupancy counter
    unsigned long long f = fresh;
    for (int o = 16; o; o >>= 1) f += __shfl_xor_sync(0xffffffffu, f, o);
    if (lane == 0 && f) atomicAdd(&sh_fresh, f);
    __syncthreads();
    if (threadIdx.x == 0 && sh_fresh) atomicAdd(&tv.occ[0], sh_fresh);
}

// ---- warp-owned slices -------------------------------------------------------------------------------------------
//
// Shared-memory CAS costs about 4 cycles per LANE on B200 (profiles/r02_ncu_pipeline_full.md: the CTA-owned kernel
// above sits at that bound whatever its instruction count).  With slices of 2^9 slots (8 KiB) ONE WARP owns a slice
// while it counts into it, so no other warp can touch the slots and plain 16-byte shared loads / stores do the work:
//   every iteration each lane reads the slot it looks at (LDS.128); the lanes that want to write it -- empty slot to
//   claim, or their key -- put their lane number into the slot's LOCK BYTE, the warp synchronises, and the lane that
//   reads its own number back owns the slot for this iteration: it stores {tag, cnt + inc} (STS.128).  The others
//   look again in the next iteration.  All reads of an iteration come before all of its writes.
// A warp takes a ticket for 32 consecutive slices, skips the empty ones, and for each: bulk-loads the 8 KiB piece
// (cp.async.bulk + its own mbarrier), pops the slice's tuples as above, bulk-stores the piece.  Warps never meet.
__device__ __noinline__ unsigned long long excl_spill(const TableView &tv, uint64_t key, unsigned long long c) {
    uint32_t fw[4], bw[4];
    unsigned long long take = 63ull << 56;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
        fw[w] = min(cnt_lane(c, w), 63u); bw[w] = min(cnt_lane(c, 4 + w), 63u);
        take |= (unsigned long long)fw[w] << (7 * w) | (unsigned long long)bw[w] << (7 * (4 + w));
    }
    remainder_add(tv, key, 63, fw, bw);
    return c - take;
}

// Lanes that wait for one slot with the same key add their increments up: the lowest lane of the group takes the
// others in lane order while the sum stays at or below 63 instances (every field of the sum then fits its 7 bits:
// a lane counts at most as many edges as instances; and a slot that cannot take 63 more has at least 65, so the
// spill of 63 never underflows).  Returns the lane's new increment, 0 = another lane counts for this one.
// All 32 lanes call it; lanes without work pass todo = false.
constexpr unsigned KQ_MERGE_MAX = 63;
__device__ __noinline__ unsigned long long warp_merge(uint64_t key, bool todo, unsigned long long ci) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned act = __ballot_sync(0xffffffffu, todo);
    const unsigned peers = __match_any_sync(0xffffffffu, todo ? key : ~0ULL - lane) & act;   // distinct dummies: no false match
    unsigned long long acc = 0; unsigned acc_n = 0; bool mine = false;
    for (unsigned m = act; m; m &= m - 1) {
        const int src = __ffs(m) - 1;
        const unsigned long long v = __shfl_sync(0xffffffffu, ci, src);
        if ((peers >> src) & 1u) {
            const unsigned n = (unsigned)(v >> 56);
            if (acc_n == 0 || acc_n + n <= KQ_MERGE_MAX) { acc += v; acc_n += n; mine |= src == (int)lane; }
        }
    }
    if (!todo) return ci;
    if (lane == (unsigned)__ffs(peers) - 1) return acc;
    return mine ? 0ULL : ci;
}

// complete insert of one (merged) increment by ONE lane while the other lanes of the warp wait: plain loads and
// stores on the warp's slice; 1 = new key
template <int TW>
__device__ __noinline__ unsigned excl_insert(uint32_t piece_a, unsigned mask, unsigned limit, const TableView &tv, uint64_t key,
                                             unsigned long long inc, uint64_t *ovf, unsigned long long ovf_cap, unsigned long long *ovf_n) {
    const unsigned long long tag = key + 1;
    unsigned slot = (unsigned)home_hash(key) & mask;
    for (unsigned probes = 0; probes < limit; ++probes, slot = (slot + 1) & mask) {
        unsigned long long t, c;
        asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(t), "=l"(c) : "r"(piece_a + (slot << 4)) : "memory");
        if (t != 0 && t != tag) continue;
        if (t == 0) c = 0;
        if (cnt_cov(c) + (unsigned)(inc >> 56) > KQ_LANE_MAX) c = excl_spill(tv, key, c);
        c += inc;
        asm volatile("st.shared.v2.u64 [%0], {%1, %2};" ::"r"(piece_a + (slot << 4)), "l"(tag), "l"(c) : "memory");
        return t == 0;
    }
    smem_give_back<TW>(ovf, ovf_cap, ovf_n, key, inc);
    return 0;
}

template <int TW>
__global__ void __launch_bounds__(256, 3)
warp_insert_kernel(TableView tv, const uint64_t *__restrict__ tuples, const unsigned long long *__restrict__ fstart,
                   const void *__restrict__ fcount_, unsigned long long cap, int stride, uint64_t *ovf,
                   unsigned long long ovf_cap, unsigned long long *ovf_n, unsigned long long *ticket, HotList hot) {
    const unsigned long long *fcount = static_cast<const unsigned long long *>(fcount_);     // level-2 layout: exact counts
    const unsigned *fcur = static_cast<const unsigned *>(fcount_);                             // one-pass layout: 32-bit cursors
    extern __shared__ __align__(128) unsigned char sm_w[];
    __shared__ unsigned long long sh_fresh;
    const unsigned nw = blockDim.x >> 5, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, lt_mask = (1u << lane) - 1;
    const unsigned slots = 1u << tv.slice_bits, mask = slots - 1, bytes = slots * (unsigned)sizeof(Slot);
    const unsigned limit = slots;                                       // once around: the slice is full
    // [nw pieces][nw x slots lock bytes][increment table][nw mbarriers]
    unsigned char *piece = sm_w + (size_t)warp * bytes;
    const uint32_t piece_a = smem_u32(piece), lock_a = smem_u32(sm_w + (size_t)nw * bytes + (size_t)warp * slots);
    unsigned long long *inc_lut = reinterpret_cast<unsigned long long *>(sm_w + (size_t)nw * (bytes + slots));
    uint64_t *bar = reinterpret_cast<uint64_t *>(sm_w + (size_t)nw * (bytes + slots) + 2048) + warp;
    for (unsigned i = threadIdx.x; i < 256; i += blockDim.x) inc_lut[i] = inc_of_edge(i);
    if (lane == 0) mbar_init(bar, 1);
    if (threadIdx.x == 0) sh_fresh = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const uint32_t lut_a = smem_u32(inc_lut);
    unsigned parity = 0, fresh = 0;
    for (;;) {
        unsigned long long g = 0;
        if (lane == 0) g = atomicAdd(ticket, 1ULL);
        g = __shfl_sync(0xffffffffu, g, 0);
        const unsigned long long s0 = g * 32;
        if (s0 >= tv.nslices) break;
        unsigned long long my_n = 0, my_start = 0;
        if (s0 + lane < tv.nslices) {
            const unsigned long long sl = s0 + lane;
            my_n = fstart ? fcount[sl] : min((unsigned long long)fcur[sl * stride], cap);
            my_start = fstart ? fstart[sl] : sl * cap;
        }
        unsigned todo_slices = __ballot_sync(0xffffffffu, my_n != 0);
        while (todo_slices) {
            const int src = __ffs(todo_slices) - 1; todo_slices &= todo_slices - 1;
            const unsigned n = (unsigned)__shfl_sync(0xffffffffu, my_n, src);      // (a batch brings fewer than 2^32 tuples)
            const unsigned long long start = __shfl_sync(0xffffffffu, my_start, src);
            if (hot.buf && n > hot.limit) {
                // a slice with many times its share of the batch holds a run of one key (poly-A, a satellite): one
                // warp would work on it long after the others are done -- hot_slices_kernel counts it with all SMs
                unsigned long long j = 0;
                if (lane == 0) j = atomicAdd(&hot.buf[0], 1ULL);
                j = __shfl_sync(0xffffffffu, j, 0);
                if (j < hot.cap) {
                    if (lane == 0) { hot.buf[1 + 3 * j] = start; hot.buf[2 + 3 * j] = n; hot.buf[3 + 3 * j] = s0 + src; }
                    continue;
                }
            }
            const uint64_t *tp = tuples + start * TW;
            Slot *gpiece = tv.T + ((s0 + src) << tv.slice_bits);
            if (lane == 0) { mbar_expect_tx(bar, bytes); bulk_g2s(piece, gpiece, bytes, bar); }
            unsigned head = 64;                                            // first tuple of the slice nobody holds yet
            ulonglong2 cur = lane < n ? t_raw<TW>(tp, lane) : make_ulonglong2(T_PAD1, T_PAD1);
            bool more = 32 + lane < n;                                     // nxt holds an element of the stream
            ulonglong2 nxt = t_raw<TW>(tp, min(32 + lane, n - 1));
            uint64_t ck; uint32_t ce; t_split<TW>(cur, ck, ce);
            mbar_wait(bar, parity); parity ^= 1;
            unsigned long long ci;
            asm volatile("ld.shared.u64 %0, [%1];" : "=l"(ci) : "r"(lut_a + ce * 8));
            unsigned slot = (unsigned)home_hash(ck) & mask, probes = 0;
            for (;;) {
                const bool live = ce != 255u;
                if (!__any_sync(0xffffffffu, live || more)) break;
                bool done = !live && more;                                 // padding inside the stream: move on
                bool want = false;
                const unsigned long long tag = ck + 1;
                unsigned long long t = 0, c = 0;
                if (live) {
                    asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(t), "=l"(c) : "r"(piece_a + (slot << 4)) : "memory");
                    if (t == 0 || t == tag) {
                        want = true;
                        asm volatile("st.shared.u8 [%0], %1;" ::"r"(lock_a + slot), "r"(lane) : "memory");
                    } else {
                        slot = (slot + 1) & mask;
                        if (++probes >= limit) { smem_give_back<TW>(ovf, ovf_cap, ovf_n, ck, ci); done = true; }
                    }
                }
                __syncwarp();
                bool lost = false;
                if (want) {
                    unsigned who;
                    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(who) : "r"(lock_a + slot) : "memory");
                    if (who == lane) {
                        if (t == 0) { c = 0; ++fresh; }
                        if (cnt_cov(c) + (unsigned)(ci >> 56) > KQ_LANE_MAX) c = excl_spill(tv, ck, c);
                        c += ci;                                           // lanes <= cov: no carry between lanes
                        asm volatile("st.shared.v2.u64 [%0], {%1, %2};" ::"r"(piece_a + (slot << 4)), "l"(tag), "l"(c) : "memory");
                        done = true;
                    } else lost = true;
                }
                __syncwarp();
                if (__popc(__ballot_sync(0xffffffffu, lost)) >= 12) {
                    // neighbours in the stream with one key (poly-A, satellites): they add up, one lane counts for all
                    // and that lane counts at once, before new tuples of the run arrive (the lock gives no lane priority:
                    // a lane that only waits with its sum would starve behind the single tuples that keep coming)
                    const bool todo = live && !done;
                    const unsigned long long inc = warp_merge(ck, todo, ci);
                    for (unsigned m = __ballot_sync(0xffffffffu, todo && inc != 0); m; m &= m - 1) {
                        if (lane == (unsigned)__ffs(m) - 1) fresh += excl_insert<TW>(piece_a, mask, limit, tv, ck, inc, ovf, ovf_cap, ovf_n);
                        __syncwarp();
                    }
                    done |= todo;
                }
                const unsigned dm = __ballot_sync(0xffffffffu, done);
                if (done) {
                    cur = more ? nxt : make_ulonglong2(T_PAD1, T_PAD1);
                    const unsigned i = head + __popc(dm & lt_mask);
                    more = i < n;
                    nxt = t_raw<TW>(tp, min(i, n - 1));                    // (always a valid address; used only if `more`)
                    t_split<TW>(cur, ck, ce);
                    asm volatile("ld.shared.u64 %0, [%1];" : "=l"(ci) : "r"(lut_a + ce * 8));
                    slot = (unsigned)home_hash(ck) & mask;
                    probes = 0;
                }
                head += __popc(dm);
            }
            fence_async_smem();                                            // every lane's stores, then one bulk store
            __syncwarp();
            if (lane == 0) { bulk_s2g(gpiece, piece, bytes); bulk_commit(); bulk_wait_read(); }
            __syncwarp();                                                  // the piece may be overwritten again
        }
    }
    if (lane == 0) bulk_wait_all();
    unsigned long long f = fresh;
    for (int o = 16; o; o >>= 1) f += __shfl_xor_sync(0xffffffffu, f, o);
    if (lane == 0 && f) atomicAdd(&sh_fresh, f);
    __syncthreads();
    if (threadIdx.x == 0 && sh_fresh) atomicAdd(&tv.occ[0], sh_fresh);
}

// ---- CTA-owned slices without atomics ---------------------------------------------------------------------------------
//
// The lock-byte protocol of warp_insert_kernel for slices of up to 2^12 slots (64 KiB), which one CTA holds: the eight
// warps run their iterations in step -- every thread reads the slot it looks at, the ones that want to write it put
// their thread number into the slot's lock byte, BARRIER, the thread that reads its own number back stores
// {tag, cnt + inc}, BARRIER (which also tells whether anybody has work left).  Measured on C2: 89.8 ms against 34.6 ms for
// shared-memory CAS (smem_insert_kernel): two CTA barriers per probe iteration cost far more than the atomics they
// replace, three CTAs per SM do not fill the waits.  Kept as a knob (cta_lock) and as a cross-check of the other kernels.  Runs of one key: a warp in which >= 12 lanes lost the lock adds the increments of equal keys up
// (warp_merge); the lanes that hold the sums count them with shared-memory CAS in a phase of their own, in which nobody
// stores without CAS.
template <int TW>
__global__ void __launch_bounds__(256, 3)
cta_insert_kernel(TableView tv, const uint64_t *__restrict__ tuples, const unsigned long long *__restrict__ fstart,
                  const void *__restrict__ fcount_, unsigned long long cap, int stride, uint64_t *ovf,
                  unsigned long long ovf_cap, unsigned long long *ovf_n, unsigned long long *ticket, HotList hot) {
    const unsigned long long *fcount = static_cast<const unsigned long long *>(fcount_);     // level-2 layout: exact counts
    const unsigned *fcur = static_cast<const unsigned *>(fcount_);                             // one-pass layout: 32-bit cursors
    extern __shared__ __align__(128) unsigned char sm_c[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ unsigned long long sh_ticket, sh_fresh, sh_n, sh_start;
    __shared__ unsigned sh_hot;
    __shared__ unsigned long long inc_lut[256];
    Slot *piece = reinterpret_cast<Slot *>(sm_c);
    const unsigned slots = 1u << tv.slice_bits, mask = slots - 1, bytes = slots * (unsigned)sizeof(Slot);
    const unsigned limit = min(slots, 1024u);                          // a longer run means the slice is (locally) full
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, lt_mask = (1u << lane) - 1;
    const uint32_t piece_a = smem_u32(piece), lock_a = smem_u32(sm_c + bytes), lut_a = smem_u32(inc_lut);
    for (unsigned i = threadIdx.x; i < 256; i += 256) inc_lut[i] = inc_of_edge(i);
    if (threadIdx.x == 0) { mbar_init(&bar, 1); sh_fresh = 0; sh_hot = 0; asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    unsigned parity = 0, fresh = 0;
    for (;;) {
        if (threadIdx.x == 0) {
            unsigned long long s, cnt = 0;
            for (;;) {
                s = atomicAdd(ticket, 1ULL); if (s >= tv.nslices) break;
                cnt = fstart ? fcount[s] : min((unsigned long long)fcur[s * stride], cap);
                if (!cnt) continue;
                if (hot.buf && cnt > hot.limit) {                       // a run of one key: left to hot_slices_kernel (all SMs)
                    const unsigned long long j = atomicAdd(&hot.buf[0], 1ULL);
                    if (j < hot.cap) { hot.buf[1 + 3 * j] = fstart ? fstart[s] : s * cap; hot.buf[2 + 3 * j] = cnt; hot.buf[3 + 3 * j] = s; continue; }
                }
                break;
            }
            sh_ticket = s; sh_n = cnt;
            if (s < tv.nslices) {
                sh_start = fstart ? fstart[s] : s * cap;
                mbar_expect_tx(&bar, bytes); bulk_g2s(piece, tv.T + (s << tv.slice_bits), bytes, &bar);
                const unsigned long long a0 = (unsigned long long)(tuples + sh_start * TW);
                const unsigned long long a = a0 & ~15ULL, e = (a0 + cnt * TW * 8) & ~15ULL;
                if (e > a) bulk_prefetch_l2(reinterpret_cast<const void *>(a), (unsigned)(e - a));
            }
        }
        __syncthreads();
        const unsigned long long s = sh_ticket;
        if (s >= tv.nslices) break;
        const uint64_t *tp = tuples + sh_start * TW;
        const unsigned n = (unsigned)sh_n;
        const unsigned per = ((n + 7) / 8 + 31u) & ~31u;                   // this warp's share [lo, hi): whole groups of 32
        const unsigned lo = min(n, warp * per), hi = min(n, lo + per);
        unsigned head = lo + 64;
        ulonglong2 cur = lo + lane < hi ? t_raw<TW>(tp, lo + lane) : make_ulonglong2(T_PAD1, T_PAD1);
        bool more = lo + 32 + lane < hi;
        ulonglong2 nxt = t_raw<TW>(tp, min(lo + 32 + lane, n - 1));
        uint64_t ck; uint32_t ce; t_split<TW>(cur, ck, ce);
        mbar_wait(&bar, parity); parity ^= 1;
        unsigned long long ci;
        asm volatile("ld.shared.u64 %0, [%1];" : "=l"(ci) : "r"(lut_a + ce * 8));
        unsigned slot = (unsigned)home_hash(ck) & mask, probes = 0;
        // pop the next tuple of the warp's share for the lanes that are done
        auto advance = [&](bool done) {
            const unsigned dm = __ballot_sync(0xffffffffu, done);
            if (done) {
                cur = more ? nxt : make_ulonglong2(T_PAD1, T_PAD1);
                const unsigned i = head + __popc(dm & lt_mask);
                more = i < hi;
                nxt = t_raw<TW>(tp, min(i, n - 1));
                t_split<TW>(cur, ck, ce);
                asm volatile("ld.shared.u64 %0, [%1];" : "=l"(ci) : "r"(lut_a + ce * 8));
                slot = (unsigned)home_hash(ck) & mask;
                probes = 0;
            }
            head += __popc(dm);
        };
        for (;;) {
            const bool live = ce != 255u;
            bool done = !live && more;                                     // padding inside the stream: move on
            bool want = false;
            const unsigned long long tag = ck + 1;
            unsigned long long t = 0, c = 0;
            if (live) {
                asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(t), "=l"(c) : "r"(piece_a + (slot << 4)) : "memory");
                if (t == 0 || t == tag) {
                    want = true;
                    asm volatile("st.shared.u8 [%0], %1;" ::"r"(lock_a + slot), "r"(threadIdx.x) : "memory");
                } else {
                    slot = (slot + 1) & mask;
                    if (++probes >= limit) { smem_give_back<TW>(ovf, ovf_cap, ovf_n, ck, ci); done = true; }
                }
            }
            __syncthreads();
            bool lost = false;
            if (want) {
                unsigned who;
                asm volatile("ld.shared.u8 %0, [%1];" : "=r"(who) : "r"(lock_a + slot) : "memory");
                if (who == threadIdx.x) {
                    if (t == 0) { c = 0; ++fresh; }
                    if (cnt_cov(c) + (unsigned)(ci >> 56) > KQ_LANE_MAX) c = excl_spill(tv, ck, c);
                    c += ci;                                               // lanes <= cov: no carry between lanes
                    asm volatile("st.shared.v2.u64 [%0], {%1, %2};" ::"r"(piece_a + (slot << 4)), "l"(tag), "l"(c) : "memory");
                    done = true;
                } else lost = true;
            }
            bool lead = false;
            if (__popc(__ballot_sync(0xffffffffu, lost)) >= 12) {
                const bool todo = live && !done;
                const unsigned long long inc = warp_merge(ck, todo, ci);
                if (todo) { if (inc) { ci = inc; lead = true; sh_hot = 1; } else done = true; }
            }
            advance(done);
            if (!__syncthreads_or((ce != 255u) || more)) break;          // (also: every store of this iteration is done)
            if (sh_hot) {
                // the sums are counted with CAS while nobody stores plainly
                if (lead) fresh += smem_insert_slow<TW>(piece, mask, limit, tv, ck, ci, ovf, ovf_cap, ovf_n);
                advance(lead);
                __syncthreads();
                if (threadIdx.x == 0) sh_hot = 0;                          // (set again only after the next barrier)
            }
        }
        if (threadIdx.x == 0) {
            fence_async_smem();
            bulk_s2g(tv.T + (s << tv.slice_bits), piece, bytes);
            bulk_commit();
            bulk_wait_read();                                             // the piece may be overwritten again
        }
    }
    if (threadIdx.x == 0) bulk_wait_all();
    unsigned long long f = fresh;
    for (int o = 16; o; o >>= 1) f += __shfl_xor_sync(0xffffffffu, f, o);
    if (lane == 0 && f) atomicAdd(&sh_fresh, f);
    __syncthreads();
    if (threadIdx.x == 0 && sh_fresh) atomicAdd(&tv.occ[0], sh_fresh);
}

// The slices warp_insert_kernel left out: every warp of the grid takes 32 tuples at a time, the lanes with one key add
// up (warp_combine) and one lane per key counts the sum into the slice in HBM (CAS through L2).
template <int TW>
__global__ void __launch_bounds__(256)
hot_slices_kernel(TableView tv, const uint64_t *__restrict__ tuples, HotList hot, uint64_t *ovf, unsigned long long ovf_cap,
                  unsigned long long *ovf_n) {
    __shared__ unsigned long long sh_fresh;
    if (threadIdx.x == 0) sh_fresh = 0;
    __syncthreads();
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const unsigned slots = 1u << tv.slice_bits, mask = slots - 1;
    const unsigned long long nh = min(hot.buf[0], (unsigned long long)hot.cap);
    unsigned fresh = 0;
    for (unsigned long long h = 0; h < nh; ++h) {
        const unsigned long long start = hot.buf[1 + 3 * h], n = hot.buf[2 + 3 * h], s = hot.buf[3 + 3 * h];
        Slot *piece = tv.T + (s << tv.slice_bits);
        const uint64_t *tp = tuples + start * TW;
        for (unsigned long long i0 = ((unsigned long long)blockIdx.x * nw + warp) * 32; i0 < n; i0 += (unsigned long long)gridDim.x * nw * 32) {
            const unsigned long long i = i0 + lane;
            uint64_t key = 0; uint32_t e = 255u;
            if (i < n) t_unpack<TW>(tp, i, key, e);
            const bool live = e != 255u;
            const unsigned long long inc = warp_combine(live ? key : ~0ULL - lane, e, live);
            if (inc) fresh += smem_insert_slow<TW>(piece, mask, slots, tv, key, inc, ovf, ovf_cap, ovf_n);
        }
    }
    unsigned long long f = fresh;
    for (int o = 16; o; o >>= 1) f += __shfl_xor_sync(0xffffffffu, f, o);
    if (lane == 0 && f) atomicAdd(&sh_fresh, f);
    __syncthreads();
    if (threadIdx.x == 0 && sh_fresh) atomicAdd(&tv.occ[0], sh_fresh);
}

// ---- launchers -------------------------------------------------------------------------------------------------------

#define KQ_TW_DISPATCH3(tw, CALL) do { if ((tw) == 1) { constexpr int TW = 1; CALL; } else { constexpr int TW = 2; CALL; } } while (0)

static size_t fine_scatter_smem(unsigned ns_max, unsigned cap, int tw, bool priv = true) {
    return (size_t)ns_max * (cap + cap / 2) * tw * 8 + (size_t)ns_max * (priv ? 16 : 8);
}

// flush chunk (tuples) of the level-2 scatter for bins of `ns_max` slices: the largest of 64 / 32 / 16 / 8 that fits
// the shared memory of `ctas_per_sm` CTAs per SM (0: none does)
unsigned fine_pick_cap(unsigned ns_max, int tw, int ctas_per_sm) {
    static int max_smem = 0;
    if (!max_smem) {
        int dev = 0; cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        if (max_smem <= 0) max_smem = 227 * 1024;
    }
    const size_t budget = ((size_t)max_smem + 1024) / (ctas_per_sm > 1 ? ctas_per_sm : 1) - 2048;   // 1 KiB per CTA is the system's
    for (unsigned cap : {64u, 32u, 16u, 8u})
        if (fine_scatter_smem(ns_max, cap, tw) <= budget) return cap;
    return 0;
}

cudaError_t launch_fine_plan(cudaStream_t st, const TableView &tv, unsigned ncoarse, const unsigned long long *start,
                             const unsigned long long *n, unsigned long long part, FineTask *tasks, unsigned task_cap,
                             unsigned *ntasks, unsigned *task_first, uint64_t *launches) {
    fine_plan_kernel<<<1, 1024, 0, st>>>(tv, ncoarse, start, n, part, tasks, task_cap, ntasks, task_first);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_route_plan(cudaStream_t st, const TableView &tv, const RouteMap &rt, int me, int cap, const unsigned long long *H,
                              unsigned long long part, FineTask *tasks, unsigned task_cap, unsigned *ntasks, unsigned *task_first,
                              uint64_t *launches) {
    route_plan_kernel<<<1, 1024, 0, st>>>(tv, rt.nranks, rt.nb_bits, me, cap, H, part, tasks, task_cap, ntasks, task_first);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_route_fine_offsets(cudaStream_t st, const TableView &tv, int nb_bits, const FineTask *tasks, const unsigned *task_first,
                                      unsigned *hist2, unsigned ns_max, unsigned long long *fcount, uint64_t *launches) {
    const unsigned grid = (unsigned)std::min<unsigned long long>((tv.nslices + 255) / 256, (unsigned long long)sm_count() * 8);
    route_fine_offsets_kernel<<<grid, 256, 0, st>>>(tv, nb_bits, tasks, task_first, hist2, ns_max, fcount);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_fine_hist(cudaStream_t st, const TableView &tv, const uint64_t *in, const FineTask *tasks,
                             const unsigned *ntasks, unsigned task_bound, unsigned ns_max, unsigned long long *fhist,
                             unsigned *hist2, uint64_t *launches) {
    if (!task_bound) return cudaSuccess;
    const unsigned grid = std::min<unsigned>(task_bound, (unsigned)sm_count() * 2);
    KQ_TW_DISPATCH3(tuple_words(tv.k), (fine_hist_kernel<TW><<<grid, FH_NT, (size_t)ns_max * 4, st>>>(tv, in, tasks, ntasks, fhist, hist2, ns_max)));
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_fine_offsets(cudaStream_t st, const TableView &tv, unsigned ncoarse, const unsigned long long *start,
                                const unsigned *task_first, unsigned *hist2, unsigned ns_max, unsigned long long *fstart,
                                unsigned long long *fcount, uint64_t *launches) {
    fine_offsets_kernel<<<ncoarse, 1024, (size_t)ns_max * 8, st>>>(tv, start, task_first, hist2, ns_max, fstart, fcount);
    ++*launches;
    return cudaGetLastError();
}

// hist2 / fstart != NULL: private cursors (local pipeline); else one global atomic per flush on `cursor`
cudaError_t launch_fine_scatter(cudaStream_t st, const TableView &tv, const uint64_t *in, const FineTask *tasks,
                                const unsigned *ntasks, unsigned task_bound, unsigned ns_max, uint64_t *out,
                                unsigned long long *cursor, const unsigned *hist2, const unsigned long long *fstart, uint64_t *launches) {
    if (!task_bound) return cudaSuccess;
    const int tw = tuple_words(tv.k);
    const bool priv = hist2 != nullptr;
    // two CTAs of 512 threads per SM (each other's barriers and flushes overlap) when bins of at least 16 tuples fit
    // half the shared memory, else one CTA of 1024 threads with all of it
    const int variant = tuning().fs_variant;
    int ctas = variant == 1 ? 1 : 2;
    unsigned cap = fine_pick_cap(ns_max, tw, ctas);
    if (ctas == 2 && cap < 16 && variant != 2) { ctas = 1; cap = fine_pick_cap(ns_max, tw, 1); }
    if (!cap) return cudaErrorInvalidConfiguration;
    const size_t smem = fine_scatter_smem(ns_max, cap, tw, priv);
    const unsigned grid = std::min<unsigned>(task_bound, (unsigned)sm_count() * ctas);
    cudaError_t e = cudaSuccess;
#define KQ_FS_GO(TWV, NTV, RV, CAPV, PV) do {                                                                          \
        e = cudaFuncSetAttribute(fine_scatter_kernel<TWV, NTV, RV, CAPV, PV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                                \
        cudaFuncSetAttribute(fine_scatter_kernel<TWV, NTV, RV, CAPV, PV>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared); \
        fine_scatter_kernel<TWV, NTV, RV, CAPV, PV><<<grid, NTV, smem, st>>>(tv, in, tasks, ntasks, out, cursor, hist2, fstart, ns_max); } while (0)
#define KQ_FS_CAP(TWV, NTV, RV, PV) do {                                                                               \
        if (cap == 64) KQ_FS_GO(TWV, NTV, RV, 64, PV); else if (cap == 32) KQ_FS_GO(TWV, NTV, RV, 32, PV);             \
        else if (cap == 16) KQ_FS_GO(TWV, NTV, RV, 16, PV); else KQ_FS_GO(TWV, NTV, RV, 8, PV); } while (0)
#define KQ_FS_SHAPE(TWV, PV) do { if (ctas == 2) KQ_FS_CAP(TWV, 512, 4, PV); else KQ_FS_CAP(TWV, 1024, 2, PV); } while (0)
    if (tw == 1) { if (priv) KQ_FS_SHAPE(1, true); else KQ_FS_SHAPE(1, false); }
    else { if (priv) KQ_FS_SHAPE(2, true); else KQ_FS_SHAPE(2, false); }
#undef KQ_FS_SHAPE
#undef KQ_FS_CAP
#undef KQ_FS_GO
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_smem_insert(cudaStream_t st, const TableView &tv, const uint64_t *tuples, const unsigned long long *fstart,
                               const void *fcount, unsigned long long cap, int stride, uint64_t *ovf,
                               unsigned long long ovf_cap, unsigned long long *ovf_n, unsigned long long *ticket, uint64_t *launches,
                               HotList hot) {
    cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    if (tuning().warp_insert && tv.slice_bits <= 10) {
        // warp-owned slices: 8 warps per CTA, as many CTAs per SM as fit (3 with 8 KiB slices)
        const unsigned slots = 1u << tv.slice_bits, nw = 8;
        const size_t smem_w = (size_t)nw * (slots * sizeof(Slot) + slots) + 2048 + nw * 8;
        const unsigned per_sm = (unsigned)std::max<size_t>(1, std::min<size_t>(3, (size_t)(227 * 1024) / (smem_w + 1024)));
        const unsigned long long groups = (tv.nslices + 31) / 32;
        const unsigned grid_w = (unsigned)std::min<unsigned long long>((groups + nw - 1) / nw, (unsigned long long)sm_count() * per_sm);
#define KQ_WI_LAUNCH(TWV) do {                                                                                         \
        e = cudaFuncSetAttribute(warp_insert_kernel<TWV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_w);   \
        if (e != cudaSuccess) return e;                                                                                \
        cudaFuncSetAttribute(warp_insert_kernel<TWV>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared); \
        warp_insert_kernel<TWV><<<grid_w, nw * 32, smem_w, st>>>(tv, tuples, fstart, fcount, cap, stride, ovf, ovf_cap, ovf_n, ticket, hot); \
        if (hot.buf) hot_slices_kernel<TWV><<<(unsigned)sm_count() * 4, 256, 0, st>>>(tv, tuples, hot, ovf, ovf_cap, ovf_n); } while (0)
        if (hot.buf) { e = cudaMemsetAsync(hot.buf, 0, sizeof(unsigned long long), st); if (e != cudaSuccess) return e; }
        if (tuple_words(tv.k) == 1) KQ_WI_LAUNCH(1); else KQ_WI_LAUNCH(2);
#undef KQ_WI_LAUNCH
        *launches += hot.buf ? 2 : 1;
        return cudaGetLastError();
    }
    if (tuning().cta_lock && tv.slice_bits <= 12) {
        const size_t smem_c = ((size_t)sizeof(Slot) << tv.slice_bits) + ((size_t)1 << tv.slice_bits);
        const unsigned grid_c = (unsigned)std::min<unsigned long long>(tv.nslices, (unsigned long long)sm_count() * 3);
        if (hot.buf) { e = cudaMemsetAsync(hot.buf, 0, sizeof(unsigned long long), st); if (e != cudaSuccess) return e; }
#define KQ_CI_LAUNCH(TWV) do {                                                                                         \
        e = cudaFuncSetAttribute(cta_insert_kernel<TWV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_c);    \
        if (e != cudaSuccess) return e;                                                                                \
        cudaFuncSetAttribute(cta_insert_kernel<TWV>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared); \
        cta_insert_kernel<TWV><<<grid_c, 256, smem_c, st>>>(tv, tuples, fstart, fcount, cap, stride, ovf, ovf_cap, ovf_n, ticket, hot); \
        if (hot.buf) hot_slices_kernel<TWV><<<(unsigned)sm_count() * 4, 256, 0, st>>>(tv, tuples, hot, ovf, ovf_cap, ovf_n); } while (0)
        if (tuple_words(tv.k) == 1) KQ_CI_LAUNCH(1); else KQ_CI_LAUNCH(2);
#undef KQ_CI_LAUNCH
        *launches += hot.buf ? 2 : 1;
        return cudaGetLastError();
    }
    const size_t smem = sizeof(Slot) << tv.slice_bits;
    const unsigned grid = (unsigned)std::min<unsigned long long>(tv.nslices, (unsigned long long)sm_count() * (unsigned)tuning().s3_ctas);
#define KQ_SI_LAUNCH(TWV, NTV) do {                                                                                    \
        e = cudaFuncSetAttribute(smem_insert_kernel<TWV, NTV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                                \
        cudaFuncSetAttribute(smem_insert_kernel<TWV, NTV>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared); \
        smem_insert_kernel<TWV, NTV><<<grid, NTV, smem, st>>>(tv, tuples, fstart, fcount, cap, stride, ovf, ovf_cap, ovf_n, ticket, hot); } while (0)
    if (hot.buf) { e = cudaMemsetAsync(hot.buf, 0, sizeof(unsigned long long), st); if (e != cudaSuccess) return e; }
    const int nt = tuning().s3_threads;
    if (tuple_words(tv.k) == 1) { if (nt == 256) KQ_SI_LAUNCH(1, 256); else if (nt == 384) KQ_SI_LAUNCH(1, 384); else KQ_SI_LAUNCH(1, 512); }
    else { if (nt == 256) KQ_SI_LAUNCH(2, 256); else if (nt == 384) KQ_SI_LAUNCH(2, 384); else KQ_SI_LAUNCH(2, 512); }
#undef KQ_SI_LAUNCH
    ++*launches;
    if (hot.buf) {
        if (tuple_words(tv.k) == 1) hot_slices_kernel<1><<<(unsigned)sm_count() * 4, 256, 0, st>>>(tv, tuples, hot, ovf, ovf_cap, ovf_n);
        else hot_slices_kernel<2><<<(unsigned)sm_count() * 4, 256, 0, st>>>(tv, tuples, hot, ovf, ovf_cap, ovf_n);
        ++*launches;
    }
    return cudaGetLastError();
}

}  // namespace kq

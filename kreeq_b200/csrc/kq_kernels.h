// kq_kernels.h -- host-callable launchers of the sm_100a kernels (kq_kernels.cu, kq_table_kernels.cu).
// Every launcher enqueues on `st`, returns the cudaError_t of the launch and adds the
// number of kernels it launched to *launches.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "kq_device.cuh"

namespace kq {

// record as moved between host and device by load / save / union (48 bytes)
constexpr uint32_t F_PLACEHOLDER = 2u;       // XRec.flags: the 8-bit map entry is a cov == 255 placeholder
struct XRec { uint64_t key; Rec32 r; uint32_t flags; };
static_assert(sizeof(XRec) == 48, "XRec layout");

// device-side statistics block filled by stats_scan
constexpr int HIST_BINS = 65536;
struct DevStats {
    unsigned long long total, unique, distinct, edges;   // over non-placeholder T entries (logical records)
    unsigned long long big_n;                            // entries with cov >= HIST_BINS appended to big[]
};

// (key, edgeBit) tuples in device memory: one 64-bit word (key << 8 | edge) while 2k + 8 <= 64,
// otherwise two words {key, edge}.
__host__ __device__ inline int tuple_words(int k) { return 2 * k + 8 <= 64 ? 1 : 2; }

// where the count pipeline takes its k-mer instances from
struct Source {
    const uint64_t *bases, *invalid;   // packed sequence (tuples == nullptr) ...
    uint64_t pos0, npos;               // ... k-mer start positions [pos0, pos0 + npos)
    const uint64_t *tuples;            // or tuples already extracted (multi-GPU: received from peers)
    const unsigned long long *ntuples; // device pointer to their number
    uint64_t ntuples_max;              // host-side upper bound of *ntuples (grid sizing)
    int ordered;                       // host side: the tuples already are in slice order (routed pipeline)
};

struct RouteArgs {           // destination of extracted tuples (multi-GPU count)
    uint64_t *tuples; unsigned long long cap_per_rank; unsigned long long *counts;
    int nranks, shift;       // owner = (key & 127) >> shift
};

// routed (multi-GPU) scatter: bins are (owner rank, 2^nb_bits equal ranges of the slice hash); the tuples of
// owner d go to out[d], the receive slab of rank d (peer-mapped; out[me] is the local one)
constexpr int KQ_MAX_RANKS = 16;
struct RouteMap {
    int nranks, shift, nb_bits;          // owner = (key & 127) >> shift
    uint64_t *out[KQ_MAX_RANKS];
};

// routed validate: where a rank's queries go (the owners' inboxes) and where an owner's answers go (the origins'
// row slabs); every pointer is mapped into this rank's address space (peer access or CUDA IPC)
struct VRoute {
    int nranks, me, shift;                         // owner = (key & 127) >> shift
    unsigned long long cap;                        // queries per (owner, sender) region of an inbox
    void *inbox[KQ_MAX_RANKS];                     // inbox[o] + (me * cap + slot) * 16
    unsigned long long *counts[KQ_MAX_RANKS];      // counts[o][me] = queries this rank left at owner o
    Base *rows[KQ_MAX_RANKS];                      // row slab of origin s: [cap rows][answer boxes: nranks x cap rows]
    unsigned long long *cursor;                    // local append cursors, one per owner
    unsigned *poslog;                              // local: poslog[o * cap + slot] = row of the query this rank left in slot `slot` at owner o
    unsigned long long *sent;                      // local: queries left at every owner in the last round (kept for the place kernel)
};

int sm_count();   // multiprocessors of the current device (cached)

// tuning knobs (defaults are the measured best on B200); kq_set_tuning changes them
struct Tuning {
    int scatter_cap;           // bin capacity of the level-1 scatter (0 = largest that fits)
    int rt_variant, sc_variant, scatter_direct;
    int insert_ctas, validate_ctas;
    int single_slice_bits;     // a table of up to 2^this slots is one slice (fused extract + insert)
    int piece_bits;            // larger tables: log2 slots per slice
    unsigned long long part;   // tuples per task of the level-2 kernels
    int smem_insert;           // 0 = count with L2 CAS (insert_tuples_kernel) also on multi-slice tables
    int direct_pct;            // batches below this share of the table's slots skip the sweep (fused kernel)
    int db_host;               // .kreeq codec on the host (cross-check of the device codec)
    int coarse_bias;           // level-1 bins: 2^v times fewer (and larger) than the square-root split
    int db_ordered;            // 1 = databases are decoded into table order before they are put / merged
    int one_pass;              // 1 = level-1 scatter into fixed regions, no histogram pass (overflow -> list)
    int cta_lock;              // 1 = CTA-owned slices are counted with the lock-byte protocol (cta_insert_kernel) instead of shared-memory CAS
    int route_smem;            // 1 = the routed slab goes through the level-2 scatter + shared-memory insert (0: L2 CAS insert over the slab)
    int warp_insert;           // 1 = slices of at most 2^10 slots are counted by one warp each (no shared-memory atomics)
    int s3_threads, s3_ctas;   // shared-memory insert: threads per CTA (256 / 384 / 512), CTAs per SM in the grid
    int fs_variant;            // level-2 scatter: 0 auto (two CTAs per SM when the bins allow), 1 one CTA, 2 two CTAs
    int pipeline;              // big tables: 4 one-pass scatter + shared-memory insert, 3 two-level scatter + shared-memory
                               // insert, 2 level-1 scatter + L2 CAS insert
    int cursor_stride;         // one-pass scatter: 4-byte words between the cursors of two slices
};
Tuning &tuning();
// shared memory the scatter kernel needs for `nslices` bins of `cap` tuples; 0 if it cannot fit
size_t scatter_smem_bytes(unsigned long long nslices, int cap, int tw);
int scatter_pick_cap(unsigned long long nslices, int tw, int max_cap = 512);   // largest power of two <= max_cap that fits, 0 = none
// routed bins flush straight into a peer's slab: chunks of at most 32 tuples (256-byte peer stores), so a
// (sender, bin) region is padded by fewer than 32 tuples
constexpr int KQ_ROUTE_MAX_CAP = 32;

// ---- count pipeline (rows A2 + A3) -------------------------------------------------------------------
// single-slice tables / small inputs: fused extract + insert
cudaError_t launch_count_direct(cudaStream_t st, const TableView &tv, const Source &src,
                                unsigned long long *instances, uint64_t *ovf, unsigned long long ovf_cap,
                                unsigned long long *ovf_n, uint64_t *launches);
// one-pass scatter (big tables): tuple -> out[s * cap + atomicAdd(cursor[s * stride], 1)], s = slice of the key;
// instances of over-subscribed slices are counted at once (L2 CAS); *instances += valid windows (may be NULL)
cudaError_t launch_scatter_fine(cudaStream_t st, const TableView &tv, const Source &src, uint64_t *out, unsigned cap,
                                unsigned *cursor, int stride, unsigned long long *instances, uint64_t *ovf,
                                unsigned long long ovf_cap, unsigned long long *ovf_n, uint64_t *launches);
// 1. tuples per slice -> hist[nslices]
cudaError_t launch_slice_hist(cudaStream_t st, const TableView &tv, const Source &src, unsigned long long *hist,
                              uint64_t *launches);
// 2. exclusive scan: cursor[p] = start[p] = first tuple index of bin p (start may be NULL; cursor is advanced by
//    the scatter, start is not); *total = number of tuples; *instances += total
cudaError_t launch_slice_offsets(cudaStream_t st, unsigned long long nbins, const unsigned long long *hist,
                                 unsigned long long *cursor, unsigned long long *start, unsigned long long *total,
                                 unsigned long long *instances, uint64_t *launches);
// 3. scatter tuples into slice order (out holds *total tuples afterwards)
// one-pass level-1 scatter: fixed region ends per coarse bin; `ovf` = list for the chunks that do not fit (as large as
// the batch); stat[0] = tuples in that list, stat[1] = sentinels written into regions, stat[2] = free for the caller
struct Regions { unsigned long long rcap; uint64_t *ovf; unsigned long long ovf_cap; unsigned long long *stat; };   // rcap = 0: exact layout
cudaError_t launch_region_init(cudaStream_t st, unsigned nbins, unsigned long long rcap, unsigned long long *cursor,
                               unsigned long long *start, unsigned long long *end, unsigned long long *stat, uint64_t *launches);
cudaError_t launch_region_counts(cudaStream_t st, unsigned nbins, unsigned long long *cursor, const unsigned long long *start,
                                 const unsigned long long *end, unsigned long long *hist, unsigned long long *total,
                                 const unsigned long long *stat, unsigned long long *instances, uint64_t *launches);
cudaError_t launch_slice_scatter(cudaStream_t st, const TableView &tv, const Source &src, uint64_t *out,
                                 unsigned long long *cursor, int cap, uint64_t *launches, Regions rg = Regions{0, nullptr, 0, nullptr});
// 4. insert slice-ordered tuples; tuples whose slice is full are appended to ovf (capacity ovf_cap,
//    ovf_n counts every failure, also those that did not fit)
cudaError_t launch_insert_tuples(cudaStream_t st, const TableView &tv, const uint64_t *tuples,
                                 const unsigned long long *n, uint64_t n_max, uint64_t *ovf,
                                 unsigned long long ovf_cap, unsigned long long *ovf_n, unsigned long long *ticket,
                                 uint64_t *launches, uint64_t n_cap = 0);
// ---- level 2 + shared-memory insert (kq_count3.cu) --------------------------------------------------------
// bins of the level-1 scatter: 2^cshift consecutive slices each
inline unsigned coarse_bins(const TableView &tv) { return (unsigned)(((tv.nslices - 1) >> tv.cshift) + 1); }
// at most `part` consecutive tuples [lo, hi) of the level-1 output whose keys live in slices [s_lo, s_lo + ns)
struct FineTask { unsigned long long lo, hi; unsigned s_lo, ns; };
constexpr int KQ_ROUTE_PLAN_BINS = 2048;      // hash ranges per owner the routed plan kernel handles
constexpr int KQ_S3_NT = 512;                 // threads per CTA of the shared-memory insert
unsigned fine_pick_cap(unsigned ns_max, int tw, int ctas_per_sm = 1);   // flush chunk of the level-2 scatter, 0 = bins do not fit
cudaError_t launch_fine_plan(cudaStream_t st, const TableView &tv, unsigned ncoarse, const unsigned long long *start,
                             const unsigned long long *n, unsigned long long part, FineTask *tasks, unsigned task_cap,
                             unsigned *ntasks, unsigned *task_first, uint64_t *launches);
// per coarse bin: hist2[t][f] -> offset of task t inside slice f's region; fstart / fcount of every slice
cudaError_t launch_fine_offsets(cudaStream_t st, const TableView &tv, unsigned ncoarse, const unsigned long long *start,
                                const unsigned *task_first, unsigned *hist2, unsigned ns_max, unsigned long long *fstart,
                                unsigned long long *fcount, uint64_t *launches);
// tasks over the receive slab of rank `me` (routed pipeline; H = all-gathered histograms, cap = flush chunk)
cudaError_t launch_route_plan(cudaStream_t st, const TableView &tv, const RouteMap &rt, int me, int cap, const unsigned long long *H,
                              unsigned long long part, FineTask *tasks, unsigned task_cap, unsigned *ntasks, unsigned *task_first, uint64_t *launches);
cudaError_t launch_route_fine_offsets(cudaStream_t st, const TableView &tv, int nb_bits, const FineTask *tasks, const unsigned *task_first,
                                      unsigned *hist2, unsigned ns_max, unsigned long long *fcount, uint64_t *launches);
// hist2 == NULL: fhist[s] += tuples of slice s; else hist2[t * ns_max + f] = tuples of slice s_lo + f in task t
cudaError_t launch_fine_hist(cudaStream_t st, const TableView &tv, const uint64_t *in, const FineTask *tasks,
                             const unsigned *ntasks, unsigned task_bound, unsigned ns_max, unsigned long long *fhist,
                             unsigned *hist2, uint64_t *launches);
// out = tuples in slice order (cursor[s] = first tuple index of slice s, advanced)
cudaError_t launch_fine_scatter(cudaStream_t st, const TableView &tv, const uint64_t *in, const FineTask *tasks,
                                const unsigned *ntasks, unsigned task_bound, unsigned ns_max, uint64_t *out,
                                unsigned long long *cursor, const unsigned *hist2, const unsigned long long *fstart, uint64_t *launches);
// count the tuples [fstart[s], fstart[s] + fcount[s]) of every slice s in shared memory
// (fstart == NULL: fcount is the one-pass scatter's array of 32-bit cursors, slice s owns [s * cap, s * cap + min(cursor[s * stride], cap)))
// slices with more than `limit` tuples of a batch are left to hot_slices_kernel: buf[0] = entries, then {start, n, slice}
struct HotList { unsigned long long *buf; unsigned cap; unsigned limit; };
constexpr unsigned KQ_HOT_CAP = 1u << 16;
cudaError_t launch_smem_insert(cudaStream_t st, const TableView &tv, const uint64_t *tuples, const unsigned long long *fstart,
                               const void *fcount, unsigned long long cap, int stride, uint64_t *ovf,
                               unsigned long long ovf_cap, unsigned long long *ovf_n, unsigned long long *ticket, uint64_t *launches, HotList hot = HotList{nullptr, 0, 0});

// routed pipeline, steps 1-3: hist[nranks << nb_bits] of this rank's tuples; from the all-gathered
// histograms H[nranks][P] the start of this rank's tuples inside every destination bin (cursor[P]) and what
// every rank receives (recv[d] = extent of d's slab with the alignment padding, recv[nranks + d] = real
// tuples in it); scatter straight into the owners' receive slabs
cudaError_t launch_route_hist(cudaStream_t st, const TableView &tv, const Source &src, const RouteMap &rt,
                              unsigned long long *hist, uint64_t *launches);
cudaError_t launch_route_offsets(cudaStream_t st, const RouteMap &rt, int me, int cap, int tw, unsigned long long slab_cap,
                                 const unsigned long long *H, unsigned long long *front, unsigned long long *back,
                                 unsigned long long *recv, uint64_t *launches);
cudaError_t launch_route_scatter(cudaStream_t st, const TableView &tv, const Source &src, const RouteMap &rt,
                                 unsigned long long *cursor, unsigned long long *back, int cap, uint64_t *launches);
// multi-GPU, NCCL transport: extract every valid window and append its tuple to the slab of its owner rank
cudaError_t launch_extract_route(cudaStream_t st, const TableView &tv, const Source &src, RouteArgs ra,
                                 uint64_t *launches);

// ---- validate / lookups (rows A5, host consumers) ------------------------------------------------------
// peers != nullptr (device array of npeers views, one per rank): probe every position in its owner's shard.
// positions [c0, c0+n) of a segment with kcount windows; the packed arrays start at segment
// position `base0` (so bit position of c is c-base0); out is indexed by c-c0.
cudaError_t launch_validate(cudaStream_t st, const TableView &tv, const uint64_t *B, const uint64_t *M,
                            uint64_t base0, uint64_t c0, uint64_t n, uint64_t kcount, uint32_t cutoff,
                            Base *out, unsigned long long *acc3, const TableView *peers, int npeers, uint64_t *launches);
// routed validate: phase 1 on the origin (queries -> owners' inboxes, counts published), phase 2 on the owner
struct WideRow { unsigned long long pos; uint32_t fw, bw, cov, isfw; };   // == kq_wide_row (24 bytes; isfw = isFw byte + padding)
cudaError_t launch_compact_rows(cudaStream_t st, const Base *rows, uint64_t n, uint64_t pos0, uint32_t *out4, void *wide,
                                unsigned long long wide_cap, unsigned long long *wide_n, uint64_t *launches);
cudaError_t launch_vroute_send(cudaStream_t st, const TableView &tv, const uint64_t *B, const uint64_t *M, uint64_t base0,
                               uint64_t c0, uint64_t n, uint64_t kcount, Base *rows, unsigned long long *acc3, const VRoute &vr,
                               uint64_t *launches, uint64_t row0 = 0, bool publish = true);
cudaError_t launch_vroute_answer(cudaStream_t st, const TableView &tv, const void *inbox, const unsigned long long *counts,
                                 uint64_t max_queries, uint32_t cutoff, unsigned long long *acc3, const VRoute &vr, uint64_t *launches);
cudaError_t launch_vroute_place(cudaStream_t st, uint64_t max_queries, const VRoute &vr, uint64_t *launches);
cudaError_t launch_lookup(cudaStream_t st, const TableView &tv, const uint64_t *keys, uint64_t n, Rec32 *rec,
                          uint8_t *found, uint64_t *launches);

// ---- table maintenance (rows A3s, A4, A8) ----------------------------------------------------------------
cudaError_t launch_stats(cudaStream_t st, const TableView &tv, DevStats *ds, unsigned long long *hist,
                         unsigned long long *big, unsigned long long big_cap, uint64_t *launches);
// export occupied slots (logical record, or F_PLACEHOLDER) in arbitrary order
cudaError_t launch_export(cudaStream_t st, const TableView &tv, XRec *out, unsigned long long cap,
                          unsigned long long *n_out, uint64_t *launches);
// move every slot of the old table into tv (grow)
cudaError_t launch_rehash(cudaStream_t st, const TableView &tv, const Slot *from, unsigned long long from_cap,
                          uint64_t *launches);
cudaError_t launch_side_rehash(cudaStream_t st, const TableView &tv, int which, const SideSlot *from,
                               unsigned long long from_cap, uint64_t *launches);
// load: unique keys, raw 8-bit-map entries (flags carry F_PLACEHOLDER) -> T
cudaError_t launch_put(cudaStream_t st, const TableView &tv, const XRec *in, uint64_t n, uint64_t *launches);
// exact union: add logical records
cudaError_t launch_add_records(cudaStream_t st, const TableView &tv, const XRec *in, uint64_t n, uint64_t *launches);
// reference-compatible union: DBG::mergeSubMaps for the 8-bit entries of one later database
cudaError_t launch_merge_ref(cudaStream_t st, const TableView &tv, const XRec *in, uint64_t n, uint64_t *launches);

// ---- .kreeq file images on the device (row A8 / N1, kq_db_kernels.cu) -----------------------------------------
// one submap of one .map.<m>.bin inside a device image: byte offsets of its header, control bytes and slots
struct DbBucket { unsigned long long hdr_off, ctrl_off, slot_off, cap, n; };
unsigned long long db_capacity_for(unsigned long long n);
// records per (map, submap) bucket (counts[128 * 256]) of the maps [lo, hi); 32-bit spill records -> hc
cudaError_t launch_db_count(cudaStream_t st, const XRec *xs, unsigned long long n, int lo, int hi,
                            unsigned long long *counts, XRec *hc, unsigned long long hc_cap, unsigned long long *hc_n,
                            uint64_t *launches);
// lay out headers + control bytes, place every record (phmap probe order), write the slots; img is zeroed
cudaError_t launch_db_build(cudaStream_t st, const XRec *xs, unsigned long long n, int lo, int hi, unsigned char *img,
                            const DbBucket *buckets, unsigned long long *pos_scratch, uint64_t *launches);
// *out += records that can create a remainder / maps32 entry (a field above a slot lane, or a placeholder)
constexpr int KQ_DB_ORDER_BITS = 12;   // a database is decoded into 4096 bands of the slice hash (= table order)
cudaError_t launch_db_band_hist(cudaStream_t st, const unsigned char *img, const DbBucket *buckets, unsigned nbuckets,
                                unsigned long long *hist, uint64_t *launches);
cudaError_t launch_db_decode_ordered(cudaStream_t st, const unsigned char *img, const DbBucket *buckets, unsigned nbuckets,
                                     XRec *out, unsigned long long cap_out, unsigned long long *cursor, uint64_t *launches);
cudaError_t launch_db_count_big(cudaStream_t st, const XRec *xs, unsigned long long n, unsigned long long *out, uint64_t *launches);
// uploaded file images -> raw records
cudaError_t launch_db_decode(cudaStream_t st, const unsigned char *img, const DbBucket *buckets, unsigned nbuckets,
                             XRec *out, unsigned long long cap_out, unsigned long long *n_out, uint64_t *launches);

}  // namespace kq

/* kreeq_b200.h -- C-ABI of libkreeq_b200.so, the B200 (sm_100a) implementation of
 * kreeq's data-parallel k-mer hot path: DBG k-mer counting, .kreeq database
 * save/load/union, and assembly k-mer validation (QV + per-base profile).
 *
 * The reference (vgl-hub/kreeq) has no FFI; its seam is the CRTP base
 * Kmap<DBG,...> (gfalibs/include/kmer.h:53-240) specialised by class DBG
 * (include/kreeq.h:150-277).  Each entry point below names the reference
 * member(s) it replaces (paths relative to the reference tree).
 *
 * Conventions: plain pointers and sizes, caller-owned HOST memory unless a
 * parameter is named d_* (device pointer on the context's GPU).  Every call
 * returns 0 or a negative KQ_E* code and never exit()s; kq_last_error() gives
 * the message.  There is NO CPU fallback: without a usable CUDA device
 * kq_create fails with KQ_ENODEV.
 * Threads: a handle may be driven by ONE thread at a time; different handles (one per GPU) may be driven by
 * different threads concurrently (the host pool the library packs and writes files with serialises its callers).
 * One handle = one GPU = one shard of the key space.  Multi-GPU = one process
 * (or handle) per GPU, kq_set_shard(rank, nranks), with tuples routed by
 * kq_owner_of().
 */
#ifndef KREEQ_B200_H
#define KREEQ_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KQ_OK        0
#define KQ_EINVAL   -1   /* bad argument (k outside 1..32, NULL, ...)              */
#define KQ_ENODEV   -2   /* no CUDA device / CUDA runtime failure at start-up      */
#define KQ_ECUDA    -3   /* CUDA error during a call                               */
#define KQ_ENOMEM   -4   /* host or device allocation failed / table cannot grow   */
#define KQ_EIO      -5   /* file missing, unreadable or not a .kreeq layout        */
#define KQ_ECORRUPT -6   /* "int32 map missing 255 value from int8 map"            */
#define KQ_ESTATE   -7   /* call not valid in the handle's current state           */

#define KQ_MAP_COUNT 128 /* Kmap::mapCount, gfalibs/include/kmer.h:68 */

typedef struct kq_ctx kq_ctx;

/* == DBGkmer32, include/kreeq.h:69-70 */
typedef struct { uint32_t fw[4], bw[4], cov; } kq_rec32;
/* == DBGbase, include/input.h:4-9 (16 bytes; isFw is a C++ bool there) */
typedef struct { uint32_t fw, bw, cov; uint8_t isFw; uint8_t pad[3]; } kq_base;
/* DBG::summary/DBstats counters, src/graph-builder.cpp:240-295 */
typedef struct { uint64_t total, unique, distinct, missing, edges; } kq_stats;
/* totMissingKmers, totKcount, totEdgeMissingKmers, src/kreeq.cpp:223-225 */
typedef struct { uint64_t missing, kcount, edge_missing; } kq_qv;

/* ---- lifetime ------------------------------------------------------------ */

/* replaces DBG knav(userInput) / Kmap ctor, kmer.h:126-153.  device = CUDA
 * ordinal.  expected_distinct sizes the first table (0 = small default; the
 * table doubles on demand). */
int  kq_create(kq_ctx **out, int k, int device, uint64_t expected_distinct);
void kq_destroy(kq_ctx *ctx);
const char *kq_last_error(kq_ctx *ctx);          /* ctx may be NULL: last create error */
int  kq_k(kq_ctx *ctx);

/* Run the context's kernels on a caller-owned stream (a cudaStream_t, e.g.
 * torch.cuda.current_stream().cuda_stream).  NULL restores the internal one. */
int  kq_set_stream(kq_ctx *ctx, void *cuda_stream);
int  kq_sync(kq_ctx *ctx);

/* Key-space shard of this handle: it owns maps [rank*128/nranks,
 * (rank+1)*128/nranks).  nranks in {1,2,4,8,...,128}. */
int  kq_set_shard(kq_ctx *ctx, int rank, int nranks);
/* owner rank of a canonical key: (key % 128) / (128 / nranks) */
int  kq_owner_of(uint64_t key, int nranks);

/* ---- count (rows A1-A3) ---------------------------------------------------- */

/* replaces Kmap::traverseInReads(std::string*) kmer.h:753-764 +
 * DBG::hashSequences gb:34-126 + DBG::processBuffers gb:128-223.
 * ascii is one batch of reads each followed by 'N' exactly as getKmers()
 * (gfalibs/include/functions.h:69-122) builds it; BORROWED, consumed (packed
 * to 2 bit into pinned staging) before return; the H2D copy is asynchronous and
 * the batch joins the device-resident pending stream (kq_set_flush_positions).
 * With nranks>1 only keys owned by this shard are inserted (every rank may be
 * fed the same batch), see also kq_dev_extract_route. */
int  kq_count_push(kq_ctx *ctx, const char *ascii, size_t len);
/* replaces Kmap::finalize kmer.h:821-828: barrier, table resident and final. */
int  kq_count_finish(kq_ctx *ctx);
/* k-mer instances inserted so far by this handle (valid windows, row A2) */
int  kq_count_instances(kq_ctx *ctx, uint64_t *n);
/* drop every record, keep allocations */
int  kq_clear(kq_ctx *ctx);

/* ---- persistence (row A8) -------------------------------------------------- */

/* replaces Kmap::dumpMap kmer.h:698-709 + dumpHighCopyKmers :563-576 + the
 * .index writer src/kreeq-output.cpp:88-94.  Writes the maps this shard owns;
 * rank 0 also writes .index; .map.hc.bin is written when nranks==1, otherwise
 * use kq_db_save_hc_part / kq_db_merge_hc_parts. */
int  kq_db_save(kq_ctx *ctx, const char *dir);
int  kq_db_save_hc_part(kq_ctx *ctx, const char *dir);       /* <dir>/.hc.part.<rank> */
int  kq_db_merge_hc_parts(const char *dir, int nranks);      /* -> <dir>/.map.hc.bin  */
/* replaces Input::loadGraph src/input.cpp:56-74 + Kmap::loadMap kmer.h:592-602 +
 * loadHighCopyKmers :605-618 (k must equal the one in <dir>/.index). */
int  kq_db_load(kq_ctx *ctx, const char *dir);
/* k stored in <dir>/.index, or a negative error */
int  kq_db_read_k(const char *dir);
/* host-only codec of the same layout (no GPU needed; tools and format tests):
 * write logical records (<=254 instances -> 8-bit record, else cov=255 placeholder
 * + 32-bit record in .map.hc.bin), read back the lookup-visible
 * records sorted by key (*n in: capacity, out: count). */
int  kq_db_write_records(const char *dir, int k, const uint64_t *keys, const kq_rec32 *recs, size_t n);
int  kq_db_read_records(const char *dir, uint64_t *keys, kq_rec32 *recs, size_t *n);

/* ---- union (row A4) ---------------------------------------------------------- */

/* replaces DBG::kunion gb:297-351 / Kmap::mergeMaps kmer.h:666-695 /
 * DBG::mergeSubMaps gb:353-432.  dirs[0] is "DB0".  ref_compatible=1
 * reproduces the reference's order-dependent treatment of high-copy keys;
 * 0 gives the exact saturating sum.  Result replaces the handle's table. */
int  kq_union(kq_ctx *ctx, const char *const *dirs, int n, int ref_compatible);

/* ---- stats (row A3s) ---------------------------------------------------------- */

/* replaces Kmap::stats kmer.h:831-852 + DBG::summary/DBstats gb:240-295 */
int  kq_stats_get(kq_ctx *ctx, kq_stats *out);
/* finalHistogram (kmer.h:895-899) sorted by coverage; *len in: capacity, out: bins */
int  kq_hist(kq_ctx *ctx, uint64_t *cov, uint64_t *n, size_t *len);

/* ---- validate (rows A5, A6) ---------------------------------------------------- */

/* replaces DBG::evaluateSegment src/kreeq.cpp:110-229 for one segment (no
 * N/n/X/x inside).  out has len entries; entries >= len-k+1 are zeroed.
 * acc ACCUMULATES.  With nranks>1 only positions whose key this shard owns are
 * filled/counted (same as the reference's mapRange test, kreeq.cpp:150).
 * `out` may be pageable memory (rows are staged through pinned buffers and copied once on the host) or
 * page-locked memory (cudaHostAlloc / cudaHostRegister: rows are copied into it directly, about twice as
 * fast since 16 bytes per base make the host copy the bottleneck). */
int  kq_validate_segment(kq_ctx *ctx, const char *seq, uint64_t len,
                         uint32_t cov_cutoff, kq_base *out, kq_qv *acc);
/* Same pass with a quarter of the bytes back over PCIe (16 bytes per base make the device-to-host copy the whole cost of
 * an end-to-end validate): out[i] = {cov, fw, bw} as bytes + flags (bit 0 isFw, bit 1 "wide"); a row with a value above
 * 254 has the wide bit set, 255 in the fields that do not fit, and its full DBGbase row with its position in `wide`
 * (*nwide in: capacity of `wide`, out: rows listed; more rows than the capacity or than 2^20 -> KQ_ENOMEM, use
 * kq_validate_segment).  kq_expand_rows rebuilds the DBGbase array on the host.  Ours; the reference has no such form
 * (its writers take DBGbase, src/kreeq-output.cpp:138-399). */
typedef struct { uint8_t cov, fw, bw, flags; } kq_base4;
typedef struct { uint64_t pos; kq_base row; } kq_wide_row;
int  kq_validate_segment_compact(kq_ctx *ctx, const char *seq, uint64_t len, uint32_t cov_cutoff, kq_base4 *out,
                                 kq_wide_row *wide, size_t *nwide, kq_qv *acc);
int  kq_expand_rows(const kq_base4 *in, uint64_t len, const kq_wide_row *wide, size_t nwide, kq_base *out);
/* errorRate src/kreeq.cpp:36-40 and QV = -10 log10(err) :87,96 */
double kq_error_rate(uint64_t missing, uint64_t total, int k);
double kq_qv_value(uint64_t missing, uint64_t total, int k);

/* ---- host consumers (subgraph / variants) ---------------------------------------- */

/* replaces the lookup idiom maps[i]->find(key) [+ maps32[i]->find(key)]
 * src/kreeq.cpp:150-166,231-244; src/subgraph.cpp:235-245; src/variants.cpp:118-133.
 * found[i]=0 leaves rec[i] untouched. */
int  kq_lookup_batch(kq_ctx *ctx, const uint64_t *keys, size_t n, kq_rec32 *rec, uint8_t *found);

/* kq_lookup_batch without any allocation (no cudaMalloc / cudaFree, no host heap): keys and answers are staged in
 * ONE file-scope device array of KQ_STATIC_LOOKUP_MAX keys (2.9 MiB per GPU that loads the library, held until
 * exit).  For callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it.  No counterpart in the reference.
 * Same answers as kq_lookup_batch (found[i]=0 leaves rec[i] untouched).  n > KQ_STATIC_LOOKUP_MAX -> KQ_EINVAL;
 * pending reads (kq_count_push without kq_sync / kq_count_finish) -> KQ_ESTATE, because counting them may grow
 * the table.  Every call shares the array: calls that overlap on one GPU (two handles, two threads) race. */
#define KQ_STATIC_LOOKUP_MAX 65536
size_t kq_lookup_static_max(void);
int  kq_lookup_batch_static(kq_ctx *ctx, const uint64_t *keys, size_t n, kq_rec32 *rec, uint8_t *found);

/* logical {key -> record} set of this shard sorted by key (parity checks,
 * writers).  *n in: capacity of keys/recs, out: records present. */
int  kq_export(kq_ctx *ctx, uint64_t *keys, kq_rec32 *recs, size_t *n);
int  kq_distinct(kq_ctx *ctx, uint64_t *n);

/* ---- packed-sequence stage API (device-resident; bench `value`, multi-GPU) ------ */

/* 2-bit packing used on the wire: base i -> bits [2*(i%32), +2) of word i/32
 * (A0 C1 G2 T3, first base least significant == the key's digit order,
 * kmer.h:785-788); invalid[i/64] bit i%64 set for every non-ACGT byte.
 * kq_packed_words(len) gives the number of uint64 words of EACH array the
 * kernels may read (includes the zero / all-invalid tail padding). */
size_t kq_packed_words(size_t len);
int    kq_pack_host(const char *ascii, size_t len, uint64_t *bases, uint64_t *invalid, int nthreads);

/* fused extract + insert of k-mer start positions [pos0, pos0+npos) of a packed
 * device-resident sequence of `len` bases (only keys this shard owns). */
int  kq_dev_count_packed(kq_ctx *ctx, const uint64_t *d_bases, const uint64_t *d_invalid,
                         uint64_t len, uint64_t pos0, uint64_t npos);
/* (key, edgeBit) tuples on the device: 8 bytes (key << 8 | edgeBit) while 2k + 8 <= 64, else 16 bytes
 * {uint64 key, uint64 edgeBit}.  kq_tuple_bytes gives the size for this handle's k. */
int  kq_tuple_bytes(kq_ctx *ctx);
/* extract only: the tuple of every valid window is appended to the slab of its owner rank
 * (kq_owner_of): slab r starts at d_tuples + r * cap_per_rank * kq_tuple_bytes; d_counts (nranks
 * uint64, zeroed by the caller) receives the number per rank.  Overflow of cap_per_rank is
 * reported by kq_sync / KQ_ENOMEM. */
int  kq_dev_extract_route(kq_ctx *ctx, const uint64_t *d_bases, const uint64_t *d_invalid,
                          uint64_t len, uint64_t pos0, uint64_t npos,
                          void *d_tuples, uint64_t cap_per_rank, unsigned long long *d_counts);
/* insert n tuples (row A3) already resident on the device (all owned by this shard) */
int  kq_dev_insert_tuples(kq_ctx *ctx, const void *d_tuples, uint64_t n);
/* ---- routed count over peer memory (count's exchange step) -----------------------
 * One process per GPU, handles sharded with kq_set_shard.  The all-to-all of (key, edgeBit) tuples the
 * reference design calls for (8e "Collective") is fused into the scatter kernel: every rank writes the
 * tuples a peer owns straight into that peer's receive slab (peer-mapped memory, NVLink), in slice order,
 * and the peer only runs the insert kernel over its slab.  Per round of k-mer start positions:
 *   kq_route_hist      tuples this rank sends into every (owner, hash range) bin -> d_hist[kq_route_bins]
 *   (caller)           all-gather the histograms of all ranks: d_all_hist[nranks][bins]; this collective
 *                      also orders the round after every rank's previous kq_route_insert
 *   kq_route_scatter   offsets from d_all_hist, capacity check (same verdict on every rank), scatter into
 *                      the owners' slabs; *nrecv = tuples this rank receives
 *   (caller)           barrier on the stream (all slabs complete)
 *   kq_route_insert    count the nrecv tuples of the local slab (row A3)
 * Set-up: kq_route_setup allocates the slab (slab_tuples tuples; nb_bits = log2 hash ranges per owner,
 * nranks << nb_bits bins must fit shared memory: <= 2048) and returns its device pointer and a 64-byte
 * CUDA IPC handle; peers open it with kq_route_open; kq_route_connect takes the nranks slab pointers
 * (own slab at index rank).  Within one process pass the pointers directly. */
int  kq_route_setup(kq_ctx *ctx, uint64_t slab_tuples, int nb_bits, void **d_slab, void *ipc_handle64);
int  kq_route_open(kq_ctx *ctx, const void *ipc_handle64, void **d_ptr);
int  kq_route_connect(kq_ctx *ctx, void *const *slabs);
int  kq_route_bins(kq_ctx *ctx);
int  kq_route_hist(kq_ctx *ctx, const uint64_t *d_bases, const uint64_t *d_invalid, uint64_t len,
                   uint64_t pos0, uint64_t npos, unsigned long long *d_hist);
int  kq_route_scatter(kq_ctx *ctx, const uint64_t *d_bases, const uint64_t *d_invalid, uint64_t len,
                      uint64_t pos0, uint64_t npos, const unsigned long long *d_all_hist, uint64_t *nrecv);
int  kq_route_insert(kq_ctx *ctx, uint64_t n);

/* ---- validate against every rank's shard over peer memory (validate's exchange step) ----------
 * Instead of routing (key, origin) queries to the owners and the answers back, a rank probes the owner's
 * table directly: kq_peer_export publishes this shard (CUDA IPC handles of its tables + geometry, 256
 * bytes; the tables then stay put until kq_peer_disconnect), kq_peer_connect takes the nranks blobs in rank
 * order.  While connected, kq_validate_segment / kq_dev_validate_packed answer EVERY position (not only the
 * ones this shard owns), so segments can be dealt to ranks and only the three QV counters are summed. */
#define KQ_PEER_BLOB 256
int  kq_peer_export(kq_ctx *ctx, void *blob256);
int  kq_peer_connect(kq_ctx *ctx, const void *blobs);
int  kq_peer_connect_local(kq_ctx *ctx, kq_ctx *const *handles);   /* shards held by this process */
int  kq_peer_disconnect(kq_ctx *ctx);

/* ---- validate across GPUs with routed queries (SURVEY 8e "Collective": all-to-all of (key, origin) queries ->
 * the owners probe their own HBM -> answers back into the origin's row slab) ----------------------------------
 * kq_vroute_setup allocates this rank's inbox (nranks regions of `chunk_positions` 16-byte queries + the
 * per-sender counts) and its row slab (chunk_positions kq_base rows) and returns their device pointers and
 * CUDA IPC handles (64 bytes each, may be NULL); kq_vroute_connect takes every rank's inbox and row-slab
 * pointers (own ones at index rank; peers' opened with kq_route_open or passed directly inside one process).
 * Per chunk of at most chunk_positions assembly positions, on EVERY rank and in this order:
 *   kq_vroute_send     positions [c0, c0+n) of a segment with kcount windows (packed arrays start at segment
 *                      position base0): queries to the owners' inboxes; rows of windows without a key are final
 *   (caller)           barrier on the streams of all ranks
 *   kq_vroute_answer   probe what arrived, store the finished rows into the origins' slabs
 *   (caller)           barrier; then rows [0, n) of kq_vroute_rows() are this rank's DBGbase rows of the chunk
 * kq_vroute_counters adds this rank's share of the three QV counters (sum over ranks = the totals). */
int  kq_vroute_setup(kq_ctx *ctx, uint64_t chunk_positions, void **d_inbox, void *ipc_inbox64, void **d_rows, void *ipc_rows64);
int  kq_vroute_connect(kq_ctx *ctx, void *const *inboxes, void *const *rows);
int  kq_vroute_send(kq_ctx *ctx, const uint64_t *d_bases, const uint64_t *d_invalid, uint64_t base0, uint64_t c0, uint64_t n, uint64_t kcount);
/* several segments in ONE round (one pair of barriers for all of them): rows of this call go to [row0, row0 + n) of the
 * row slab, the counts are published to the owners with the call that passes last = 1 (row0 + n <= chunk of
 * kq_vroute_setup).  kq_vroute_send == kq_vroute_send_at(.., 0, 1). */
int  kq_vroute_send_at(kq_ctx *ctx, const uint64_t *d_bases, const uint64_t *d_invalid, uint64_t base0, uint64_t c0,
                       uint64_t n, uint64_t kcount, uint64_t row0, int last);
int  kq_vroute_answer(kq_ctx *ctx, uint32_t cov_cutoff);
/* after the barrier that follows kq_vroute_answer: the owners stored their answers in this rank's answer boxes in
 * query order (coalesced peer stores); this puts them into their rows.  Then kq_vroute_rows / kq_vroute_copy_rows. */
int  kq_vroute_place(kq_ctx *ctx);
int  kq_vroute_rows(kq_ctx *ctx, kq_base **d_rows);
int  kq_vroute_copy_rows(kq_ctx *ctx, kq_base *host_dst, uint64_t n);      /* slab rows [0, n) -> host, async on the stream */
int  kq_vroute_counters(kq_ctx *ctx, kq_qv *acc, int reset);

/* ---- all GPUs of one box behind one handle (SURVEY 5.6 `--gpus N`, 8b kq_create(..., ngpus, ...)) ---------------
 * One process, one host thread per GPU while a collective step runs; shard g owns maps [g*128/G, (g+1)*128/G).
 * Same meaning as the single-GPU calls above: count_push deals the batches to the GPUs (tuples are routed to
 * their owners over peer memory inside the scatter kernel), save / load / union / stats are shard-local and the
 * results are those of one GPU (same database files, same counters).  ngpus: a power of two, 0 = all visible;
 * devices: CUDA ordinals (NULL = 0..ngpus-1; an ordinal may repeat, e.g. every shard on one GPU). */
typedef struct kq_multi kq_multi;
int  kq_multi_create(kq_multi **out, int k, int ngpus, const int *devices, uint64_t expected_distinct);
void kq_multi_destroy(kq_multi *m);
const char *kq_multi_last_error(kq_multi *m);
int  kq_multi_ngpus(kq_multi *m);
kq_ctx *kq_multi_shard(kq_multi *m, int g);
/* positions a GPU routes per count round / assembly positions per validate chunk (0 = keep; before first use) */
int  kq_multi_set_round(kq_multi *m, uint64_t round_positions, uint64_t validate_chunk);
int  kq_multi_count_push(kq_multi *m, const char *ascii, size_t len);
int  kq_multi_count_finish(kq_multi *m);
int  kq_multi_count_instances(kq_multi *m, uint64_t *n);
int  kq_multi_clear(kq_multi *m);
int  kq_multi_stats_get(kq_multi *m, kq_stats *out);
int  kq_multi_hist(kq_multi *m, uint64_t *cov, uint64_t *n, size_t *len);
int  kq_multi_export(kq_multi *m, uint64_t *keys, kq_rec32 *recs, size_t *n);
int  kq_multi_lookup_batch(kq_multi *m, const uint64_t *keys, size_t n, kq_rec32 *rec, uint8_t *found);
int  kq_multi_db_save(kq_multi *m, const char *dir);
int  kq_multi_db_load(kq_multi *m, const char *dir);
int  kq_multi_union(kq_multi *m, const char *const *dirs, int n, int ref_compatible);
int  kq_multi_validate_segment(kq_multi *m, const char *seq, uint64_t len, uint32_t cov_cutoff, kq_base *out, kq_qv *acc);

/* probe positions [0, len-k+1) of one packed device-resident segment;
 * d_out (len entries) and d_acc (kq_qv, accumulates) are device memory. */
int  kq_dev_validate_packed(kq_ctx *ctx, const uint64_t *d_bases, const uint64_t *d_invalid,
                            uint64_t len, uint32_t cov_cutoff, kq_base *d_out, kq_qv *d_acc);
/* make sure the table can take `extra` more distinct keys without growing
 * inside a timed region (also tells the count pipeline that the table is sized: it then plans
 * large batches) */
int  kq_reserve(kq_ctx *ctx, uint64_t extra_distinct);
/* kq_count_push only appends the packed batch to a device-resident pending stream; the count
 * pipeline runs once this many k-mer start positions are pending (0 = default 2^29; the next batches are copied in
 * while it runs), at
 * kq_count_finish / kq_sync, and before anything that reads the table. */
int  kq_set_flush_positions(kq_ctx *ctx, uint64_t positions);

/* process-wide tuning knobs (defaults: measured best on B200; the KQ_* environment variables of
 * the same names are read once, at first use).  Geometry: "single_slice_bits" (a table of up to 2^v slots is one
 * slice), "piece_bits" (log2 slots per slice of larger tables: 12 = one CTA's shared memory, <= 10: one warp's),
 * "coarse_bias".  Count pipeline: "pipeline" (2 L2 CAS insert, 3 two scatter levels + shared-memory insert, 4 one-pass
 * scatter), "smem_insert", "warp_insert", "cta_lock", "one_pass" (2: also small batches, almost no slack: tests),
 * "fine_part", "direct_pct", "scatter_cap", "insert_ctas", "s3_threads", "s3_ctas", "cursor_stride"; multi-GPU:
 * "route_smem" (1: the routed slab goes through level 2 + shared-memory insert instead of L2 CAS); validate:
 * "validate_ctas"; databases: "db_host", "db_ordered".  Geometry knobs apply to tables created or grown afterwards.
 * Unknown name or value out of range -> KQ_EINVAL. */
int  kq_set_tuning(const char *name, long long value);

/* counters for bench.py: kernels launched by this handle, bytes copied */
int  kq_counters(kq_ctx *ctx, uint64_t *kernel_launches, uint64_t *h2d_bytes, uint64_t *d2h_bytes);
/* CUDA-event time (ms) on the launching stream since the last call with reset=1:
 * which = 0 whole count pipeline (per batch), 1 validate/probe kernel, 2 insert kernel,
 * 3 level-1 histogram + scatter (or extract_route / kq_route_hist + kq_route_scatter), 4 level-2 histogram + scatter,
 * 5 the scatter kernel alone (level 1 / one-pass), 6 the level-2 scatter kernel alone, 7 the load / union kernels
 * (image decode + put / merge), 8 / 9 the query and the answer kernel of the routed validate.
 * Only measured when kq_set_timing(ctx,1). */
int  kq_set_timing(kq_ctx *ctx, int on);
int  kq_kernel_time_ms(kq_ctx *ctx, int which, double *ms, uint64_t *launches, int reset);

#ifdef __cplusplus
}
#endif
#endif /* KREEQ_B200_H */

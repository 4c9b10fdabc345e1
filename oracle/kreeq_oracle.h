/* kreeq_oracle.h -- CPU restatement of kreeq's k-mer hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under kreeq_b200/ (the product) may
 * include, link or dlopen this.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline leg use it, and only as the checker.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this restatement
 * against (i) the reference's own golden stdout (validateFiles/test.0-35.tst,
 * committed as tests/golden/ref_golden.json), (ii) the decoded reference
 * fixture databases testFiles/{test1,test2}.kreeq, and (iii) outputs of the
 * unmodified reference binary (oracle/_ref/kreeq) on generated inputs.
 *
 * Every function cites the reference file:line (relative to /root/reference)
 * whose algorithm it follows.
 */
#ifndef KREEQ_ORACLE_H
#define KREEQ_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* DBGkmer   include/kreeq.h:20-21  (9 bytes, 8-bit saturating record)      */
typedef struct { uint8_t fw[4], bw[4], cov; } ko_rec8;
/* DBGkmer32 include/kreeq.h:69-70  (36 bytes, 32-bit spill record)         */
typedef struct { uint32_t fw[4], bw[4], cov; } ko_rec32;
/* DBGbase   include/input.h:4-9    (validate output, 16 bytes)             */
typedef struct { uint32_t fw, bw, cov; uint8_t isFw; uint8_t pad[3]; } ko_base;
/* DBG::summary / DBG::DBstats  src/graph-builder.cpp:240-295               */
typedef struct { uint64_t total, unique, distinct, missing, edges; } ko_stats;
/* totMissingKmers / totKcount / totEdgeMissingKmers  src/kreeq.cpp:223-225 */
typedef struct { uint64_t missing, kcount, edge_missing; } ko_qv;

typedef struct ko_table ko_table;

ko_table *ko_table_new(int k);
void      ko_table_free(ko_table *t);
int       ko_table_k(const ko_table *t);

/* Kmap::hash  gfalibs/include/kmer.h:781-794.  codes[i] in 0..3. */
uint64_t ko_hash(const uint8_t *codes, int k, int *is_fw);

/* DBG::hashSequences  src/graph-builder.cpp:75-113: emit one (key, edgeBit)
 * tuple per valid k-window of one 'N'-joined batch.  keys/edges may be NULL
 * to only count.  Returns the number of tuples. */
size_t ko_extract(int k, const char *ascii, size_t len, uint64_t *keys, uint8_t *edges);

/* DBG::processBuffers  src/graph-builder.cpp:160-206: apply one tuple. */
void ko_insert_tuple(ko_table *t, uint64_t key, uint8_t edge_bits);

/* extract + insert of one batch (what hashSequences -> processBuffers do). */
size_t ko_count_batch(ko_table *t, const char *ascii, size_t len);

/* raw record access, used to load decoded .kreeq databases */
void ko_put8(ko_table *t, uint64_t key, const ko_rec8 *r);
void ko_put32(ko_table *t, uint64_t key, const ko_rec32 *r);

/* DBG::reloadMap32  src/graph-builder.cpp:225-238: after a map is dumped, every
 * 32-bit key is (re)inserted in the 8-bit map as a cov=255 placeholder. */
void ko_finalize(ko_table *t);

/* number of keys present in the 8-bit map / the 32-bit map */
size_t ko_size8(const ko_table *t);
size_t ko_size32(const ko_table *t);
/* dump the raw maps sorted by key (returns n written, at most cap) */
size_t ko_dump8(const ko_table *t, uint64_t *keys, ko_rec8 *recs, size_t cap);
size_t ko_dump32(const ko_table *t, uint64_t *keys, ko_rec32 *recs, size_t cap);

/* lookup idiom of src/kreeq.cpp:150-166: 8-bit map, then 32-bit map when
 * cov==255.  Returns 1 found, 0 not found, -1 inconsistent (placeholder
 * without 32-bit record; the reference exit()s there). */
int ko_lookup(const ko_table *t, uint64_t key, ko_rec32 *out);

/* logical records: every key found by ko_lookup, sorted by key */
size_t ko_export(const ko_table *t, uint64_t *keys, ko_rec32 *recs, size_t cap);

/* DBG::summary + DBG::DBstats  src/graph-builder.cpp:240-295 (incl. the
 * "Total edges" precedence quirk and the double-precision "Missing kmers"). */
void ko_stats_get(const ko_table *t, ko_stats *out);
/* coverage histogram (finalHistogram, kmer.h:895-899) sorted by coverage */
size_t ko_hist(const ko_table *t, uint64_t *cov, uint64_t *n, size_t cap);

/* DBG::kunion + Kmap::mergeMaps + DBG::mergeSubMaps
 * src/graph-builder.cpp:297-432, gfalibs/include/kmer.h:666-695.
 * dbs[0] is "DB0"; order matters exactly as in the reference. */
ko_table *ko_union(ko_table *const *dbs, int n);

/* DBG::evaluateSegment  src/kreeq.cpp:110-229 for one segment; out has len
 * entries (zero-initialised by the callee); acc accumulates.  Returns 0, or
 * -1 for the reference's "int32 map missing 255 value" exit. */
int ko_validate_segment(const ko_table *t, const char *seq, uint64_t len,
                        uint32_t cov_cutoff, ko_base *out, ko_qv *acc);

/* errorRate  src/kreeq.cpp:36-40 and QV = -10*log10(err)  :87,96 */
double ko_error_rate(uint64_t missing, uint64_t total, int k);
double ko_qv_value(uint64_t missing, uint64_t total, int k);

#ifdef __cplusplus
}
#endif
#endif

/* kreeq_oracle.c -- CPU restatement of kreeq's k-mer hot path (count, union,
 * validate, stats).  TEST INFRASTRUCTURE ONLY: see kreeq_oracle.h.
 *
 * Parity status: PINNED against the reference's golden vectors and against
 * the unmodified reference binary (tests/test_oracle_golden.py).
 *
 * This is a restatement, not a copy: one flat open-addressing table stands in
 * for the reference's 128 x 256 phmap submaps (the partitioning does not
 * change any result), and every update rule below cites the reference lines
 * it follows (paths relative to /root/reference).
 */
#include "kreeq_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define KO_LARGEST 4294967295u /* include/kreeq.h:68 */

/* ---- table ------------------------------------------------------------- */

typedef struct {
    uint64_t key;
    ko_rec32 r32;   /* entry of maps32[key%128]  (valid iff has32) */
    ko_rec8  r8;    /* entry of maps[key%128]    (valid iff has8)  */
    uint8_t  has8, has32, used;
} ko_slot;

struct ko_table {
    int      k;
    size_t   cap;   /* power of two */
    size_t   n;     /* used slots */
    ko_slot *s;
};

static uint64_t mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
    x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
    x ^= x >> 33; return x;
}

ko_table *ko_table_new(int k) {
    ko_table *t = (ko_table *)calloc(1, sizeof *t);
    t->k = k; t->cap = 1024; t->n = 0;
    t->s = (ko_slot *)calloc(t->cap, sizeof(ko_slot));
    return t;
}

void ko_table_free(ko_table *t) { if (t) { free(t->s); free(t); } }
int  ko_table_k(const ko_table *t) { return t->k; }

static ko_slot *find_slot(const ko_table *t, uint64_t key) {
    size_t i = mix64(key) & (t->cap - 1);
    while (t->s[i].used) {
        if (t->s[i].key == key) return &t->s[i];
        i = (i + 1) & (t->cap - 1);
    }
    return NULL;
}

static void grow(ko_table *t) {
    size_t ocap = t->cap; ko_slot *os = t->s;
    t->cap = ocap * 2;
    t->s = (ko_slot *)calloc(t->cap, sizeof(ko_slot));
    for (size_t j = 0; j < ocap; ++j) {
        if (!os[j].used) continue;
        size_t i = mix64(os[j].key) & (t->cap - 1);
        while (t->s[i].used) i = (i + 1) & (t->cap - 1);
        t->s[i] = os[j];
    }
    free(os);
}

/* find or create the slot for key (neither map entry is created here) */
static ko_slot *get_slot(ko_table *t, uint64_t key) {
    if ((t->n + 1) * 10 > t->cap * 6) grow(t);
    size_t i = mix64(key) & (t->cap - 1);
    while (t->s[i].used) {
        if (t->s[i].key == key) return &t->s[i];
        i = (i + 1) & (t->cap - 1);
    }
    memset(&t->s[i], 0, sizeof(ko_slot));
    t->s[i].used = 1; t->s[i].key = key; t->n++;
    return &t->s[i];
}

/* ---- key and edge extraction ------------------------------------------- */

/* ctoi  gfalibs/include/kmer.h:97-114: A/a 0, C/c 1, G/g 2, T/t 3, else 4 */
static uint8_t ctoi(unsigned char c) {
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': return 3;
        default: return 4;
    }
}

/* Kmap::hash  gfalibs/include/kmer.h:781-794.  First base is the least
 * significant base-4 digit of fw; rv is the reverse complement read the same
 * way; the key is the smaller, isFw = fw < rv (a tie counts as not-forward). */
uint64_t ko_hash(const uint8_t *codes, int k, int *is_fw) {
    uint64_t fw = 0, rv = 0, p = 1;
    for (int c = 0; c < k; ++c) { fw += (uint64_t)codes[c] * p; p *= 4; }
    p = 1;
    for (int c = k - 1; c >= 0; --c) { rv += (uint64_t)(3 - codes[c]) * p; p *= 4; }
    if (is_fw) *is_fw = fw < rv;
    return fw < rv ? fw : rv;
}

/* edgeBit::assign  include/kreeq.h:10-12 */
static uint8_t edge_assign(uint8_t bits, int e) { return (uint8_t)(bits | (1u << (7 - e))); }
/* edgeBit::read    include/kreeq.h:14-16 */
static int edge_read(uint8_t bits, int e) { return (bits >> (7 - e)) & 1; }

/* DBG::hashSequences  src/graph-builder.cpp:62-113.  The reference walks the
 * batch, restarting after every non-ACGT byte (:77-91); that is the same as
 * visiting every k-window made only of ACGT.  Edge byte (:98-112): the base
 * after the window (ascii[p+k]; at the very end this is the std::string's
 * NUL terminator, which maps to 4) and the base before it (p>0). */
size_t ko_extract(int k, const char *ascii, size_t len, uint64_t *keys, uint8_t *edges) {
    if (len < (size_t)k) return 0;                 /* :60-64 */
    size_t n = 0, run = 0;
    uint8_t *codes = (uint8_t *)malloc(len + 1);
    for (size_t i = 0; i < len; ++i) codes[i] = ctoi((unsigned char)ascii[i]);
    codes[len] = 4;                                /* c_str() terminator */
    for (size_t i = 0; i < len; ++i) {
        run = codes[i] <= 3 ? run + 1 : 0;
        if (run < (size_t)k) continue;
        size_t p = i + 1 - (size_t)k;              /* window start */
        int is_fw;
        uint64_t key = ko_hash(codes + p, k, &is_fw);
        uint8_t nxt = codes[p + k];
        uint8_t prv = p > 0 ? codes[p - 1] : 4;
        uint8_t e = 0;
        if (is_fw) {
            if (nxt <= 3) e = edge_assign(e, nxt);
            if (prv <= 3) e = edge_assign(e, 4 + prv);
        } else {
            if (prv <= 3) e = edge_assign(e, 3 - prv);
            if (nxt <= 3) e = edge_assign(e, 4 + 3 - nxt);
        }
        if (keys)  keys[n] = key;
        if (edges) edges[n] = e;
        ++n;
    }
    free(codes);
    return n;
}

/* ---- count: record update ---------------------------------------------- */

static void widen(ko_rec32 *d, const ko_rec8 *s) {   /* DBGkmer32(const DBGkmer&) include/kreeq.h:74-78 */
    for (int w = 0; w < 4; ++w) { d->fw[w] = s->fw[w]; d->bw[w] = s->bw[w]; }
    d->cov = s->cov;
}

/* DBG::processBuffers  src/graph-builder.cpp:160-206, one 9-byte tuple. */
void ko_insert_tuple(ko_table *t, uint64_t key, uint8_t e) {
    ko_slot *s = get_slot(t, key);
    if (!s->has8) { s->has8 = 1; memset(&s->r8, 0, sizeof s->r8); }   /* map[hash] :163 */
    ko_rec8 *r = &s->r8;
    int overflow = r->cov >= 254;                                     /* :164 */
    for (int w = 0; w < 4; ++w)                                       /* :166-172 */
        if (r->fw[w] + 1 == 255 || r->bw[w] + 1 == 255) { overflow = 1; break; }
    if (!overflow) {                                                  /* :174-182 */
        for (int w = 0; w < 4; ++w) {
            r->fw[w] = (uint8_t)(r->fw[w] + edge_read(e, w));
            r->bw[w] = (uint8_t)(r->bw[w] + edge_read(e, 4 + w));
        }
        ++r->cov;
        return;
    }
    if (!s->has32) { s->has32 = 1; memset(&s->r32, 0, sizeof s->r32); } /* map32[hash] :186 */
    ko_rec32 *q = &s->r32;
    if (q->cov == 0) { widen(q, r); r->cov = 255; }                   /* :188-192 */
    for (int w = 0; w < 4; ++w) {                                     /* :194-200 */
        if (KO_LARGEST - q->fw[w] >= (uint32_t)edge_read(e, w))     q->fw[w] += (uint32_t)edge_read(e, w);
        if (KO_LARGEST - q->bw[w] >= (uint32_t)edge_read(e, 4 + w)) q->bw[w] += (uint32_t)edge_read(e, 4 + w);
    }
    if (q->cov < KO_LARGEST) ++q->cov;                                /* :201-202 */
}

size_t ko_count_batch(ko_table *t, const char *ascii, size_t len) {
    size_t n = ko_extract(t->k, ascii, len, NULL, NULL);
    if (n == 0) return 0;
    uint64_t *keys = (uint64_t *)malloc(n * sizeof(uint64_t));
    uint8_t *edges = (uint8_t *)malloc(n);
    ko_extract(t->k, ascii, len, keys, edges);
    for (size_t i = 0; i < n; ++i) ko_insert_tuple(t, keys[i], edges[i]);
    free(keys); free(edges);
    return n;
}

void ko_put8(ko_table *t, uint64_t key, const ko_rec8 *r) {
    ko_slot *s = get_slot(t, key); s->has8 = 1; s->r8 = *r;
}
void ko_put32(ko_table *t, uint64_t key, const ko_rec32 *r) {
    ko_slot *s = get_slot(t, key); s->has32 = 1; s->r32 = *r;
}

/* DBG::reloadMap32  src/graph-builder.cpp:225-238 (insert never overwrites) */
void ko_finalize(ko_table *t) {
    for (size_t i = 0; i < t->cap; ++i) {
        ko_slot *s = &t->s[i];
        if (s->used && s->has32 && !s->has8) {
            s->has8 = 1; memset(&s->r8, 0, sizeof s->r8); s->r8.cov = 255;
        }
    }
}

/* ---- raw dumps, lookup, export ------------------------------------------ */

static int cmp_slot_ptr(const void *a, const void *b) {
    const ko_slot *x = *(const ko_slot *const *)a, *y = *(const ko_slot *const *)b;
    return x->key < y->key ? -1 : x->key > y->key;
}

/* all used slots sorted by key; caller frees */
static const ko_slot **sorted_slots(const ko_table *t, size_t *n_out) {
    const ko_slot **v = (const ko_slot **)malloc((t->n + 1) * sizeof *v);
    size_t n = 0;
    for (size_t i = 0; i < t->cap; ++i) if (t->s[i].used) v[n++] = &t->s[i];
    qsort(v, n, sizeof *v, cmp_slot_ptr);
    *n_out = n;
    return v;
}

size_t ko_size8(const ko_table *t) {
    size_t n = 0; for (size_t i = 0; i < t->cap; ++i) n += t->s[i].used && t->s[i].has8; return n;
}
size_t ko_size32(const ko_table *t) {
    size_t n = 0; for (size_t i = 0; i < t->cap; ++i) n += t->s[i].used && t->s[i].has32; return n;
}

size_t ko_dump8(const ko_table *t, uint64_t *keys, ko_rec8 *recs, size_t cap) {
    size_t n, m = 0; const ko_slot **v = sorted_slots(t, &n);
    for (size_t i = 0; i < n && m < cap; ++i)
        if (v[i]->has8) { keys[m] = v[i]->key; recs[m] = v[i]->r8; ++m; }
    free(v); return m;
}
size_t ko_dump32(const ko_table *t, uint64_t *keys, ko_rec32 *recs, size_t cap) {
    size_t n, m = 0; const ko_slot **v = sorted_slots(t, &n);
    for (size_t i = 0; i < n && m < cap; ++i)
        if (v[i]->has32) { keys[m] = v[i]->key; recs[m] = v[i]->r32; ++m; }
    free(v); return m;
}

/* lookup idiom  src/kreeq.cpp:150-166 (also subgraph.cpp:235-245 etc.) */
static int slot_lookup(const ko_slot *s, ko_rec32 *out) {
    if (!s || !s->has8) return 0;                  /* map->find(key) == end() */
    if (s->r8.cov == 255) {                        /* :158 */
        if (!s->has32) return -1;                  /* :161-164 exit(EXIT_FAILURE) */
        *out = s->r32; return 1;
    }
    widen(out, &s->r8); return 1;
}

int ko_lookup(const ko_table *t, uint64_t key, ko_rec32 *out) {
    return slot_lookup(find_slot(t, key), out);
}

size_t ko_export(const ko_table *t, uint64_t *keys, ko_rec32 *recs, size_t cap) {
    size_t n, m = 0; const ko_slot **v = sorted_slots(t, &n);
    for (size_t i = 0; i < n && m < cap; ++i) {
        ko_rec32 r;
        if (slot_lookup(v[i], &r) == 1) { keys[m] = v[i]->key; recs[m] = r; ++m; }
    }
    free(v); return m;
}

/* ---- stats -------------------------------------------------------------- */

/* `edgeCount += fw[w] > 0 ? 1 : 0 + bw[w] > 0 ? 1 : 0;`
 * src/graph-builder.cpp:253-254,262-263 parses as fw>0 ? 1 : (bw>0 ? 1 : 0). */
static uint64_t edge_quirk(uint32_t fw, uint32_t bw) { return fw > 0 ? 1 : (0 + bw > 0 ? 1 : 0); }

/* DBG::summary  src/graph-builder.cpp:240-282 over all maps, then
 * DBG::DBstats :284-295.  8-bit records with cov==255 are skipped (:246-247);
 * every 32-bit record counts (:259-266). */
void ko_stats_get(const ko_table *t, ko_stats *o) {
    memset(o, 0, sizeof *o);
    for (size_t i = 0; i < t->cap; ++i) {
        const ko_slot *s = &t->s[i];
        if (!s->used) continue;
        if (s->has8 && s->r8.cov != 255) {
            if (s->r8.cov == 1) ++o->unique;
            for (int w = 0; w < 4; ++w) o->edges += edge_quirk(s->r8.fw[w], s->r8.bw[w]);
            ++o->distinct; o->total += s->r8.cov;          /* tot += cov*hist[cov] :277-278 */
        }
        if (s->has32) {
            for (int w = 0; w < 4; ++w) o->edges += edge_quirk(s->r32.fw[w], s->r32.bw[w]);
            ++o->distinct; o->total += s->r32.cov;
        }
    }
    o->missing = (uint64_t)(pow(4, t->k) - o->distinct);   /* :286, evaluated in double */
}

static int cmp_u64(const void *a, const void *b) {
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b; return x < y ? -1 : x > y;
}

size_t ko_hist(const ko_table *t, uint64_t *cov, uint64_t *cnt, size_t cap) {
    uint64_t *all = (uint64_t *)malloc((2 * t->n + 1) * sizeof(uint64_t)); size_t n = 0;
    for (size_t i = 0; i < t->cap; ++i) {
        const ko_slot *s = &t->s[i];
        if (!s->used) continue;
        if (s->has8 && s->r8.cov != 255) all[n++] = s->r8.cov;
        if (s->has32) all[n++] = s->r32.cov;
    }
    qsort(all, n, sizeof(uint64_t), cmp_u64);
    size_t m = 0;
    for (size_t i = 0; i < n;) {
        size_t j = i; while (j < n && all[j] == all[i]) ++j;
        if (m < cap) { cov[m] = all[i]; cnt[m] = j - i; }
        ++m; i = j;
    }
    free(all); return m;
}

/* ---- union -------------------------------------------------------------- */

static uint32_t sat_add(uint32_t a, uint32_t b) { return KO_LARGEST - a >= b ? a + b : KO_LARGEST; }

static void sat_add_rec32(ko_rec32 *d, const ko_rec32 *s) {   /* src/graph-builder.cpp:315-330 */
    for (int w = 0; w < 4; ++w) { d->fw[w] = sat_add(d->fw[w], s->fw[w]); d->bw[w] = sat_add(d->bw[w], s->bw[w]); }
    d->cov = sat_add(d->cov, s->cov);
}
static void sat_add_rec8(ko_rec32 *d, const ko_rec8 *s) {     /* src/graph-builder.cpp:410-427 */
    for (int w = 0; w < 4; ++w) { d->fw[w] = sat_add(d->fw[w], s->fw[w]); d->bw[w] = sat_add(d->bw[w], s->bw[w]); }
    d->cov = sat_add(d->cov, s->cov);
}

/* DBG::mergeSubMaps  src/graph-builder.cpp:353-432 for one (key, a) of a
 * later database; M is the merged table (its 8-bit side is "submap2", its
 * 32-bit side is maps32[m]). */
static void merge_one(ko_table *M, uint64_t key, const ko_rec8 *a) {
    if (a->cov == 255) return;                                /* :367-368 */
    int overflow = 0;
    ko_slot *s = get_slot(M, key);
    if (s->has32) {                                           /* :370-372 */
        overflow = 1;
    } else if (!s->has8) {                                    /* :375-377 insert */
        s->has8 = 1; s->r8 = *a;
    } else {
        ko_rec8 *b = &s->r8;
        if (255 - b->cov <= a->cov) overflow = 1;             /* :382-383 */
        for (int w = 0; w < 4; ++w)                           /* :385-391 */
            if (255 - b->fw[w] <= a->fw[w] || 255 - b->bw[w] <= a->bw[w]) { overflow = 1; break; }
        if (!overflow) {                                      /* :393-400 */
            for (int w = 0; w < 4; ++w) { b->fw[w] = (uint8_t)(b->fw[w] + a->fw[w]); b->bw[w] = (uint8_t)(b->bw[w] + a->bw[w]); }
            b->cov = (uint8_t)(b->cov + a->cov);
        }
    }
    if (overflow) {                                           /* :404-428 */
        if (!s->has32) { s->has32 = 1; memset(&s->r32, 0, sizeof s->r32); }
        if (s->r32.cov == 0) { widen(&s->r32, &s->r8); s->r8.cov = 255; }
        sat_add_rec8(&s->r32, a);
    }
}

/* DBG::kunion  src/graph-builder.cpp:297-351 + Kmap::mergeMaps kmer.h:666-695 */
ko_table *ko_union(ko_table *const *dbs, int n) {
    ko_table *M = ko_table_new(dbs[0]->k);
    /* 1. saturating sum of every database's high-copy map (:301-336) */
    for (int i = 0; i < n; ++i)
        for (size_t j = 0; j < dbs[i]->cap; ++j) {
            const ko_slot *s = &dbs[i]->s[j];
            if (!s->used || !s->has32) continue;
            ko_slot *d = get_slot(M, s->key);
            if (!d->has32) { d->has32 = 1; memset(&d->r32, 0, sizeof d->r32); }
            sat_add_rec32(&d->r32, &s->r32);
        }
    /* 2. DB0's 8-bit maps are loaded raw (kmer.h:668-672) */
    for (size_t j = 0; j < dbs[0]->cap; ++j) {
        const ko_slot *s = &dbs[0]->s[j];
        if (!s->used || !s->has8) continue;
        ko_slot *d = get_slot(M, s->key);
        d->has8 = 1; d->r8 = s->r8;
    }
    /* 3. every later database is merged in order (kmer.h:674-686) */
    for (int i = 1; i < n; ++i)
        for (size_t j = 0; j < dbs[i]->cap; ++j) {
            const ko_slot *s = &dbs[i]->s[j];
            if (s->used && s->has8) merge_one(M, s->key, &s->r8);
        }
    return M;
}

/* ---- validate ----------------------------------------------------------- */

/* DBG::evaluateSegment  src/kreeq.cpp:110-229, single map range [0,128).
 * Segments never contain N/n/X/x (gfalibs/src/gfa.cpp:170-203).  Any other
 * non-ACGT byte makes the reference index fw[4]/fw[-1] (undefined); here such
 * a window is "not found" and such a neighbour gives no edge. */
int ko_validate_segment(const ko_table *t, const char *seq, uint64_t len,
                        uint32_t cutoff, ko_base *out, ko_qv *acc) {
    const int k = t->k;
    memset(out, 0, (size_t)len * sizeof(ko_base));
    if (len < (uint64_t)k) return 0;                          /* :123-124 */
    uint64_t kcount = len - (uint64_t)k + 1;
    uint8_t *str = (uint8_t *)malloc((size_t)len);
    for (uint64_t i = 0; i < len; ++i) str[i] = ctoi((unsigned char)seq[i]);   /* :131-132 */
    uint64_t bad = 0;                                         /* non-ACGT bytes in the current window */
    for (int i = 0; i < k - 1; ++i) bad += str[i] > 3;
    int rc = 0;
    for (uint64_t c = 0; c < kcount; ++c) {
        bad += str[c + k - 1] > 3;
        if (c > 0) bad -= str[c - 1] > 3;
        ko_rec32 kh; memset(&kh, 0, sizeof kh);               /* DBGkmer32 khmer :154 */
        int is_fw = 0, found = 0;
        if (!bad) {
            uint64_t key = ko_hash(str + c, k, &is_fw);       /* :145 */
            found = ko_lookup(t, key, &kh);                   /* :152-166 */
            if (found < 0) { rc = -1; break; }
        }
        if (found) { out[c].cov = kh.cov; out[c].isFw = (uint8_t)is_fw; }   /* :168-169 */
        if (out[c].cov == 0) ++acc->missing;                  /* :172-173 */
        else if (out[c].cov < cutoff) ++acc->missing;         /* :174-175 */
        else {
            int no_left = 0, no_right = 0;
            uint8_t nxt = c < kcount - 1 ? str[c + k] : 4, prv = c > 0 ? str[c - 1] : 4;
            if (out[c].isFw) {                                /* :178-192 */
                if (c < kcount - 1) { if (nxt <= 3 && kh.fw[nxt] != 0) out[c].fw = kh.fw[nxt]; else no_right = 1; }
                if (c > 0)          { if (prv <= 3 && kh.bw[prv] != 0) out[c].bw = kh.bw[prv]; else no_left = 1; }
            } else {                                          /* :194-210 */
                if (c > 0)          { if (prv <= 3 && kh.fw[3 - prv] != 0) out[c].fw = kh.fw[3 - prv]; else no_left = 1; }
                if (c < kcount - 1) { if (nxt <= 3 && kh.bw[3 - nxt] != 0) out[c].bw = kh.bw[3 - nxt]; else no_right = 1; }
            }
            if (no_left && no_right) ++acc->edge_missing;     /* :211-214 */
        }
        ++acc->kcount;                                        /* :216 */
    }
    free(str);
    return rc;
}

/* errorRate  src/kreeq.cpp:36-40 */
double ko_error_rate(uint64_t missing, uint64_t total, int k) {
    return 1 - pow(1 - (double)missing / total, (double)1 / k);
}
/* src/kreeq.cpp:87,96 */
double ko_qv_value(uint64_t missing, uint64_t total, int k) {
    return -10 * log10(ko_error_rate(missing, total, k));
}

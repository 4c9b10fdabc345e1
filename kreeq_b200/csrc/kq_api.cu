// kq_api.cu -- implementation of the C-ABI in include/kreeq_b200.h.
//
// One kq_ctx = one GPU = one shard of the key space.  kq_count_push packs a host batch to 2 bit on the
// host pool into pinned staging and copies it (cudaMemcpyAsync, ring of 3) to the END of a device-
// resident "pending reads" stream; the stream is turned into table updates by the count pipeline
// (histogram -> scan -> scatter into slice order -> L2-resident insert) once enough k-mers are
// pending to pay for one sweep over the table, or at kq_count_finish.  Everything is enqueued on one
// stream (the caller's after kq_set_stream).  There is no CPU fallback anywhere in this file.
#include "../../include/kreeq_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <sys/stat.h>

#include "kq_host.h"
#include "kq_kernels.h"

using namespace kq;

static_assert(sizeof(kq_rec32) == sizeof(Rec32), "kq_rec32 layout");
static_assert(sizeof(kq_base) == sizeof(Base) && sizeof(kq_base) == 16, "kq_base layout");

namespace {

constexpr int NBUF = 3;                          // staging ring depth
constexpr uint64_t CHUNK = 16ull << 20;          // bases per staged chunk (multiple of 64)
constexpr double MAX_LOAD = 0.80;                // never plan a batch that could push T beyond this
constexpr double TARGET_LOAD = 0.50;             // load right after a grow / reserve
constexpr uint64_t MIN_SLOTS = 1ull << 12;
constexpr uint64_t MIN_SCAP = 1ull << 12;
constexpr uint64_t OVF_CAP = 4ull << 20;         // tuples kept when their slice is full
constexpr uint64_t DEFAULT_FLUSH = 1ull << 29;   // pending k-mer start positions that trigger the pipeline
constexpr int NTIMERS = 10;                       // 0 count pipeline, 1 validate, 2 insert kernel, 3 level-1 hist + scatter, 4 level-2 hist + scatter,
                                                 // 5 scatter kernel alone (level 1 / one-pass), 6 level-2 scatter kernel alone, 7 load / union kernels

thread_local std::string g_create_error;

struct Stage {                                   // one staging slot (count and validate share them)
    uint64_t *h_bases = nullptr, *h_inv = nullptr;    // pinned
    uint64_t *d_bases = nullptr, *d_inv = nullptr;    // validate only
    Base *h_out = nullptr, *d_out = nullptr;          // validate results (pinned / device)
    size_t words = 0, dwords = 0, out_n = 0;
    cudaEvent_t done = nullptr, kernel_done = nullptr; bool busy = false;
};

struct TimedRange { cudaEvent_t a, b; int which; };

}  // namespace

struct kq_ctx {
    int k = 0, device = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    Slot *T = nullptr; uint64_t nslices = 0; int slice_bits = 0, cshift = 0;
    SideSlot *R = nullptr; uint64_t rcap = 0;
    SideSlot *M = nullptr; uint64_t mcap = 0;
    unsigned long long *d_occ = nullptr;         // [0] T, [1] R, [2] M
    unsigned long long *d_inst = nullptr;        // k-mer instances inserted from packed sequences ...
    uint64_t inst_tuples = 0;                    // ... and from tuple sources (kq_dev_insert_tuples)
    unsigned long long *d_acc = nullptr;         // validate counters (3)
    int *d_err = nullptr;
    uint64_t occ_known = 0;                      // T occupancy at the last synchronisation
    uint64_t occ_bound = 0;                      // upper bound of it including batches in flight
    uint64_t big_seen = 0;                       // records fed by load / union that could own a side-table entry
    uint64_t r_bound = 0;                        // upper bound of R's occupancy (a key needs 128 instances to get there)
    bool optimistic = false;                     // the caller sized the table (kq_create / kq_reserve)
    int rank = 0, nranks = 1;
    bool loaded_db = false;                      // table holds literal database entries (load/union)
    // pending reads (device-resident packed stream)
    uint64_t *d_pb = nullptr, *d_pi = nullptr; uint64_t pend_cap = 0, pend_len = 0, flush_positions = DEFAULT_FLUSH;
    // ... double-buffered: batches are copied in on copy_stream while the pipeline reads the other buffer
    uint64_t *d_pb_alt = nullptr, *d_pi_alt = nullptr; uint64_t pend_cap_alt = 0;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copied = nullptr;                 // last host-to-device copy into the current buffer
    cudaEvent_t ev_read = nullptr, ev_read_alt = nullptr;   // pipeline done with the current / the other buffer
    bool read_armed = false, read_armed_alt = false;
    // pipeline scratch
    uint64_t *d_tuples = nullptr; uint64_t tup_cap = 0;          // tuples (not words)
    unsigned long long *d_hist = nullptr, *d_cursor = nullptr; uint64_t hist_cap = 0;
    unsigned long long *d_total = nullptr, *d_nsrc = nullptr, *d_ticket = nullptr;
    unsigned long long *d_cend = nullptr, *d_rstat = nullptr;     // one-pass level-1 scatter: region ends, statistics (Regions)
    unsigned long long *d_hot = nullptr;                          // slices left to hot_slices_kernel (HotList)
    uint64_t *d_ovf = nullptr; unsigned long long *d_ovf_n = nullptr; bool ovf_pending = false;
    uint64_t *ovf_list = nullptr; uint64_t ovf_list_cap = 0;    // where the last batch put the tuples of full slices
    // level 2 of the pipeline: tuples in slice order, per-slice counts / offsets, tasks
    uint64_t *d_tuples2 = nullptr; uint64_t tup2_cap = 0;
    unsigned long long *d_cstart = nullptr, *d_fhist = nullptr, *d_fcursor = nullptr, *d_fstart = nullptr; uint64_t fine_cap = 0;
    FineTask *d_tasks = nullptr; uint64_t task_cap = 0; unsigned *d_ntasks = nullptr, *d_task_first = nullptr;
    unsigned *d_hist2 = nullptr; uint64_t hist2_cap = 0;
    // one-pass scatter: cursor per slice, big list for the tuples of slices that are full
    unsigned *d_scur = nullptr; uint64_t scur_cap = 0;
    uint64_t *d_ovf2 = nullptr; uint64_t ovf2_cap = 0;
    // routed (multi-GPU) pipeline: receive slab the owners' tuples are written into by every rank
    uint64_t *d_slab = nullptr; uint64_t slab_cap = 0;          // tuples
    RouteMap route{}; bool route_ready = false;
    uint64_t route_extent = 0; unsigned route_task_bound = 0, route_ns_max = 0;   // tasks over the slab of the last kq_route_scatter (0 = none)
    std::vector<void *> ipc_opened;
    unsigned long long *d_rcursor = nullptr, *d_rback = nullptr, *d_rtotals = nullptr; uint64_t route_real = 0;
    // peer tables (multi-GPU validate): views of every rank's shard, mapped into this process
    TableView *d_peers = nullptr; std::vector<void *> peer_opened; bool exported = false;
    // routed validate: inbox the peers' queries arrive in, row slab their answers arrive in, local cursors / counters
    unsigned char *d_vinbox = nullptr; uint64_t vcap = 0; Base *d_vrows = nullptr;
    unsigned long long *d_vcursor = nullptr, *d_vacc = nullptr; VRoute vroute{}; bool vroute_ready = false;
    uint32_t *d_out4[8] = {}; kq_wide_row *d_wide = nullptr; unsigned long long *d_wide_n = nullptr;   // compact validate rows
    unsigned *d_vpos = nullptr;                                       // rows of the queries sent in the current round (VRoute::poslog)
    bool hold_pending = false;                   // kq_multi routes the pending reads itself: nothing else may flush them
    Stage st[NBUF]; int next_stage = 0;
    uint64_t launches = 0, h2d = 0, d2h = 0;
    bool timing = false; std::vector<TimedRange> ranges; std::vector<cudaEvent_t> ev_pool;
    double time_ms[NTIMERS] = {}; uint64_t time_n[NTIMERS] = {};
    std::string err;
};

namespace {

int fail(kq_ctx *c, int code, const std::string &msg) { if (c) c->err = msg; else g_create_error = msg; return code; }

#define CU(call)                                                                                     \
    do { cudaError_t e_ = (call);                                                                   \
         if (e_ != cudaSuccess) return fail(c, e_ == cudaErrorMemoryAllocation ? KQ_ENOMEM : KQ_ECUDA, \
                                            std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

inline uint64_t table_slots(const kq_ctx *c) { return c->nslices << c->slice_bits; }

TableView view(const kq_ctx *c) {
    TableView tv;
    tv.T = c->T; tv.nslices = c->nslices; tv.slice_bits = c->slice_bits; tv.cshift = c->cshift;
    tv.R = c->R; tv.rcap = c->rcap; tv.M = c->M; tv.mcap = c->mcap;
    tv.occ = c->d_occ; tv.err = c->d_err; tv.k = c->k;
    tv.shard_rank = c->nranks > 1 ? c->rank : -1;
    int shift = 0; while ((128 >> shift) > c->nranks) ++shift;       // 128 >> shift == nranks
    tv.shard_shift = shift;                                          // owner = (key & 127) >> log2(128/nranks)
    return tv;
}

// check the sticky device error word (requires a synchronised stream)
int check_device_error(kq_ctx *c) {
    int e = 0;
    CU(cudaMemcpyAsync(&e, c->d_err, sizeof e, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (!e) return KQ_OK;
    CU(cudaMemsetAsync(c->d_err, 0, sizeof(int), c->stream));
    if (e & 4) return fail(c, KQ_ECORRUPT, "Error: int32 map missing 255 value from int8 map");
    if (e & 8) return fail(c, KQ_ENOMEM, "route buffer overflow (cap_per_rank too small)");
    if (e & 2) return fail(c, KQ_ENOMEM, "high-copy side table full");
    return fail(c, KQ_ENOMEM, "k-mer table full");
}

int read_occ(kq_ctx *c, uint64_t occ[3]) {
    unsigned long long h[3];
    CU(cudaMemcpyAsync(h, c->d_occ, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    occ[0] = h[0]; occ[1] = h[1]; occ[2] = h[2];
    c->occ_known = h[0];
    return KQ_OK;
}

// geometry of a table with at least `need` slots: one slice while it fits 2^21 slots (32 MiB, L2 resident),
// else slices of 2^12 slots (64 KiB: what one CTA of the shared-memory insert holds), grouped 2^cshift at a
// time into the bins of the level-1 scatter (about sqrt(nslices) bins at either level)
void table_geometry(uint64_t need, int tw, uint64_t &nslices, int &slice_bits, int &cshift) {
    need = std::max(need, MIN_SLOTS);
    const int max_bits = tuning().single_slice_bits;
    int b = std::min(12, max_bits); while (b < max_bits && (1ull << b) < need) ++b;
    cshift = 0;
    if ((1ull << b) >= need) { nslices = 1; slice_bits = b; return; }
    // smallest slices (from the piece_bits knob up to 2^12 slots) whose coarse bins fit the level-1 scatter and whose
    // slices per coarse bin fit the level-2 scatter
    for (slice_bits = tuning().piece_bits;; ++slice_bits) {
        nslices = (need + (1ull << slice_bits) - 1) >> slice_bits;
        int lg = 0; while ((2ull << lg) <= nslices) ++lg;            // floor(log2 nslices)
        cshift = lg / 2;
        while ((((nslices - 1) >> cshift) + 1) > 3000) ++cshift;      // bins of the level-1 scatter must fit shared memory
        // fewer, larger coarse bins while the level-2 scatter still has room for them (knob; measured in DESIGN.md)
        for (int b2 = 0; b2 < tuning().coarse_bias && fine_pick_cap(2u << cshift, tw) != 0 && (1ull << (cshift + 1)) < nslices; ++b2) ++cshift;
        if (slice_bits >= 12 || fine_pick_cap(1u << cshift, tw) != 0) break;
    }
}

int grow_table(kq_ctx *c, uint64_t need) {
    if (c->exported) return fail(c, KQ_ESTATE, "the table is mapped by peers (kq_peer_export): call kq_peer_disconnect first");
    Slot *old = c->T; const uint64_t old_slots = table_slots(c);
    uint64_t ns; int sb, cs; table_geometry(need, tuple_words(c->k), ns, sb, cs);
    Slot *nt = nullptr;
    CU(cudaMalloc(&nt, (ns << sb) * sizeof(Slot)));
    CU(cudaMemsetAsync(nt, 0, (ns << sb) * sizeof(Slot), c->stream));
    c->T = nt; c->nslices = ns; c->slice_bits = sb; c->cshift = cs;
    if (old) {
        CU(launch_rehash(c->stream, view(c), old, old_slots, &c->launches));
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaFree(old));
    }
    return KQ_OK;
}

int grow_side(kq_ctx *c, int which, uint64_t new_scap) {
    if (c->exported) return fail(c, KQ_ESTATE, "the table is mapped by peers (kq_peer_export): call kq_peer_disconnect first");
    SideSlot *&S = which == 1 ? c->R : c->M; uint64_t &cap = which == 1 ? c->rcap : c->mcap;
    SideSlot *old = S; const uint64_t old_cap = cap;
    SideSlot *ns = nullptr;
    CU(cudaMalloc(&ns, new_scap * sizeof(SideSlot)));
    CU(cudaMemsetAsync(ns, 0, new_scap * sizeof(SideSlot), c->stream));
    S = ns; cap = new_scap;
    if (old) {
        CU(cudaMemsetAsync(c->d_occ + which, 0, sizeof(unsigned long long), c->stream));   // re-counted by the rehash
        CU(launch_side_rehash(c->stream, view(c), which, old, old_cap, &c->launches));
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaFree(old));
    }
    return KQ_OK;
}

int stage_reserve(kq_ctx *c, Stage &s, size_t words, size_t dwords, size_t out_n) {
    if (words > s.words) {
        if (s.h_bases) { cudaFreeHost(s.h_bases); cudaFreeHost(s.h_inv); }
        s.h_bases = s.h_inv = nullptr; s.words = 0;
        CU(cudaMallocHost(&s.h_bases, words * 8)); CU(cudaMallocHost(&s.h_inv, words * 8));
        s.words = words;
    }
    if (dwords > s.dwords) {
        if (s.d_bases) { cudaFree(s.d_bases); cudaFree(s.d_inv); }
        s.d_bases = s.d_inv = nullptr; s.dwords = 0;
        CU(cudaMalloc(&s.d_bases, dwords * 8)); CU(cudaMalloc(&s.d_inv, dwords * 8));
        s.dwords = dwords;
    }
    if (out_n > s.out_n) {
        if (s.h_out) { cudaFreeHost(s.h_out); cudaFree(s.d_out); }
        s.h_out = s.d_out = nullptr; s.out_n = 0;
        CU(cudaMallocHost(&s.h_out, out_n * sizeof(Base))); CU(cudaMalloc(&s.d_out, out_n * sizeof(Base)));
        s.out_n = out_n;
    }
    return KQ_OK;
}

// wait until the GPU is done with staging slot s
int stage_acquire(kq_ctx *c, Stage &s) {
    if (s.busy) { CU(cudaEventSynchronize(s.done)); s.busy = false; }
    return KQ_OK;
}

cudaEvent_t get_event(kq_ctx *c) {
    if (!c->ev_pool.empty()) { cudaEvent_t e = c->ev_pool.back(); c->ev_pool.pop_back(); return e; }
    cudaEvent_t e = nullptr; cudaEventCreate(&e); return e;
}
// CUDA-event bracket on the launching stream; returns the index of the open range or -1
int time_begin(kq_ctx *c, int which) {
    if (!c->timing || which < 0) return -1;
    TimedRange r{get_event(c), get_event(c), which};
    cudaEventRecord(r.a, c->stream); c->ranges.push_back(r);
    return (int)c->ranges.size() - 1;
}
void time_end(kq_ctx *c, int idx) {
    if (idx >= 0) cudaEventRecord(c->ranges[idx].b, c->stream);
}
void time_collect(kq_ctx *c) {
    for (auto &r : c->ranges) {
        float ms = 0; cudaEventSynchronize(r.b); cudaEventElapsedTime(&ms, r.a, r.b);
        c->time_ms[r.which] += ms; ++c->time_n[r.which];
        c->ev_pool.push_back(r.a); c->ev_pool.push_back(r.b);
    }
    c->ranges.clear();
}

}  // namespace

// ---- count pipeline -------------------------------------------------------------------------------------

namespace {

int pipeline_scratch(kq_ctx *c, uint64_t tuples, bool level2 = false, uint64_t tuples2 = 0) {
    const uint64_t nbins = coarse_bins(view(c));
    if (c->hist_cap < nbins) {
        cudaFree(c->d_hist); cudaFree(c->d_cursor); cudaFree(c->d_cstart); cudaFree(c->d_cend); c->d_hist = c->d_cursor = c->d_cstart = c->d_cend = nullptr; c->hist_cap = 0;
        CU(cudaMalloc(&c->d_hist, nbins * 8)); CU(cudaMalloc(&c->d_cursor, nbins * 8)); CU(cudaMalloc(&c->d_cstart, nbins * 8)); CU(cudaMalloc(&c->d_cend, nbins * 8));
        c->hist_cap = nbins;
    }
    if (!c->d_ovf) CU(cudaMalloc(&c->d_ovf, OVF_CAP * 16));
    if (c->tup_cap < tuples) {
        cudaFree(c->d_tuples); c->d_tuples = nullptr; c->tup_cap = 0;
        CU(cudaMalloc(&c->d_tuples, tuples * 8 * tuple_words(c->k)));
        c->tup_cap = tuples;
    }
    if (!level2) return KQ_OK;
    if (!tuples2) tuples2 = tuples;
    if (c->tup2_cap < tuples2) {
        cudaFree(c->d_tuples2); c->d_tuples2 = nullptr; c->tup2_cap = 0;
        CU(cudaMalloc(&c->d_tuples2, tuples2 * 8 * tuple_words(c->k)));
        c->tup2_cap = tuples2;
    }
    if (c->fine_cap < c->nslices) {
        cudaFree(c->d_fhist); cudaFree(c->d_fcursor); cudaFree(c->d_fstart); c->d_fhist = c->d_fcursor = c->d_fstart = nullptr; c->fine_cap = 0;
        CU(cudaMalloc(&c->d_fhist, c->nslices * 8)); CU(cudaMalloc(&c->d_fcursor, c->nslices * 8)); CU(cudaMalloc(&c->d_fstart, c->nslices * 8));
        c->fine_cap = c->nslices;
    }
    const uint64_t tasks = std::max(tuples, tuples2) / tuning().part + std::max<uint64_t>(nbins, KQ_MAX_COARSE) + 1;
    if (c->task_cap < tasks) {
        cudaFree(c->d_tasks); c->d_tasks = nullptr; c->task_cap = 0;
        CU(cudaMalloc(&c->d_tasks, tasks * sizeof(FineTask)));
        c->task_cap = tasks;
    }
    if (!c->d_hot) CU(cudaMalloc(&c->d_hot, (1 + 3 * (size_t)KQ_HOT_CAP) * 8));
    if (!c->d_rstat) CU(cudaMalloc(&c->d_rstat, 3 * 8));
    if (c->ovf2_cap < OVF_CAP) {
        cudaFree(c->d_ovf2); c->d_ovf2 = nullptr; c->ovf2_cap = 0;
        CU(cudaMalloc(&c->d_ovf2, OVF_CAP * 16));
        c->ovf2_cap = OVF_CAP;
    }
    if (!c->d_ntasks) { CU(cudaMalloc(&c->d_ntasks, sizeof(unsigned))); CU(cudaMalloc(&c->d_task_first, (KQ_MAX_COARSE + 2) * sizeof(unsigned))); }
    const uint64_t h2 = c->task_cap * std::max<uint64_t>(1ull << c->cshift, c->route_ns_max);   // per-task histograms of the level-2 pass
    if (c->hist2_cap < h2) {
        cudaFree(c->d_hist2); c->d_hist2 = nullptr; c->hist2_cap = 0;
        CU(cudaMalloc(&c->d_hist2, h2 * sizeof(unsigned)));
        c->hist2_cap = h2;
    }
    return KQ_OK;
}

// After a synchronisation: put the tuples back whose slice was full (after doubling the table), and
// grow when the table got fuller than planned.
int settle(kq_ctx *c) {
    uint64_t occ[3];
    int rc = read_occ(c, occ); if (rc) return rc;
    rc = check_device_error(c); if (rc) return rc;
    c->occ_bound = occ[0];
    c->r_bound = occ[1];
    if ((double)occ[1] > 0.5 * (double)c->rcap) { rc = grow_side(c, 1, c->rcap * 4); if (rc) return rc; }
    if (!c->ovf_pending) return KQ_OK;
    c->ovf_pending = false;
    unsigned long long n = 0;
    CU(cudaMemcpyAsync(&n, c->d_ovf_n, 8, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (!n) return KQ_OK;
    // the shared-memory insert keeps the tuples of full slices in the (batch-sized) level-1 buffer, so a size
    // hint that was too small costs a grow and a second insert, never the batch; the other kernels have the
    // 4 Mi-tuple list
    uint64_t *list = c->ovf_list ? c->ovf_list : c->d_ovf;
    const uint64_t list_cap = c->ovf_list ? c->ovf_list_cap : OVF_CAP;
    if (n > list_cap)
        return fail(c, KQ_ENOMEM, "k-mer table too small for the batch (" + std::to_string(n) +
                                  " instances found their slice full): pass a larger expected_distinct to kq_create / kq_reserve");
    rc = grow_table(c, std::max<uint64_t>(2 * table_slots(c), (uint64_t)((double)(occ[0] + n) / TARGET_LOAD))); if (rc) return rc;
    CU(cudaMemsetAsync(c->d_ovf_n, 0, 8, c->stream));
    CU(cudaMemcpyAsync(c->d_nsrc, &n, 8, cudaMemcpyHostToDevice, c->stream));
    // second-chance failures go to the fixed list (a different buffer than the one being read)
    CU(launch_insert_tuples(c->stream, view(c), list, c->d_nsrc, n, c->d_ovf, OVF_CAP, c->d_ovf_n, c->d_ticket, &c->launches));
    CU(cudaMemcpyAsync(&n, c->d_ovf_n, 8, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (n) return fail(c, KQ_ENOMEM, "k-mer table full after growing");
    rc = read_occ(c, occ); if (rc) return rc;
    c->occ_bound = occ[0];
    return KQ_OK;
}

// a slice that receives more than eight times its share of a batch (and more than 4096 tuples) is a hot one
HotList hot_list(kq_ctx *c, uint64_t bound) {
    return HotList{c->d_hot, KQ_HOT_CAP, (unsigned)std::min<uint64_t>(1u << 30, std::max<uint64_t>(4096, 8 * (bound / c->nslices)))};
}

// one batch of the pipeline: every instance of src goes into the table
int run_batch(kq_ctx *c, Source src) {
    const uint64_t bound = src.tuples ? src.ntuples_max : src.npos;
    if (!bound) return KQ_OK;
    const TableView tv = view(c);
    const int tw = tuple_words(c->k);
    const unsigned ncoarse = coarse_bins(tv);
    const int cap = c->nslices > 1 ? scatter_pick_cap(ncoarse, tw) : 0;
    { int rc = pipeline_scratch(c, 0); if (rc) return rc; }
    c->ovf_list = nullptr; c->ovf_list_cap = 0;
    const int t0 = time_begin(c, 0);
    // big table, batch worth a sweep: one-pass scatter into per-slice regions, then the shared-memory insert
    if (tuning().pipeline == 4 && c->nslices > 1 && c->slice_bits <= 13 &&
        (double)bound * 100.0 >= (double)tuning().direct_pct * (double)table_slots(c)) {
        const int stride = tuning().cursor_stride;
        const uint64_t cap_t = ((bound / c->nslices) * 5 / 4 + 64 + 31) & ~31ull;      // tuples per slice region
        const uint64_t region = c->nslices * cap_t;
        if (c->tup2_cap < region) {
            cudaFree(c->d_tuples2); c->d_tuples2 = nullptr; c->tup2_cap = 0;
            CU(cudaMalloc(&c->d_tuples2, region * 8 * tw));
            c->tup2_cap = region;
        }
        if (c->scur_cap < c->nslices * (uint64_t)stride) {
            cudaFree(c->d_scur); c->d_scur = nullptr; c->scur_cap = 0;
            CU(cudaMalloc(&c->d_scur, c->nslices * (uint64_t)stride * 4));
            c->scur_cap = c->nslices * (uint64_t)stride;
        }
        const uint64_t want_ovf = std::max<uint64_t>(OVF_CAP, bound / 8);
        if (c->ovf2_cap < want_ovf) {
            cudaFree(c->d_ovf2); c->d_ovf2 = nullptr; c->ovf2_cap = 0;
            CU(cudaMalloc(&c->d_ovf2, want_ovf * 8 * tw));
            c->ovf2_cap = want_ovf;
        }
        CU(cudaMemsetAsync(c->d_scur, 0, c->nslices * (uint64_t)stride * 4, c->stream));
        const int t5 = time_begin(c, 5);
        CU(launch_scatter_fine(c->stream, tv, src, c->d_tuples2, (unsigned)cap_t, c->d_scur, stride, src.tuples ? nullptr : c->d_inst,
                               c->d_ovf2, c->ovf2_cap, c->d_ovf_n, &c->launches));
        time_end(c, t5);
        const int t2 = time_begin(c, 2);
        CU(launch_smem_insert(c->stream, tv, c->d_tuples2, nullptr, c->d_scur, cap_t, stride, c->d_ovf2, c->ovf2_cap, c->d_ovf_n, c->d_ticket, &c->launches));
        time_end(c, t2);
        c->ovf_list = c->d_ovf2; c->ovf_list_cap = c->ovf2_cap;
        c->ovf_pending = true;
        c->route_task_bound = 0;
        time_end(c, t0);
        return KQ_OK;
    }
    // routed pipeline: the slab is in bin order already, its tasks were planned by kq_route_scatter.  Knob route_smem,
    // off by default: the L2 CAS insert over the slab (below; a bin's hash range is an L2-sized window of the table)
    // takes 69 ms per 2.6 G tuples at any GPU count, level-2 scatter + shared-memory insert 70 / 82 / 116 ms at 2 / 4 / 8 GPUs
    // (the slab's bins get wider as the ranks share 2 048 bins)
    if (tuning().route_smem && tuning().pipeline == 3 && src.ordered && c->route_task_bound && bound == c->route_extent && c->nslices > 1 && tuning().smem_insert && c->slice_bits <= 13 &&
        fine_pick_cap(c->route_ns_max, tw) != 0) {
        int rc = pipeline_scratch(c, 0, true, bound); if (rc) return rc;
        const int t4 = time_begin(c, 4);
        // per-task histograms -> private cursors, as in the local pipeline (a slice may lie across two bins of the slab)
        CU(launch_fine_hist(c->stream, tv, src.tuples, c->d_tasks, c->d_ntasks, c->route_task_bound, c->route_ns_max, nullptr, c->d_hist2, &c->launches));
        CU(launch_route_fine_offsets(c->stream, tv, c->route.nb_bits, c->d_tasks, c->d_task_first, c->d_hist2, c->route_ns_max, c->d_fhist, &c->launches));
        CU(launch_slice_offsets(c->stream, c->nslices, c->d_fhist, c->d_fcursor, c->d_fstart, c->d_total, nullptr, &c->launches));
        const int t6 = time_begin(c, 6);
        CU(launch_fine_scatter(c->stream, tv, src.tuples, c->d_tasks, c->d_ntasks, c->route_task_bound, c->route_ns_max, c->d_tuples2, nullptr, c->d_hist2, c->d_fstart, &c->launches));
        time_end(c, t6);
        time_end(c, t4);
        const int t2 = time_begin(c, 2);
        // tuples of full slices go back into the slab (read completely by now, as large as the batch)
        CU(launch_smem_insert(c->stream, tv, c->d_tuples2, c->d_fstart, c->d_fhist, 0, 1, c->d_slab, c->slab_cap, c->d_ovf_n, c->d_ticket, &c->launches,
                              hot_list(c, bound)));
        time_end(c, t2);
        c->ovf_list = c->d_slab; c->ovf_list_cap = c->slab_cap;
        c->ovf_pending = true;
        c->route_task_bound = 0;
        time_end(c, t0);
        return KQ_OK;
    }
    // a sweep over the table pays off once the batch touches a good part of it
    if (src.ordered || c->nslices == 1 || cap == 0 || (double)bound * 100.0 < (double)tuning().direct_pct * (double)table_slots(c)) {
        if (src.tuples) {
            const int t2 = time_begin(c, src.ordered ? 2 : -1);
            CU(launch_insert_tuples(c->stream, tv, src.tuples, src.ntuples, bound, c->d_ovf, OVF_CAP, c->d_ovf_n, c->d_ticket, &c->launches));
            time_end(c, t2);
            c->ovf_pending = true;
        } else {
            CU(launch_count_direct(c->stream, tv, src, c->d_inst, c->d_ovf, OVF_CAP, c->d_ovf_n, &c->launches));
            c->ovf_pending = true;
        }
        time_end(c, t0);
        return KQ_OK;
    }
    // level 2 + shared-memory insert when a slice fits one CTA's shared memory and its bins the scatter's
    const unsigned ns_max = 1u << c->cshift;
    const bool level2 = tuning().pipeline == 3 && tuning().smem_insert && c->slice_bits <= 13 && fine_pick_cap(ns_max, tw) != 0;
    // one-pass form (knob, off: measured no faster -- the scatter kernel takes 51.6 instead of 41.8 ms on C2, as much as
    // the histogram pass it saves): the hash spreads the keys evenly, so every coarse bin gets a fixed region of its share
    // + 3 % and no histogram pass is needed; chunks that do not fit (runs of one key) go to a list and are counted with L2 CAS
    // (knob value 2: also for small batches and with almost no slack, so that tests reach the overflow paths)
    const bool one_pass = level2 && (tuning().one_pass == 2 || (tuning().one_pass == 1 && bound >= (1ull << 24)));
    // (an odd number of 64-byte units: region starts that are a large power of two apart would put the write streams of
    // all bins on the same memory channels at the same time)
    const uint64_t rcap = one_pass ? (((bound / ncoarse + bound / ncoarse / 32 + (tuning().one_pass == 2 ? 64 : 65536) + 15) & ~15ull) + 8) : 0;
    int rc = pipeline_scratch(c, one_pass ? rcap * ncoarse : bound, level2, bound + (one_pass ? (uint64_t)ncoarse * 64 : 0)); if (rc) return rc;
    const int t3 = time_begin(c, 3);
    if (one_pass) {
        CU(launch_region_init(c->stream, ncoarse, rcap, c->d_cursor, c->d_cstart, c->d_cend, c->d_rstat, &c->launches));
        const int t5 = time_begin(c, 5);
        // (overflow list: the level-2 output buffer, free until the level-2 scatter and as large as the batch)
        CU(launch_slice_scatter(c->stream, tv, src, c->d_tuples, c->d_cursor, cap, &c->launches, Regions{rcap, c->d_tuples2, c->tup2_cap, c->d_rstat}));
        time_end(c, t5);
        // its tuples are counted with L2 CAS right away; the ones whose slice is full wait in the small list (stat[2])
        CU(launch_insert_tuples(c->stream, tv, c->d_tuples2, c->d_rstat, std::max<uint64_t>(bound / 64, 4096), c->d_ovf2, c->ovf2_cap, c->d_rstat + 2,
                                c->d_ticket, &c->launches, c->tup2_cap));
        CU(launch_region_counts(c->stream, ncoarse, c->d_cursor, c->d_cstart, c->d_cend, c->d_hist, c->d_total, c->d_rstat,
                                src.tuples ? nullptr : c->d_inst, &c->launches));
    } else {
        CU(cudaMemsetAsync(c->d_hist, 0, (size_t)ncoarse * 8, c->stream));
        CU(launch_slice_hist(c->stream, tv, src, c->d_hist, &c->launches));
        CU(launch_slice_offsets(c->stream, ncoarse, c->d_hist, c->d_cursor, c->d_cstart, c->d_total,
                                src.tuples ? nullptr : c->d_inst, &c->launches));
        const int t5 = time_begin(c, 5);
        CU(launch_slice_scatter(c->stream, tv, src, c->d_tuples, c->d_cursor, cap, &c->launches));
        time_end(c, t5);
    }
    time_end(c, t3);
    if (!level2) {
        const int t2 = time_begin(c, 2);
        CU(launch_insert_tuples(c->stream, tv, c->d_tuples, c->d_total, bound, c->d_ovf, OVF_CAP, c->d_ovf_n, c->d_ticket, &c->launches));
        time_end(c, t2);
        c->ovf_pending = true;
        time_end(c, t0);
        return KQ_OK;
    }
    const int t4 = time_begin(c, 4);
    const unsigned task_bound = (unsigned)std::min<uint64_t>(c->task_cap, (bound + (uint64_t)ncoarse * 64) / tuning().part + ncoarse + 1);
    CU(launch_fine_plan(c->stream, tv, ncoarse, c->d_cstart, c->d_hist, tuning().part, c->d_tasks, task_bound, c->d_ntasks, c->d_task_first, &c->launches));
    CU(launch_fine_hist(c->stream, tv, c->d_tuples, c->d_tasks, c->d_ntasks, task_bound, ns_max, nullptr, c->d_hist2, &c->launches));
    // (one-pass form: the level-2 output is dense, bin b starts at the scan of the extents left in d_cursor)
    CU(launch_fine_offsets(c->stream, tv, ncoarse, one_pass ? c->d_cursor : c->d_cstart, c->d_task_first, c->d_hist2, ns_max, c->d_fstart, c->d_fhist, &c->launches));
    const int t6 = time_begin(c, 6);
    CU(launch_fine_scatter(c->stream, tv, c->d_tuples, c->d_tasks, c->d_ntasks, task_bound, ns_max, c->d_tuples2, nullptr, c->d_hist2, c->d_fstart, &c->launches));
    time_end(c, t6);
    time_end(c, t4);
    const int t2 = time_begin(c, 2);
    // tuples of full slices go back into the level-1 buffer (free by now, as large as the batch)
    CU(launch_smem_insert(c->stream, tv, c->d_tuples2, c->d_fstart, c->d_fhist, 0, 1, c->d_tuples, c->tup_cap, c->d_ovf_n, c->d_ticket, &c->launches,
                          hot_list(c, bound)));
    // tuples of the one-pass scatter's overflow list whose slice was full: same list, same counter as the insert above
    if (one_pass) CU(launch_insert_tuples(c->stream, tv, c->d_ovf2, c->d_rstat + 2, 4096, c->d_tuples, c->tup_cap, c->d_ovf_n, c->d_ticket, &c->launches, c->ovf2_cap));
    time_end(c, t2);
    c->ovf_list = c->d_tuples; c->ovf_list_cap = c->tup_cap;
    c->ovf_pending = true;
    time_end(c, t0);
    return KQ_OK;
}

// Count the k-mers of start positions [pos0, pos0+npos) of a packed device-resident sequence (or of
// ntuples tuples).  Batches are planned so that the table cannot overfill: without a size hint a
// batch brings at most as many instances as there is room below MAX_LOAD (every instance could be a
// new key); with a hint (kq_create / kq_reserve) four times that, backed by the overflow list.
int count_source(kq_ctx *c, Source src) {
    uint64_t left = src.tuples ? src.ntuples_max : src.npos;
    while (left) {
        if (c->ovf_pending || c->occ_bound != c->occ_known) { int rc = settle(c); if (rc) return rc; }
        int64_t room = (int64_t)(MAX_LOAD * (double)table_slots(c)) - (int64_t)c->occ_known;
        const int64_t want = (int64_t)std::min<uint64_t>(left, 1ull << 20);
        if (room < want || (double)c->occ_known > 0.6 * (double)table_slots(c)) {
            int rc = grow_table(c, std::max<uint64_t>(2 * table_slots(c), (uint64_t)((double)(c->occ_known + want) / TARGET_LOAD)));
            if (rc) return rc;
            continue;
        }
        uint64_t nb = c->optimistic ? (uint64_t)room * 4 : (uint64_t)room;
        nb = std::min(nb, left);
        nb = std::min<uint64_t>(nb, 1ull << 32);                       // 32-bit cursors and positions inside a batch
        Source b = src;
        if (!src.tuples) { b.pos0 = src.pos0 + (src.npos - left); b.npos = nb; }
        else {
            b.tuples = src.tuples + (src.ntuples_max - left) * tuple_words(c->k); b.ntuples_max = nb; b.ntuples = c->d_nsrc;
            const unsigned long long v = nb;
            CU(cudaMemcpyAsync(c->d_nsrc, &v, 8, cudaMemcpyHostToDevice, c->stream));
        }
        // R cannot overflow inside the batch: a key reaches it with its 128th instance at the earliest
        c->r_bound += nb / 128 + 1;
        if ((double)c->r_bound * 1.25 + 4096 > (double)c->rcap) {
            int rc = grow_side(c, 1, std::max<uint64_t>(2 * c->rcap, (uint64_t)((double)c->r_bound * 1.5) + 8192)); if (rc) return rc;
        }
        int rc = run_batch(c, b); if (rc) return rc;
        c->occ_bound = c->occ_known + nb;
        left -= nb;
    }
    return KQ_OK;
}

int pending_reserve(kq_ctx *c, uint64_t bases) {
    if (bases <= c->pend_cap) return KQ_OK;
    const uint64_t cap = (std::max<uint64_t>(std::max<uint64_t>(bases, 2 * c->pend_cap), 1ull << 26) + 63) & ~63ull;
    uint64_t *nb = nullptr, *ni = nullptr;
    const size_t bw = cap / 32 + 16, iw = cap / 64 + 16;
    CU(cudaMalloc(&nb, bw * 8)); CU(cudaMalloc(&ni, iw * 8));
    CU(cudaMemsetAsync(nb, 0, bw * 8, c->stream));
    CU(cudaMemsetAsync(ni, 0xFF, iw * 8, c->stream));                // everything not written is "not ACGT"
    CU(cudaStreamSynchronize(c->copy_stream));                       // copies into the old buffer have landed
    if (c->pend_len) {
        CU(cudaMemcpyAsync(nb, c->d_pb, (c->pend_len / 32) * 8, cudaMemcpyDeviceToDevice, c->stream));
        CU(cudaMemcpyAsync(ni, c->d_pi, (c->pend_len / 64) * 8, cudaMemcpyDeviceToDevice, c->stream));
    }
    CU(cudaStreamSynchronize(c->stream));
    cudaFree(c->d_pb); cudaFree(c->d_pi);
    c->d_pb = nb; c->d_pi = ni; c->pend_cap = cap;
    return KQ_OK;
}

// run the pipeline over everything pending
int pending_flush(kq_ctx *c) {
    if (!c->pend_len || c->hold_pending) return KQ_OK;
    const uint64_t len = c->pend_len;
    CU(cudaStreamWaitEvent(c->stream, c->ev_copied, 0));               // every batch of this buffer has arrived
    if (len >= (uint64_t)c->k) {
        Source src{c->d_pb, c->d_pi, 0, len - c->k + 1, nullptr, nullptr, 0};
        int rc = count_source(c, src); if (rc) return rc;
    }
    CU(cudaMemsetAsync(c->d_pi, 0xFF, (len / 64 + 1) * 8, c->stream));
    CU(cudaEventRecord(c->ev_read, c->stream));                        // the buffer may be refilled after this point
    c->read_armed = true;
    c->pend_len = 0;
    // the next batches go to the other buffer, so their copies overlap the kernels launched above
    std::swap(c->d_pb, c->d_pb_alt); std::swap(c->d_pi, c->d_pi_alt); std::swap(c->pend_cap, c->pend_cap_alt);
    std::swap(c->ev_read, c->ev_read_alt); std::swap(c->read_armed, c->read_armed_alt);
    return KQ_OK;
}

}  // namespace

// ---- lifetime -----------------------------------------------------------------------------------------

extern "C" int kq_create(kq_ctx **out, int k, int device, uint64_t expected_distinct) {
    kq_ctx *c = nullptr;
    if (!out) return fail(c, KQ_EINVAL, "kq_create: out is NULL");
    *out = nullptr;
    if (k < 1 || k > 32) return fail(c, KQ_EINVAL, "Invalid kmer length.");            // src/input.cpp:142
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(c, KQ_ENODEV, std::string("no CUDA device: ") + (e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e)));
    if (device < 0 || device >= ndev) return fail(c, KQ_ENODEV, "CUDA device ordinal out of range");
    if (cudaSetDevice(device) != cudaSuccess) return fail(c, KQ_ENODEV, "cudaSetDevice failed");
    kq_ctx *ctx = new kq_ctx;
    ctx->k = k; ctx->device = device;
    c = ctx;
    auto bail = [&](int rc) { g_create_error = ctx->err; kq_destroy(ctx); return rc; };
#define CC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_); \
                      return bail(e_ == cudaErrorMemoryAllocation ? KQ_ENOMEM : KQ_ENODEV); } } while (0)
    CC(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
    ctx->stream = ctx->own_stream;
    CC(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    CC(cudaEventCreateWithFlags(&ctx->ev_copied, cudaEventDisableTiming));
    CC(cudaEventCreateWithFlags(&ctx->ev_read, cudaEventDisableTiming));
    CC(cudaEventCreateWithFlags(&ctx->ev_read_alt, cudaEventDisableTiming));
    CC(cudaMalloc(&ctx->d_occ, 3 * sizeof(unsigned long long)));
    CC(cudaMalloc(&ctx->d_inst, sizeof(unsigned long long)));
    CC(cudaMalloc(&ctx->d_acc, 3 * sizeof(unsigned long long)));
    CC(cudaMalloc(&ctx->d_total, sizeof(unsigned long long)));
    CC(cudaMalloc(&ctx->d_nsrc, sizeof(unsigned long long)));
    CC(cudaMalloc(&ctx->d_ticket, sizeof(unsigned long long)));
    CC(cudaMalloc(&ctx->d_ovf_n, sizeof(unsigned long long)));
    CC(cudaMalloc(&ctx->d_err, sizeof(int)));
    CC(cudaMemsetAsync(ctx->d_occ, 0, 3 * sizeof(unsigned long long), ctx->stream));
    CC(cudaMemsetAsync(ctx->d_inst, 0, sizeof(unsigned long long), ctx->stream));
    CC(cudaMemsetAsync(ctx->d_acc, 0, 3 * sizeof(unsigned long long), ctx->stream));
    CC(cudaMemsetAsync(ctx->d_ovf_n, 0, sizeof(unsigned long long), ctx->stream));
    CC(cudaMemsetAsync(ctx->d_err, 0, sizeof(int), ctx->stream));
    for (auto &s : ctx->st) { CC(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming)); CC(cudaEventCreateWithFlags(&s.kernel_done, cudaEventDisableTiming)); }
#undef CC
    ctx->optimistic = expected_distinct != 0;
    const uint64_t want = expected_distinct ? (uint64_t)((double)expected_distinct / TARGET_LOAD) : (1ull << 20);
    int rc = grow_table(ctx, want); if (rc) return bail(rc);
    rc = grow_side(ctx, 1, MIN_SCAP); if (rc) return bail(rc);
    rc = grow_side(ctx, 2, MIN_SCAP); if (rc) return bail(rc);
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) { ctx->err = "stream sync failed"; return bail(KQ_ECUDA); }
    *out = ctx;
    return KQ_OK;
}

extern "C" void kq_destroy(kq_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (auto &s : c->st) {
        if (s.h_bases) { cudaFreeHost(s.h_bases); cudaFreeHost(s.h_inv); }
        if (s.d_bases) { cudaFree(s.d_bases); cudaFree(s.d_inv); }
        if (s.h_out) { cudaFreeHost(s.h_out); cudaFree(s.d_out); }
        if (s.done) cudaEventDestroy(s.done);
        if (s.kernel_done) cudaEventDestroy(s.kernel_done);
    }
    for (auto &r : c->ranges) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    for (auto e : c->ev_pool) cudaEventDestroy(e);
    cudaFree(c->T); cudaFree(c->R); cudaFree(c->M); cudaFree(c->d_occ); cudaFree(c->d_inst); cudaFree(c->d_acc); cudaFree(c->d_err);
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    if (c->ev_copied) cudaEventDestroy(c->ev_copied);
    if (c->ev_read) cudaEventDestroy(c->ev_read);
    if (c->ev_read_alt) cudaEventDestroy(c->ev_read_alt);
    cudaFree(c->d_pb_alt); cudaFree(c->d_pi_alt);
    cudaFree(c->d_pb); cudaFree(c->d_pi); cudaFree(c->d_tuples); cudaFree(c->d_hist); cudaFree(c->d_cursor);
    cudaFree(c->d_tuples2); cudaFree(c->d_cstart); cudaFree(c->d_fhist); cudaFree(c->d_fcursor); cudaFree(c->d_fstart);
    cudaFree(c->d_tasks); cudaFree(c->d_ntasks); cudaFree(c->d_task_first); cudaFree(c->d_hist2); cudaFree(c->d_scur); cudaFree(c->d_ovf2);
    cudaFree(c->d_total); cudaFree(c->d_nsrc); cudaFree(c->d_ticket); cudaFree(c->d_hot); cudaFree(c->d_cend); cudaFree(c->d_rstat); cudaFree(c->d_ovf); cudaFree(c->d_ovf_n);
    for (void *p : c->ipc_opened) cudaIpcCloseMemHandle(p);
    for (void *p : c->peer_opened) cudaIpcCloseMemHandle(p);
    cudaFree(c->d_peers);
    cudaFree(c->d_slab); cudaFree(c->d_rcursor); cudaFree(c->d_rback); cudaFree(c->d_rtotals);
    cudaFree(c->d_vinbox); cudaFree(c->d_vrows); cudaFree(c->d_vcursor); cudaFree(c->d_vacc); cudaFree(c->d_vpos); for (auto *p : c->d_out4) cudaFree(p); cudaFree(c->d_wide); cudaFree(c->d_wide_n);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

extern "C" const char *kq_last_error(kq_ctx *c) { return c ? c->err.c_str() : g_create_error.c_str(); }
extern "C" int kq_k(kq_ctx *c) { return c ? c->k : KQ_EINVAL; }

extern "C" int kq_set_stream(kq_ctx *c, void *stream) {
    if (!c) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    for (auto &s : c->st) s.busy = false;
    c->stream = stream ? (cudaStream_t)stream : c->own_stream;
    return KQ_OK;
}

// flush pending reads, wait, settle the table
extern "C" int kq_sync(kq_ctx *c) {
    if (!c) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    int rc = pending_flush(c); if (rc) return rc;
    CU(cudaStreamSynchronize(c->stream));
    for (auto &s : c->st) s.busy = false;
    return settle(c);
}

extern "C" int kq_owner_of(uint64_t key, int nranks) {
    if (nranks < 1 || nranks > 128 || (nranks & (nranks - 1))) return KQ_EINVAL;
    return (int)((key % 128) / (128 / (unsigned)nranks));
}

extern "C" int kq_set_shard(kq_ctx *c, int rank, int nranks) {
    if (!c) return KQ_EINVAL;
    if (nranks < 1 || nranks > 128 || (nranks & (nranks - 1)) || rank < 0 || rank >= nranks)
        return fail(c, KQ_EINVAL, "kq_set_shard: nranks must be a power of two <= 128 and 0 <= rank < nranks");
    c->rank = rank; c->nranks = nranks;
    return KQ_OK;
}

extern "C" int kq_set_timing(kq_ctx *c, int on) { if (!c) return KQ_EINVAL; c->timing = on != 0; return KQ_OK; }

extern "C" int kq_kernel_time_ms(kq_ctx *c, int which, double *ms, uint64_t *launches, int reset) {
    if (!c || which < 0 || which >= NTIMERS) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    time_collect(c);
    if (ms) *ms = c->time_ms[which];
    if (launches) *launches = c->time_n[which];
    if (reset) { c->time_ms[which] = 0; c->time_n[which] = 0; }
    return KQ_OK;
}

extern "C" int kq_counters(kq_ctx *c, uint64_t *kernel_launches, uint64_t *h2d_bytes, uint64_t *d2h_bytes) {
    if (!c) return KQ_EINVAL;
    if (kernel_launches) *kernel_launches = c->launches;
    if (h2d_bytes) *h2d_bytes = c->h2d;
    if (d2h_bytes) *d2h_bytes = c->d2h;
    return KQ_OK;
}

extern "C" int kq_reserve(kq_ctx *c, uint64_t extra) {
    if (!c) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    int rc = kq_sync(c); if (rc) return rc;
    c->optimistic = true;
    if ((double)(c->occ_known + extra) > TARGET_LOAD * (double)table_slots(c)) {
        rc = grow_table(c, (uint64_t)((double)(c->occ_known + extra) / TARGET_LOAD)); if (rc) return rc;
    }
    return KQ_OK;
}

// process-wide tuning knobs (kq_kernels.h Tuning); geometry knobs apply to tables created or grown afterwards
extern "C" int kq_set_tuning(const char *name, long long value) {
    if (!name) return KQ_EINVAL;
    Tuning &t = tuning();
    const std::string n(name);
    if (n == "single_slice_bits" && value >= 8 && value <= 26) t.single_slice_bits = (int)value;
    else if (n == "piece_bits" && value >= 6 && value <= 13) t.piece_bits = (int)value;
    else if (n == "fine_part" && value >= 64) t.part = (unsigned long long)value;
    else if (n == "smem_insert" && (value == 0 || value == 1)) t.smem_insert = (int)value;
    else if (n == "direct_pct" && value >= 0 && value <= 100) t.direct_pct = (int)value;
    else if (n == "scatter_cap" && value >= 0 && value <= 512) t.scatter_cap = (int)value;
    else if (n == "insert_ctas" && value >= 1 && value <= 8) t.insert_ctas = (int)value;
    else if (n == "validate_ctas" && value >= 1 && value <= 8) t.validate_ctas = (int)value;
    else if (n == "db_host" && (value == 0 || value == 1)) t.db_host = (int)value;
    else if (n == "coarse_bias" && value >= 0 && value <= 4) t.coarse_bias = (int)value;
    else if (n == "db_ordered" && (value == 0 || value == 1)) t.db_ordered = (int)value;
    else if (n == "one_pass" && value >= 0 && value <= 2) t.one_pass = (int)value;
    else if (n == "cta_lock" && (value == 0 || value == 1)) t.cta_lock = (int)value;
    else if (n == "route_smem" && (value == 0 || value == 1)) t.route_smem = (int)value;
    else if (n == "warp_insert" && (value == 0 || value == 1)) t.warp_insert = (int)value;
    else if (n == "s3_threads" && (value == 256 || value == 384 || value == 512)) t.s3_threads = (int)value;
    else if (n == "s3_ctas" && value >= 1 && value <= 8) t.s3_ctas = (int)value;
    else if (n == "pipeline" && value >= 2 && value <= 4) t.pipeline = (int)value;
    else if (n == "cursor_stride" && value >= 1 && value <= 32) t.cursor_stride = (int)value;
    else return KQ_EINVAL;
    return KQ_OK;
}

extern "C" int kq_set_flush_positions(kq_ctx *c, uint64_t positions) {
    if (!c) return KQ_EINVAL;
    c->flush_positions = positions ? positions : DEFAULT_FLUSH;
    return KQ_OK;
}

extern "C" int kq_clear(kq_ctx *c) {
    if (!c) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    if (c->pend_len) {
        CU(cudaStreamSynchronize(c->copy_stream));                     // batches still on their way into the buffer
        CU(cudaMemsetAsync(c->d_pi, 0xFF, (c->pend_len / 64 + 1) * 8, c->stream)); c->pend_len = 0;
        CU(cudaEventRecord(c->ev_read, c->stream)); c->read_armed = true;
    }
    CU(cudaMemsetAsync(c->T, 0, table_slots(c) * sizeof(Slot), c->stream));
    CU(cudaMemsetAsync(c->R, 0, c->rcap * sizeof(SideSlot), c->stream));
    CU(cudaMemsetAsync(c->M, 0, c->mcap * sizeof(SideSlot), c->stream));
    CU(cudaMemsetAsync(c->d_occ, 0, 3 * sizeof(unsigned long long), c->stream));
    CU(cudaMemsetAsync(c->d_inst, 0, sizeof(unsigned long long), c->stream));
    CU(cudaMemsetAsync(c->d_ovf_n, 0, sizeof(unsigned long long), c->stream));
    CU(cudaMemsetAsync(c->d_err, 0, sizeof(int), c->stream));
    c->occ_known = c->occ_bound = 0; c->r_bound = 0; c->big_seen = 0; c->inst_tuples = 0; c->ovf_pending = false; c->loaded_db = false;
    return KQ_OK;
}

// ---- count ------------------------------------------------------------------------------------------------

extern "C" size_t kq_packed_words(size_t len) { return packed_words(len); }

extern "C" int kq_pack_host(const char *ascii, size_t len, uint64_t *bases, uint64_t *invalid, int nthreads) {
    if ((!ascii && len) || !bases || !invalid) return KQ_EINVAL;
    pack_sequence(ascii, len, bases, invalid, nthreads == 1 ? nullptr : &global_pool());
    return KQ_OK;
}

extern "C" int kq_count_push(kq_ctx *c, const char *ascii, size_t len) {
    if (!c || (!ascii && len)) return KQ_EINVAL;
    if (c->loaded_db) return fail(c, KQ_ESTATE, "kq_count_push: the table holds a loaded/merged database");
    if (len < (size_t)c->k) return KQ_OK;                              // src/graph-builder.cpp:60-64
    CU(cudaSetDevice(c->device));
    if (c->pend_len && c->pend_len + len >= c->flush_positions) { int rc = pending_flush(c); if (rc) return rc; }
    int rc = pending_reserve(c, c->pend_len + len + 128); if (rc) return rc;
    // the batch goes to the end of the pending stream, 64-base aligned; at least one invalid position
    // separates it from the next batch (== the '\0' / no-previous-base cases of graph-builder.cpp:98-112)
    const uint64_t at = c->pend_len;
    if (at == 0 && c->read_armed) {                                    // first batch into a buffer the pipeline has read before
        CU(cudaStreamWaitEvent(c->copy_stream, c->ev_read, 0));
        c->read_armed = false;
    }
    for (uint64_t c0 = 0; c0 < len; c0 += CHUNK) {
        const uint64_t n = std::min<uint64_t>(CHUNK, len - c0);
        Stage &s = c->st[c->next_stage]; c->next_stage = (c->next_stage + 1) % NBUF;
        rc = stage_acquire(c, s); if (rc) return rc;
        rc = stage_reserve(c, s, packed_words(CHUNK), 0, 0); if (rc) return rc;
        pack_sequence(ascii + c0, n, s.h_bases, s.h_inv, &global_pool());
        const size_t bw = (n + 31) / 32, iw = (n + 63) / 64;
        CU(cudaMemcpyAsync(c->d_pb + (at + c0) / 32, s.h_bases, bw * 8, cudaMemcpyHostToDevice, c->copy_stream));
        CU(cudaMemcpyAsync(c->d_pi + (at + c0) / 64, s.h_inv, iw * 8, cudaMemcpyHostToDevice, c->copy_stream));
        c->h2d += (bw + iw) * 8;
        CU(cudaEventRecord(s.done, c->copy_stream)); s.busy = true;
    }
    CU(cudaEventRecord(c->ev_copied, c->copy_stream));
    c->pend_len = (at + len + 64) & ~63ull;                            // next 64-aligned position strictly after the end
    return KQ_OK;
}

extern "C" int kq_count_finish(kq_ctx *c) { return kq_sync(c); }

extern "C" int kq_count_instances(kq_ctx *c, uint64_t *n) {
    if (!c || !n) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    int rc = pending_flush(c); if (rc) return rc;
    unsigned long long v = 0;
    CU(cudaMemcpyAsync(&v, c->d_inst, sizeof v, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *n = v + c->inst_tuples;
    return KQ_OK;
}

extern "C" int kq_distinct(kq_ctx *c, uint64_t *n) {
    if (!c || !n) return KQ_EINVAL;
    int rc = kq_sync(c); if (rc) return rc;
    *n = c->occ_known;
    return KQ_OK;
}

extern "C" int kq_dev_count_packed(kq_ctx *c, const uint64_t *d_bases, const uint64_t *d_invalid,
                                   uint64_t len, uint64_t pos0, uint64_t npos) {
    if (!c || !d_bases || !d_invalid) return KQ_EINVAL;
    if (c->loaded_db) return fail(c, KQ_ESTATE, "kq_dev_count_packed: the table holds a loaded/merged database");
    if (len < (uint64_t)c->k) return KQ_OK;
    if (pos0 + npos > len - c->k + 1) return fail(c, KQ_EINVAL, "kq_dev_count_packed: positions beyond len-k+1");
    CU(cudaSetDevice(c->device));
    int rc = pending_flush(c); if (rc) return rc;
    return count_source(c, Source{d_bases, d_invalid, pos0, npos, nullptr, nullptr, 0});
}

extern "C" int kq_tuple_bytes(kq_ctx *c) { return c ? 8 * tuple_words(c->k) : KQ_EINVAL; }

extern "C" int kq_dev_extract_route(kq_ctx *c, const uint64_t *d_bases, const uint64_t *d_invalid,
                                    uint64_t len, uint64_t pos0, uint64_t npos,
                                    void *d_tuples, uint64_t cap_per_rank, unsigned long long *d_counts) {
    if (!c || !d_bases || !d_invalid || !d_tuples || !d_counts) return KQ_EINVAL;
    if (len < (uint64_t)c->k) return KQ_OK;
    if (pos0 + npos > len - c->k + 1) return fail(c, KQ_EINVAL, "kq_dev_extract_route: positions beyond len-k+1");
    CU(cudaSetDevice(c->device));
    const TableView tv = view(c);
    RouteArgs ra{static_cast<uint64_t *>(d_tuples), cap_per_rank, d_counts, c->nranks, tv.shard_shift};
    const int t = time_begin(c, 3);
    CU(launch_extract_route(c->stream, tv, Source{d_bases, d_invalid, pos0, npos, nullptr, nullptr, 0}, ra, &c->launches));
    time_end(c, t);
    return KQ_OK;
}

extern "C" int kq_dev_insert_tuples(kq_ctx *c, const void *d_tuples, uint64_t n) {
    if (!c || (n && !d_tuples)) return KQ_EINVAL;
    if (c->loaded_db) return fail(c, KQ_ESTATE, "kq_dev_insert_tuples: the table holds a loaded/merged database");
    if (!n) return KQ_OK;
    CU(cudaSetDevice(c->device));
    int rc = pending_flush(c); if (rc) return rc;
    rc = count_source(c, Source{nullptr, nullptr, 0, 0, static_cast<const uint64_t *>(d_tuples), c->d_nsrc, n}); if (rc) return rc;
    c->inst_tuples += n;
    return KQ_OK;
}

// ---- routed count over peer memory (row E) ---------------------------------------------------------------
// One process per GPU.  Every rank owns a receive slab; the scatter kernel of every rank writes the tuples
// a slab's owner must count straight into it (peer-mapped pointers, NVLink), already in slice order, so
// the all-to-all is the scatter kernel's own stores and the receiver only runs the insert kernel.

extern "C" int kq_route_setup(kq_ctx *c, uint64_t slab_tuples, int nb_bits, void **d_slab, void *ipc_handle) {
    if (!c || !slab_tuples || nb_bits < 0 || nb_bits > 12) return KQ_EINVAL;
    if (c->nranks > KQ_MAX_RANKS) return fail(c, KQ_EINVAL, "kq_route_setup: too many ranks");
    CU(cudaSetDevice(c->device));
    const int tw = tuple_words(c->k);
    if (scatter_pick_cap((unsigned long long)c->nranks << nb_bits, tw, KQ_ROUTE_MAX_CAP) == 0)
        return fail(c, KQ_EINVAL, "kq_route_setup: nranks << nb_bits bins do not fit shared memory");
    CU(cudaStreamSynchronize(c->stream));
    for (void *p : c->ipc_opened) cudaIpcCloseMemHandle(p);
    c->ipc_opened.clear();
    cudaFree(c->d_slab); c->d_slab = nullptr; c->slab_cap = 0; c->route_ready = false;
    CU(cudaMalloc(&c->d_slab, slab_tuples * 8 * tw));
    c->slab_cap = slab_tuples;
    if (!c->d_rcursor) {
        CU(cudaMalloc(&c->d_rcursor, (size_t)KQ_MAX_RANKS * 4096 * 8));
        CU(cudaMalloc(&c->d_rback, (size_t)KQ_MAX_RANKS * 4096 * 8));
        CU(cudaMalloc(&c->d_rtotals, 2 * KQ_MAX_RANKS * 8));
    }
    c->route = RouteMap{};
    c->route.nranks = c->nranks; c->route.nb_bits = nb_bits;
    c->route.shift = view(c).shard_shift;
    if (d_slab) *d_slab = c->d_slab;
    if (ipc_handle) {
        cudaIpcMemHandle_t h;
        CU(cudaIpcGetMemHandle(&h, c->d_slab));
        static_assert(sizeof h == 64, "CUDA IPC handles are 64 bytes");
        std::memcpy(ipc_handle, &h, sizeof h);
    }
    return KQ_OK;
}

extern "C" int kq_route_open(kq_ctx *c, const void *ipc_handle, void **d_ptr) {
    if (!c || !ipc_handle || !d_ptr) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, ipc_handle, sizeof h);
    void *p = nullptr;
    CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->ipc_opened.push_back(p);
    *d_ptr = p;
    return KQ_OK;
}

extern "C" int kq_route_connect(kq_ctx *c, void *const *slabs) {
    if (!c || !slabs || !c->d_slab) return KQ_EINVAL;
    for (int r = 0; r < c->nranks; ++r) {
        if (!slabs[r]) return fail(c, KQ_EINVAL, "kq_route_connect: NULL slab pointer");
        c->route.out[r] = static_cast<uint64_t *>(slabs[r]);
    }
    if (c->route.out[c->rank] != c->d_slab) return fail(c, KQ_EINVAL, "kq_route_connect: slabs[rank] is not this handle's slab");
    c->route_ready = true;
    return KQ_OK;
}

extern "C" int kq_route_bins(kq_ctx *c) { return c && c->d_slab ? c->nranks << c->route.nb_bits : KQ_EINVAL; }

extern "C" int kq_route_hist(kq_ctx *c, const uint64_t *d_bases, const uint64_t *d_invalid, uint64_t len,
                             uint64_t pos0, uint64_t npos, unsigned long long *d_hist) {
    if (!c || !d_bases || !d_invalid || !d_hist) return KQ_EINVAL;
    if (!c->d_slab) return fail(c, KQ_ESTATE, "kq_route_hist: call kq_route_setup first");
    if (len < (uint64_t)c->k) npos = 0;
    else if (pos0 + npos > len - c->k + 1) return fail(c, KQ_EINVAL, "kq_route_hist: positions beyond len-k+1");
    CU(cudaSetDevice(c->device));
    const int t = time_begin(c, 3);
    CU(launch_route_hist(c->stream, view(c), Source{d_bases, d_invalid, pos0, npos, nullptr, nullptr, 0, 0}, c->route, d_hist, &c->launches));
    time_end(c, t);
    return KQ_OK;
}

extern "C" int kq_route_scatter(kq_ctx *c, const uint64_t *d_bases, const uint64_t *d_invalid, uint64_t len,
                                uint64_t pos0, uint64_t npos, const unsigned long long *d_all_hist, uint64_t *nrecv) {
    if (!c || !d_bases || !d_invalid || !d_all_hist || !nrecv) return KQ_EINVAL;
    if (!c->route_ready) return fail(c, KQ_ESTATE, "kq_route_scatter: call kq_route_connect first");
    if (len < (uint64_t)c->k) npos = 0;
    else if (pos0 + npos > len - c->k + 1) return fail(c, KQ_EINVAL, "kq_route_scatter: positions beyond len-k+1");
    CU(cudaSetDevice(c->device));
    const int t = time_begin(c, 3);
    const int tw = tuple_words(c->k);
    const int cap = scatter_pick_cap((unsigned long long)c->nranks << c->route.nb_bits, tw, KQ_ROUTE_MAX_CAP);
    unsigned long long totals[2 * KQ_MAX_RANKS];
    // the padding sentinels go to the peers' slabs from this kernel: like the scatter, it runs after the
    // all-gather that ordered this round behind every rank's previous insert
    CU(launch_route_offsets(c->stream, c->route, c->rank, cap, tw, c->slab_cap, d_all_hist, c->d_rcursor, c->d_rback, c->d_rtotals, &c->launches));
    CU(cudaMemcpyAsync(totals, c->d_rtotals, 2 * c->nranks * 8, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    for (int r = 0; r < c->nranks; ++r)                  // every rank sees the same totals and takes the same decision
        if (totals[r] > c->slab_cap) {
            time_end(c, t);
            return fail(c, KQ_ENOMEM, "kq_route_scatter: rank " + std::to_string(r) + " would receive " + std::to_string(totals[r]) +
                                      " tuples, its slab holds " + std::to_string(c->slab_cap) + " (smaller rounds or a larger slab)");
        }
    CU(launch_route_scatter(c->stream, view(c), Source{d_bases, d_invalid, pos0, npos, nullptr, nullptr, 0, 0}, c->route, c->d_rcursor, c->d_rback, cap, &c->launches));
    time_end(c, t);
    // what this rank receives goes through the level-2 scatter and the shared-memory insert: plan its tasks from the
    // same histograms (bin extents inside the slab, slices a bin's hash range maps to on the local table)
    c->route_task_bound = 0;
    if (tuning().route_smem && tuning().pipeline == 3 && c->nslices > 1 && totals[c->rank] && (1 << c->route.nb_bits) <= KQ_ROUTE_PLAN_BINS) {
        const unsigned nb = 1u << c->route.nb_bits;
        c->route_ns_max = (unsigned)(c->nslices / nb + 2);
        int rc = pipeline_scratch(c, 0, true, totals[c->rank]); if (rc) return rc;
        const unsigned bound = (unsigned)std::min<uint64_t>(c->task_cap, totals[c->rank] / tuning().part + nb + 1);
        CU(launch_route_plan(c->stream, view(c), c->route, c->rank, cap, d_all_hist, tuning().part, c->d_tasks, bound, c->d_ntasks, c->d_task_first, &c->launches));
        c->route_task_bound = bound; c->route_extent = totals[c->rank];
    }
    c->route_real = totals[c->nranks + c->rank];
    *nrecv = totals[c->rank];
    return KQ_OK;
}

extern "C" int kq_route_insert(kq_ctx *c, uint64_t n) {
    if (!c || !c->d_slab || n > c->slab_cap) return KQ_EINVAL;
    if (c->loaded_db) return fail(c, KQ_ESTATE, "kq_route_insert: the table holds a loaded/merged database");
    if (!n) return KQ_OK;
    CU(cudaSetDevice(c->device));
    int rc = pending_flush(c); if (rc) return rc;
    rc = count_source(c, Source{nullptr, nullptr, 0, 0, c->d_slab, c->d_nsrc, n, 1}); if (rc) return rc;
    c->inst_tuples += c->route_real;                     // n is the slab extent including the alignment padding
    return KQ_OK;
}

// ---- peer tables: validate against every rank's shard over NVLink (row E) ----------------------------------

namespace {
struct PeerBlob {                     // what a rank publishes about its shard (kq_peer_export)
    cudaIpcMemHandle_t T, R, M;
    unsigned long long nslices, rcap, mcap; int slice_bits, k, rank, nranks;
};
static_assert(sizeof(PeerBlob) <= 256, "PeerBlob fits KQ_PEER_BLOB bytes");

int peers_install(kq_ctx *c, const std::vector<TableView> &views) {
    if (!c->d_peers) CU(cudaMalloc(&c->d_peers, KQ_MAX_RANKS * sizeof(TableView)));
    CU(cudaMemcpyAsync(c->d_peers, views.data(), views.size() * sizeof(TableView), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return KQ_OK;
}
}  // namespace

extern "C" int kq_peer_export(kq_ctx *c, void *blob) {
    if (!c || !blob) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    int rc = kq_sync(c); if (rc) return rc;
    PeerBlob b{};
    CU(cudaIpcGetMemHandle(&b.T, c->T)); CU(cudaIpcGetMemHandle(&b.R, c->R)); CU(cudaIpcGetMemHandle(&b.M, c->M));
    b.nslices = c->nslices; b.rcap = c->rcap; b.mcap = c->mcap; b.slice_bits = c->slice_bits;
    b.k = c->k; b.rank = c->rank; b.nranks = c->nranks;
    std::memset(blob, 0, 256); std::memcpy(blob, &b, sizeof b);
    c->exported = true;               // the tables must stay where they are until kq_peer_disconnect
    return KQ_OK;
}

extern "C" int kq_peer_connect(kq_ctx *c, const void *blobs) {
    if (!c || !blobs) return KQ_EINVAL;
    if (c->nranks > KQ_MAX_RANKS) return fail(c, KQ_EINVAL, "kq_peer_connect: too many ranks");
    CU(cudaSetDevice(c->device));
    std::vector<TableView> views(c->nranks, view(c));
    for (int r = 0; r < c->nranks; ++r) {
        if (r == c->rank) continue;
        PeerBlob b; std::memcpy(&b, static_cast<const char *>(blobs) + (size_t)r * 256, sizeof b);
        if (b.k != c->k || b.rank != r || b.nranks != c->nranks) return fail(c, KQ_EINVAL, "kq_peer_connect: blob " + std::to_string(r) + " does not describe rank " + std::to_string(r));
        void *T = nullptr, *R = nullptr, *M = nullptr;
        CU(cudaIpcOpenMemHandle(&T, b.T, cudaIpcMemLazyEnablePeerAccess)); c->peer_opened.push_back(T);
        CU(cudaIpcOpenMemHandle(&R, b.R, cudaIpcMemLazyEnablePeerAccess)); c->peer_opened.push_back(R);
        CU(cudaIpcOpenMemHandle(&M, b.M, cudaIpcMemLazyEnablePeerAccess)); c->peer_opened.push_back(M);
        TableView &v = views[r];
        v.T = static_cast<Slot *>(T); v.R = static_cast<SideSlot *>(R); v.M = static_cast<SideSlot *>(M);
        v.nslices = b.nslices; v.rcap = b.rcap; v.mcap = b.mcap; v.slice_bits = b.slice_bits;
    }
    return peers_install(c, views);
}

// same-process form: the other shards are handles of this process (tests; several GPUs driven by one process)
extern "C" int kq_peer_connect_local(kq_ctx *c, kq_ctx *const *handles) {
    if (!c || !handles) return KQ_EINVAL;
    if (c->nranks > KQ_MAX_RANKS) return fail(c, KQ_EINVAL, "kq_peer_connect_local: too many ranks");
    CU(cudaSetDevice(c->device));
    std::vector<TableView> views(c->nranks, view(c));
    for (int r = 0; r < c->nranks; ++r) {
        kq_ctx *o = handles[r];
        if (!o || o->k != c->k || o->rank != r || o->nranks != c->nranks) return fail(c, KQ_EINVAL, "kq_peer_connect_local: handle " + std::to_string(r) + " is not rank " + std::to_string(r));
        if (o == c) continue;
        int rc = kq_sync(o); if (rc) return fail(c, rc, kq_last_error(o));
        CU(cudaSetDevice(c->device));
        if (o->device != c->device) { cudaError_t e = cudaDeviceEnablePeerAccess(o->device, 0); if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CU(e); cudaGetLastError(); }
        TableView &v = views[r];
        v.T = o->T; v.R = o->R; v.M = o->M; v.nslices = o->nslices; v.rcap = o->rcap; v.mcap = o->mcap; v.slice_bits = o->slice_bits;
        o->exported = true;
    }
    return peers_install(c, views);
}

extern "C" int kq_peer_disconnect(kq_ctx *c) {
    if (!c) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    for (void *p : c->peer_opened) cudaIpcCloseMemHandle(p);
    c->peer_opened.clear();
    cudaFree(c->d_peers); c->d_peers = nullptr;
    c->exported = false;
    return KQ_OK;
}

// ---- routed validate: queries to the owners, rows back (row E) ------------------------------------------------

extern "C" int kq_vroute_setup(kq_ctx *c, uint64_t chunk, void **d_inbox, void *ipc_inbox, void **d_rows, void *ipc_rows) {
    if (!c || !chunk || chunk > 0xFFFFFFFFull) return KQ_EINVAL;
    if (c->nranks > KQ_MAX_RANKS) return fail(c, KQ_EINVAL, "kq_vroute_setup: too many ranks");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    cudaFree(c->d_vinbox); cudaFree(c->d_vrows); c->d_vinbox = nullptr; c->d_vrows = nullptr; c->vroute_ready = false;
    const size_t qbytes = (size_t)c->nranks * chunk * 16;
    CU(cudaMalloc(&c->d_vinbox, qbytes + KQ_MAX_RANKS * 8));
    CU(cudaMemsetAsync(c->d_vinbox + qbytes, 0, KQ_MAX_RANKS * 8, c->stream));
    // row slab + one answer box per owner (the owners store their answers there in query order, kq_vroute_place sorts them in)
    CU(cudaMalloc(&c->d_vrows, ((size_t)c->nranks + 1) * chunk * sizeof(Base)));
    cudaFree(c->d_vpos); c->d_vpos = nullptr;
    CU(cudaMalloc(&c->d_vpos, (size_t)c->nranks * chunk * sizeof(unsigned)));
    if (!c->d_vcursor) {
        CU(cudaMalloc(&c->d_vcursor, 2 * KQ_MAX_RANKS * 8)); CU(cudaMalloc(&c->d_vacc, 3 * 8));    // cursors, then `sent`
        CU(cudaMemsetAsync(c->d_vcursor, 0, 2 * KQ_MAX_RANKS * 8, c->stream)); CU(cudaMemsetAsync(c->d_vacc, 0, 3 * 8, c->stream));
    }
    CU(cudaStreamSynchronize(c->stream));
    c->vcap = chunk;
    if (d_inbox) *d_inbox = c->d_vinbox;
    if (d_rows) *d_rows = c->d_vrows;
    cudaIpcMemHandle_t h;
    if (ipc_inbox) { CU(cudaIpcGetMemHandle(&h, c->d_vinbox)); std::memcpy(ipc_inbox, &h, sizeof h); }
    if (ipc_rows) { CU(cudaIpcGetMemHandle(&h, c->d_vrows)); std::memcpy(ipc_rows, &h, sizeof h); }
    return KQ_OK;
}

extern "C" int kq_vroute_connect(kq_ctx *c, void *const *inboxes, void *const *rows) {
    if (!c || !inboxes || !rows || !c->d_vinbox) return KQ_EINVAL;
    VRoute &v = c->vroute;
    v = VRoute{};
    v.nranks = c->nranks; v.me = c->rank; v.shift = view(c).shard_shift; v.cap = c->vcap; v.cursor = c->d_vcursor;
    v.sent = c->d_vcursor + KQ_MAX_RANKS; v.poslog = c->d_vpos;
    for (int r = 0; r < c->nranks; ++r) {
        if (!inboxes[r] || !rows[r]) return fail(c, KQ_EINVAL, "kq_vroute_connect: NULL pointer");
        v.inbox[r] = inboxes[r];
        v.counts[r] = reinterpret_cast<unsigned long long *>(static_cast<unsigned char *>(inboxes[r]) + (size_t)c->nranks * c->vcap * 16);
        v.rows[r] = static_cast<Base *>(rows[r]);
    }
    if (v.inbox[c->rank] != c->d_vinbox || v.rows[c->rank] != c->d_vrows) return fail(c, KQ_EINVAL, "kq_vroute_connect: entry [rank] is not this handle's buffer");
    c->vroute_ready = true;
    return KQ_OK;
}

extern "C" int kq_vroute_send_at(kq_ctx *c, const uint64_t *d_bases, const uint64_t *d_invalid, uint64_t base0, uint64_t c0, uint64_t n,
                                uint64_t kcount, uint64_t row0, int last) {
    if (!c || (n && (!d_bases || !d_invalid))) return KQ_EINVAL;
    if (!c->vroute_ready) return fail(c, KQ_ESTATE, "kq_vroute_send: call kq_vroute_connect first");
    if (row0 + n > c->vcap) return fail(c, KQ_EINVAL, "kq_vroute_send: more positions in a round than given to kq_vroute_setup");
    CU(cudaSetDevice(c->device));
    const int t = time_begin(c, 1), t8 = time_begin(c, 8);
    CU(launch_vroute_send(c->stream, view(c), d_bases, d_invalid, base0, c0, n, kcount, c->d_vrows, c->d_vacc, c->vroute, &c->launches, row0, last != 0));
    time_end(c, t8);
    time_end(c, t);
    return KQ_OK;
}
extern "C" int kq_vroute_send(kq_ctx *c, const uint64_t *d_bases, const uint64_t *d_invalid, uint64_t base0, uint64_t c0, uint64_t n, uint64_t kcount) {
    return kq_vroute_send_at(c, d_bases, d_invalid, base0, c0, n, kcount, 0, 1);
}

extern "C" int kq_vroute_answer(kq_ctx *c, uint32_t cutoff) {
    if (!c) return KQ_EINVAL;
    if (!c->vroute_ready) return fail(c, KQ_ESTATE, "kq_vroute_answer: call kq_vroute_connect first");
    CU(cudaSetDevice(c->device));
    const int t = time_begin(c, 1), t9 = time_begin(c, 9);
    CU(launch_vroute_answer(c->stream, view(c), c->d_vinbox, c->vroute.counts[c->rank], (uint64_t)c->nranks * c->vcap, cutoff, c->d_vacc, c->vroute, &c->launches));
    time_end(c, t9);
    time_end(c, t);
    return KQ_OK;
}

// after the barrier that follows kq_vroute_answer: the answers the owners left in this rank's boxes go to their rows
extern "C" int kq_vroute_place(kq_ctx *c) {
    if (!c) return KQ_EINVAL;
    if (!c->vroute_ready) return fail(c, KQ_ESTATE, "kq_vroute_place: call kq_vroute_connect first");
    CU(cudaSetDevice(c->device));
    const int t = time_begin(c, 1), t9 = time_begin(c, 9);
    CU(launch_vroute_place(c->stream, c->vcap, c->vroute, &c->launches));
    time_end(c, t9);
    time_end(c, t);
    return KQ_OK;
}

extern "C" int kq_vroute_rows(kq_ctx *c, kq_base **d_rows) {
    if (!c || !d_rows || !c->d_vrows) return KQ_EINVAL;
    *d_rows = reinterpret_cast<kq_base *>(c->d_vrows);
    return KQ_OK;
}

// rows [0, n) of this rank's slab -> host memory, asynchronously on the handle's stream
extern "C" int kq_vroute_copy_rows(kq_ctx *c, kq_base *host_dst, uint64_t n) {
    if (!c || !host_dst || !c->d_vrows || n > c->vcap) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(host_dst, c->d_vrows, n * sizeof(Base), cudaMemcpyDeviceToHost, c->stream));
    c->d2h += n * sizeof(Base);
    return KQ_OK;
}

// this rank's share of the QV counters since the last reset (sum over all ranks = the three totals)
extern "C" int kq_vroute_counters(kq_ctx *c, kq_qv *acc, int reset) {
    if (!c || !acc || !c->d_vacc) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    unsigned long long h[3];
    CU(cudaMemcpyAsync(h, c->d_vacc, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    if (reset) CU(cudaMemsetAsync(c->d_vacc, 0, sizeof h, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    acc->missing += h[0]; acc->kcount += h[1]; acc->edge_missing += h[2];
    return check_device_error(c);
}

// ---- validate -----------------------------------------------------------------------------------------------

extern "C" double kq_error_rate(uint64_t missing, uint64_t total, int k) {      // src/kreeq.cpp:36-40
    return 1 - pow(1 - (double)missing / total, (double)1 / k);
}
extern "C" double kq_qv_value(uint64_t missing, uint64_t total, int k) {        // src/kreeq.cpp:87,96
    return -10 * log10(kq_error_rate(missing, total, k));
}

extern "C" int kq_dev_validate_packed(kq_ctx *c, const uint64_t *d_bases, const uint64_t *d_invalid,
                                      uint64_t len, uint32_t cutoff, kq_base *d_out, kq_qv *d_acc) {
    if (!c || !d_bases || !d_invalid || !d_out || !d_acc) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    if (c->pend_len || c->ovf_pending) { int rc = kq_sync(c); if (rc) return rc; }
    const uint64_t kcount = len >= (uint64_t)c->k ? len - c->k + 1 : 0;
    const int t = time_begin(c, 1);
    CU(launch_validate(c->stream, view(c), d_bases, d_invalid, 0, 0, len, kcount, cutoff,
                       reinterpret_cast<Base *>(d_out), reinterpret_cast<unsigned long long *>(d_acc), c->d_peers, c->nranks, &c->launches));
    time_end(c, t);
    return KQ_OK;
}

namespace {

constexpr unsigned long long WIDE_CAP = 1ull << 20;                  // rows of one call that do not fit a byte per field
static_assert(sizeof(kq_wide_row) == sizeof(WideRow) && sizeof(kq_wide_row) == 24 && sizeof(kq_base4) == 4, "compact row layouts");

// kq_validate_segment (16-byte DBGbase rows) and kq_validate_segment_compact (4-byte rows + a list of the rare rows that
// do not fit a byte per field) share the staging loop: `out` receives `rb` bytes per position
int validate_impl(kq_ctx *c, const char *seq, uint64_t len, uint32_t cutoff, void *out_v, size_t rb, kq_wide_row *wide, size_t *nwide, kq_qv *acc) {
    unsigned char *out = static_cast<unsigned char *>(out_v);
    const bool compact = rb == 4;
    if (!c || (!seq && len) || (!out && len) || !acc) return KQ_EINVAL;
    if (len == 0) return KQ_OK;
    CU(cudaSetDevice(c->device));
    if (c->pend_len || c->ovf_pending) { int rc = kq_sync(c); if (rc) return rc; }
    const uint64_t k = (uint64_t)c->k;
    const uint64_t kcount = len >= k ? len - k + 1 : 0;                // src/kreeq.cpp:123-126
    CU(cudaMemsetAsync(c->d_acc, 0, 3 * sizeof(unsigned long long), c->stream));
    const uint64_t VCH = 4ull << 20;
    if (compact) {
        if (!c->d_out4[0]) for (int i = 0; i < NBUF; ++i) CU(cudaMalloc(&c->d_out4[i], VCH * 4));
        if (!c->d_wide) { CU(cudaMalloc(&c->d_wide, WIDE_CAP * sizeof(kq_wide_row))); CU(cudaMalloc(&c->d_wide_n, 8)); }
        CU(cudaMemsetAsync(c->d_wide_n, 0, 8, c->stream));
    }                                   // output entries per staged chunk
    // A caller that hands in page-locked memory (cudaHostAlloc / cudaHostRegister) gets the rows copied straight
    // into it; pageable memory goes through the pinned staging slot and one host copy (which, at 16 bytes per
    // base, costs more than the PCIe transfer).
    bool direct = false;
    {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, out) == cudaSuccess) direct = at.type == cudaMemoryTypeHost;
        else cudaGetLastError();
    }
    struct Pending { Stage *s; uint64_t c0, n; };
    Pending pend[NBUF]; int npend = 0;
    auto drain_one = [&](Pending &p) -> int {                          // pinned -> caller memory, on the pool
        int rc = stage_acquire(c, *p.s); if (rc) return rc;
        if (direct) return KQ_OK;
        const size_t grain = 1 << 18, tasks = (p.n + grain - 1) / grain;
        const unsigned char *src = reinterpret_cast<const unsigned char *>(p.s->h_out); unsigned char *dst = out + p.c0 * rb; const uint64_t n = p.n;
        global_pool().parallel_for(tasks, [&](size_t t) {
            const size_t b = t * grain, e = std::min<size_t>(n, b + grain);
            std::memcpy(dst + b * rb, src + b * rb, (e - b) * rb);
        });
        return KQ_OK;
    };
    for (uint64_t c0 = 0; c0 < len; c0 += VCH) {
        const uint64_t n = std::min(VCH, len - c0);
        if (npend == NBUF) { int rc = drain_one(pend[0]); if (rc) return rc; for (int i = 1; i < NBUF; ++i) pend[i - 1] = pend[i]; --npend; }
        Stage &s = c->st[c->next_stage]; c->next_stage = (c->next_stage + 1) % NBUF;
        int rc = stage_acquire(c, s); if (rc) return rc;
        const size_t need = packed_words(std::max(CHUNK, VCH + k + 128));
        rc = stage_reserve(c, s, need, need, VCH); if (rc) return rc;
        // slice of the segment this chunk needs: one base before c0 (64-aligned down) .. k bases after the last start
        const uint64_t a0 = c0 == 0 ? 0 : ((c0 - 1) & ~63ull);
        const uint64_t a1 = std::min<uint64_t>(len, c0 + n + k);
        const size_t words = packed_words(a1 - a0);
        pack_sequence(seq + a0, a1 - a0, s.h_bases, s.h_inv, &global_pool());
        CU(cudaMemcpyAsync(s.d_bases, s.h_bases, words * 8, cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemcpyAsync(s.d_inv, s.h_inv, (words / 2 + 1) * 8, cudaMemcpyHostToDevice, c->stream));
        c->h2d += words * 8 + (words / 2 + 1) * 8;
        const int t = time_begin(c, 1);
        CU(launch_validate(c->stream, view(c), s.d_bases, s.d_inv, a0, c0, n, kcount, cutoff, s.d_out, c->d_acc, c->d_peers, c->nranks, &c->launches));
        time_end(c, t);
        // results leave on the copy stream, so the next chunk's kernel runs under this chunk's transfer
        CU(cudaEventRecord(s.kernel_done, c->stream));
        CU(cudaStreamWaitEvent(c->copy_stream, s.kernel_done, 0));
        const void *d_rows = s.d_out;
        if (compact) {                              // 16-byte rows -> 4-byte rows (+ the wide list) on the device: a quarter of the bytes cross PCIe
            uint32_t *d4 = c->d_out4[&s - c->st];
            CU(launch_compact_rows(c->stream, s.d_out, n, c0, d4, c->d_wide, WIDE_CAP, c->d_wide_n, &c->launches));
            CU(cudaEventRecord(s.kernel_done, c->stream));
            CU(cudaStreamWaitEvent(c->copy_stream, s.kernel_done, 0));
            d_rows = d4;
        }
        CU(cudaMemcpyAsync(direct ? (void *)(out + c0 * rb) : (void *)s.h_out, d_rows, n * rb, cudaMemcpyDeviceToHost, c->copy_stream));
        c->d2h += n * rb;
        CU(cudaEventRecord(s.done, c->copy_stream)); s.busy = true;
        pend[npend++] = Pending{&s, c0, n};
    }
    for (int i = 0; i < npend; ++i) { int rc = drain_one(pend[i]); if (rc) return rc; }
    unsigned long long h[3];
    CU(cudaMemcpyAsync(h, c->d_acc, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->d2h += sizeof h;
    int rc = check_device_error(c); if (rc) return rc;
    acc->missing += h[0]; acc->kcount += h[1]; acc->edge_missing += h[2];
    if (compact) {
        unsigned long long nw = 0;
        CU(cudaMemcpy(&nw, c->d_wide_n, 8, cudaMemcpyDeviceToHost));
        const size_t cap = nwide ? *nwide : 0;
        if (nwide) *nwide = (size_t)nw;
        if (nw > WIDE_CAP || nw > cap)
            return fail(c, KQ_ENOMEM, "kq_validate_segment_compact: " + std::to_string(nw) + " rows have a value above 254 (list capacity " +
                                      std::to_string(std::min<size_t>(cap, WIDE_CAP)) + "): use kq_validate_segment or a larger list");
        if (nw) { CU(cudaMemcpy(wide, c->d_wide, nw * sizeof(kq_wide_row), cudaMemcpyDeviceToHost)); c->d2h += nw * sizeof(kq_wide_row); }
    }
    return KQ_OK;
}

}  // namespace

extern "C" int kq_validate_segment(kq_ctx *c, const char *seq, uint64_t len, uint32_t cutoff, kq_base *out, kq_qv *acc) {
    return validate_impl(c, seq, len, cutoff, out, sizeof(kq_base), nullptr, nullptr, acc);
}

extern "C" int kq_expand_rows(const kq_base4 *in, uint64_t len, const kq_wide_row *wide, size_t nwide, kq_base *out) {
    if ((len && (!in || !out)) || (nwide && !wide)) return KQ_EINVAL;
    const size_t grain = 1 << 18, tasks = (size_t)((len + grain - 1) / grain);
    global_pool().parallel_for(tasks, [&](size_t t) {
        const uint64_t b = (uint64_t)t * grain, e = std::min<uint64_t>(len, b + grain);
        for (uint64_t i = b; i < e; ++i) {
            kq_base r; r.fw = in[i].fw; r.bw = in[i].bw; r.cov = in[i].cov; r.isFw = in[i].flags & 1; r.pad[0] = r.pad[1] = r.pad[2] = 0;
            out[i] = r;
        }
    });
    for (size_t j = 0; j < nwide; ++j) { if (wide[j].pos >= len) return KQ_EINVAL; out[wide[j].pos] = wide[j].row; }
    return KQ_OK;
}

extern "C" int kq_validate_segment_compact(kq_ctx *c, const char *seq, uint64_t len, uint32_t cutoff, kq_base4 *out,
                                           kq_wide_row *wide, size_t *nwide, kq_qv *acc) {
    if (len && (!nwide || (*nwide && !wide))) return KQ_EINVAL;
    return validate_impl(c, seq, len, cutoff, out, sizeof(kq_base4), wide, nwide, acc);
}

extern "C" int kq_lookup_batch(kq_ctx *c, const uint64_t *keys, size_t n, kq_rec32 *rec, uint8_t *found) {
    if (!c || (n && (!keys || !rec || !found))) return KQ_EINVAL;
    if (!n) return KQ_OK;
    CU(cudaSetDevice(c->device));
    if (c->pend_len || c->ovf_pending) { int rc = kq_sync(c); if (rc) return rc; }
    struct Scratch {                                                   // freed on every way out, error returns included
        uint64_t *dk = nullptr; Rec32 *dr = nullptr; uint8_t *df = nullptr;
        ~Scratch() { cudaFree(dk); cudaFree(dr); cudaFree(df); }
    } scratch;
    uint64_t *&dk = scratch.dk; Rec32 *&dr = scratch.dr; uint8_t *&df = scratch.df;
    CU(cudaMalloc(&dk, n * 8)); CU(cudaMalloc(&dr, n * sizeof(Rec32))); CU(cudaMalloc(&df, n));
    CU(cudaMemcpyAsync(dk, keys, n * 8, cudaMemcpyHostToDevice, c->stream));
    CU(launch_lookup(c->stream, view(c), dk, n, dr, df, &c->launches));
    std::vector<Rec32> hr(n);
    CU(cudaMemcpyAsync(hr.data(), dr, n * sizeof(Rec32), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(found, df, n, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->h2d += n * 8; c->d2h += n * (sizeof(Rec32) + 1);
    for (size_t i = 0; i < n; ++i) if (found[i]) std::memcpy(&rec[i], &hr[i], sizeof(Rec32));
    return check_device_error(c);
}

// ---- stats / export ---------------------------------------------------------------------------------------------

namespace {

// copy the literal maps32 table to the host
int fetch_map32(kq_ctx *c, std::vector<SideSlot> &side) {
    side.resize(c->mcap);
    CU(cudaMemcpyAsync(side.data(), c->M, c->mcap * sizeof(SideSlot), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->d2h += c->mcap * sizeof(SideSlot);
    return KQ_OK;
}

inline uint32_t clamp_u32(unsigned long long v) { return v > KQ_LARGEST ? KQ_LARGEST : (uint32_t)v; }

struct FullStats { kq_stats s; std::vector<std::pair<uint64_t, uint64_t>> hist; };

// DBG::summary over all maps + DBG::DBstats, src/graph-builder.cpp:240-295
int compute_stats(kq_ctx *c, FullStats &fs) {
    int rc = kq_sync(c); if (rc) return rc;
    DevStats *d_ds = nullptr; unsigned long long *d_hist = nullptr, *d_big = nullptr;
    const unsigned long long big_cap = c->rcap + 16;                  // a coverage >= 65536 has a remainder record
    CU(cudaMalloc(&d_ds, sizeof(DevStats))); CU(cudaMalloc(&d_hist, HIST_BINS * 8)); CU(cudaMalloc(&d_big, big_cap * 8));
    CU(cudaMemsetAsync(d_ds, 0, sizeof(DevStats), c->stream));
    CU(cudaMemsetAsync(d_hist, 0, HIST_BINS * 8, c->stream));
    CU(launch_stats(c->stream, view(c), d_ds, d_hist, d_big, big_cap, &c->launches));
    DevStats ds; std::vector<unsigned long long> hist(HIST_BINS);
    CU(cudaMemcpyAsync(&ds, d_ds, sizeof ds, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(hist.data(), d_hist, HIST_BINS * 8, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (ds.big_n > big_cap) { cudaFree(d_ds); cudaFree(d_hist); cudaFree(d_big); return fail(c, KQ_ENOMEM, "stats: high-coverage list overflow"); }
    std::vector<unsigned long long> big(ds.big_n);
    if (ds.big_n) CU(cudaMemcpy(big.data(), d_big, ds.big_n * 8, cudaMemcpyDeviceToHost));
    cudaFree(d_ds); cudaFree(d_hist); cudaFree(d_big);
    c->d2h += sizeof ds + HIST_BINS * 8 + ds.big_n * 8;

    std::vector<std::pair<uint64_t, uint64_t>> extra;                 // coverages outside the device histogram
    uint64_t total = ds.total, uniq = ds.unique, distinct = ds.distinct, edges = ds.edges;
    for (auto v : big) { total += v; extra.emplace_back(v, 1); }
    // every maps32 entry counts (src/graph-builder.cpp:259-266)
    std::vector<SideSlot> side; rc = fetch_map32(c, side); if (rc) return rc;
    for (const auto &q : side) {
        if (!q.tag) continue;
        const uint32_t cov = clamp_u32(q.cov);
        for (int w = 0; w < 4; ++w) edges += (q.fw[w] > 0 || q.bw[w] > 0);
        ++distinct; total += cov;
        if (cov < (uint32_t)HIST_BINS) ++hist[cov]; else extra.emplace_back(cov, 1);
    }
    fs.hist.clear();
    for (int i = 0; i < HIST_BINS; ++i) if (hist[i]) fs.hist.emplace_back((uint64_t)i, (uint64_t)hist[i]);
    std::sort(extra.begin(), extra.end());
    for (auto &e : extra) { if (!fs.hist.empty() && fs.hist.back().first == e.first) fs.hist.back().second += e.second; else fs.hist.push_back(e); }
    fs.s.total = total; fs.s.unique = uniq; fs.s.distinct = distinct; fs.s.edges = edges;
    fs.s.missing = (uint64_t)(pow(4, c->k) - distinct);               // evaluated in double, src/graph-builder.cpp:286
    return KQ_OK;
}

// all occupied T slots as XRec (unsorted) + the literal maps32 table
int fetch_records(kq_ctx *c, std::vector<XRec> &recs, std::vector<SideSlot> &side) {
    int rc = kq_sync(c); if (rc) return rc;
    XRec *d_out = nullptr; unsigned long long *d_n = nullptr;
    const unsigned long long cap = c->occ_known + 1;
    CU(cudaMalloc(&d_out, cap * sizeof(XRec))); CU(cudaMalloc(&d_n, 8));
    CU(cudaMemsetAsync(d_n, 0, 8, c->stream));
    CU(launch_export(c->stream, view(c), d_out, cap, d_n, &c->launches));
    unsigned long long n = 0;
    CU(cudaMemcpyAsync(&n, d_n, 8, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (n > cap) { cudaFree(d_out); cudaFree(d_n); return fail(c, KQ_ECUDA, "export: occupancy counter out of sync"); }
    recs.resize(n);
    if (n) CU(cudaMemcpy(recs.data(), d_out, n * sizeof(XRec), cudaMemcpyDeviceToHost));
    cudaFree(d_out); cudaFree(d_n);
    c->d2h += n * sizeof(XRec);
    return fetch_map32(c, side);
}

const SideSlot *host_side_find(const std::vector<SideSlot> &side, uint64_t key) {
    const uint64_t scap = side.size();
    uint64_t i = (uint64_t)(((unsigned __int128)mix64(key ^ 0x9E3779B97F4A7C15ULL) * scap) >> 64);
    for (uint64_t n = 0; n < scap; ++n) {
        if (side[i].tag == key + 1) return &side[i];
        if (side[i].tag == 0) return nullptr;
        if (++i == scap) i = 0;
    }
    return nullptr;
}

}  // namespace

extern "C" int kq_stats_get(kq_ctx *c, kq_stats *out) {
    if (!c || !out) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    FullStats fs; int rc = compute_stats(c, fs); if (rc) return rc;
    *out = fs.s;
    return KQ_OK;
}

extern "C" int kq_hist(kq_ctx *c, uint64_t *cov, uint64_t *n, size_t *len) {
    if (!c || !len) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    FullStats fs; int rc = compute_stats(c, fs); if (rc) return rc;
    const size_t cap = *len; *len = fs.hist.size();
    if (cov && n) for (size_t i = 0; i < fs.hist.size() && i < cap; ++i) { cov[i] = fs.hist[i].first; n[i] = fs.hist[i].second; }
    return KQ_OK;
}

extern "C" int kq_export(kq_ctx *c, uint64_t *keys, kq_rec32 *recs, size_t *n) {
    if (!c || !n) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    std::vector<XRec> xs; std::vector<SideSlot> side;
    int rc = fetch_records(c, xs, side); if (rc) return rc;
    std::sort(xs.begin(), xs.end(), [](const XRec &a, const XRec &b) { return a.key < b.key; });
    const size_t cap = *n; size_t m = 0;
    for (const auto &x : xs) {
        Rec32 r = x.r;
        if (x.flags & F_PLACEHOLDER) {                                 // resolve through maps32 (src/kreeq.cpp:158-166)
            const SideSlot *q = host_side_find(side, x.key);
            if (!q) return fail(c, KQ_ECORRUPT, "Error: int32 map missing 255 value from int8 map");
            for (int w = 0; w < 4; ++w) { r.fw[w] = clamp_u32(q->fw[w]); r.bw[w] = clamp_u32(q->bw[w]); }
            r.cov = clamp_u32(q->cov);
        }
        if (keys && recs && m < cap) { keys[m] = x.key; std::memcpy(&recs[m], &r, sizeof r); }
        ++m;
    }
    *n = m;
    return KQ_OK;
}

// ---- persistence -------------------------------------------------------------------------------------------------

namespace {

std::string map_path(const std::string &dir, int m) { return dir + "/.map." + std::to_string(m) + ".bin"; }
void owned_maps(const kq_ctx *c, int &lo, int &hi) { const int per = 128 / c->nranks; lo = c->rank * per; hi = lo + per; }

// split the resident database into the reference's on-disk representation
int gather_for_save(kq_ctx *c, std::vector<std::vector<Rec8Entry>> &maps, std::vector<Rec32Entry> &hc) {
    std::vector<XRec> xs; std::vector<SideSlot> side;
    int rc = fetch_records(c, xs, side); if (rc) return rc;
    maps.assign(128, {});
    for (const auto &x : xs) {
        Rec8Entry e; e.key = x.key;
        if (!(x.flags & F_PLACEHOLDER) && x.r.cov <= 254) {            // exact 8-bit record
            for (int w = 0; w < 4; ++w) { e.fw[w] = (uint8_t)x.r.fw[w]; e.bw[w] = (uint8_t)x.r.bw[w]; }
            e.cov = (uint8_t)x.r.cov;
        } else {                                                       // cov == 255 placeholder (fw/bw unobservable: zeros)
            std::memset(e.fw, 0, 4); std::memset(e.bw, 0, 4); e.cov = 255;
            if (!(x.flags & F_PLACEHOLDER)) {                          // >= 255 instances counted here: 32-bit record
                Rec32Entry h; h.key = x.key; std::memcpy(h.fw, x.r.fw, 16); std::memcpy(h.bw, x.r.bw, 16); h.cov = x.r.cov;
                hc.push_back(h);
            }
        }
        maps[x.key % 128].push_back(e);
    }
    for (const auto &q : side) {                                       // literal maps32 entries of a loaded / merged database
        if (!q.tag) continue;
        Rec32Entry h; h.key = q.tag - 1;
        for (int w = 0; w < 4; ++w) { h.fw[w] = clamp_u32(q.fw[w]); h.bw[w] = clamp_u32(q.bw[w]); }
        h.cov = clamp_u32(q.cov);
        hc.push_back(h);
    }
    return KQ_OK;
}

std::string hc_part_path(const std::string &dir, int rank) { return dir + "/.hc.part." + std::to_string(rank); }

int write_hc_part(kq_ctx *c, const std::string &dir, const std::vector<Rec32Entry> &hc) {
    FILE *f = std::fopen(hc_part_path(dir, c->rank).c_str(), "wb");
    if (!f) return fail(c, KQ_EIO, "cannot write " + hc_part_path(dir, c->rank));
    const uint64_t n = hc.size();
    bool ok = std::fwrite(&n, 8, 1, f) == 1 && (n == 0 || std::fwrite(hc.data(), sizeof(Rec32Entry), n, f) == n);
    ok = std::fclose(f) == 0 && ok;
    return ok ? KQ_OK : fail(c, KQ_EIO, "short write on " + hc_part_path(dir, c->rank));
}

int save_maps_gpu(kq_ctx *c, const std::string &dir, std::vector<Rec32Entry> &hc);
int load_maps_gpu(kq_ctx *c, const std::string &dir, XRec **d_xs, uint64_t *n_out);

// KQ_DB_HOST=1 keeps the file codec on the host (the first implementation; useful to cross-check the device one)
bool db_on_host() { return tuning().db_host != 0; }

int save_impl(kq_ctx *c, const char *dir_c, bool maps_too) {
    if (!c || !dir_c) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    const std::string dir(dir_c);
    mkdir(dir.c_str(), 0777);
    std::vector<Rec32Entry> hc;
    if (maps_too && !db_on_host()) {
        int rc = save_maps_gpu(c, dir, hc); if (rc) return rc;
        std::vector<SideSlot> side; rc = fetch_map32(c, side); if (rc) return rc;
        for (const auto &q : side) {                                   // literal maps32 entries of a loaded / merged database
            if (!q.tag) continue;
            Rec32Entry h; h.key = q.tag - 1;
            for (int w = 0; w < 4; ++w) { h.fw[w] = clamp_u32(q.fw[w]); h.bw[w] = clamp_u32(q.bw[w]); }
            h.cov = clamp_u32(q.cov);
            hc.push_back(h);
        }
        if (c->rank == 0) { const std::string e = write_index(dir, c->k); if (!e.empty()) return fail(c, KQ_EIO, e); }
    } else {
        std::vector<std::vector<Rec8Entry>> maps;
        int rc = gather_for_save(c, maps, hc); if (rc) return rc;
        if (maps_too) {
            int lo, hi; owned_maps(c, lo, hi);
            std::vector<std::string> errs(128);
            global_pool().parallel_for((size_t)(hi - lo), [&](size_t i) { errs[lo + i] = write_map8(map_path(dir, lo + (int)i), maps[lo + i]); });
            for (const auto &e : errs) if (!e.empty()) return fail(c, KQ_EIO, e);
            if (c->rank == 0) { const std::string e = write_index(dir, c->k); if (!e.empty()) return fail(c, KQ_EIO, e); }
        }
    }
    if (c->nranks == 1 && maps_too) {
        const std::string e = write_map32(dir + "/.map.hc.bin", hc);
        return e.empty() ? KQ_OK : fail(c, KQ_EIO, e);
    }
    return write_hc_part(c, dir, hc);
}

// upload literal maps32 entries as side table M
int upload_map32(kq_ctx *c, const std::vector<Rec32Entry> &hc, uint64_t min_scap) {
    uint64_t scap = std::max<uint64_t>(min_scap, MIN_SCAP);
    while (scap < 4 * hc.size()) scap *= 2;
    std::vector<SideSlot> side(scap);
    std::memset(side.data(), 0, scap * sizeof(SideSlot));
    for (const auto &h : hc) {
        uint64_t i = (uint64_t)(((unsigned __int128)mix64(h.key ^ 0x9E3779B97F4A7C15ULL) * scap) >> 64);
        while (side[i].tag && side[i].tag != h.key + 1) if (++i == scap) i = 0;
        SideSlot &q = side[i];
        q.tag = h.key + 1; q.cov = h.cov;
        for (int w = 0; w < 4; ++w) { q.fw[w] = h.fw[w]; q.bw[w] = h.bw[w]; }
    }
    CU(cudaFree(c->M)); c->M = nullptr;
    CU(cudaMalloc(&c->M, scap * sizeof(SideSlot))); c->mcap = scap;
    CU(cudaMemcpyAsync(c->M, side.data(), scap * sizeof(SideSlot), cudaMemcpyHostToDevice, c->stream));
    const unsigned long long n = hc.size();
    CU(cudaMemcpyAsync(c->d_occ + 2, &n, 8, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->h2d += scap * sizeof(SideSlot);
    return KQ_OK;
}

// read the 8-bit maps this shard owns of one database
int read_owned_maps(kq_ctx *c, const std::string &dir, std::vector<Rec8Entry> &all) {
    int lo, hi; owned_maps(c, lo, hi);
    std::vector<std::vector<Rec8Entry>> per((size_t)(hi - lo)); std::vector<std::string> errs((size_t)(hi - lo));
    global_pool().parallel_for((size_t)(hi - lo), [&](size_t i) { errs[i] = read_map8(map_path(dir, lo + (int)i), per[i]); });
    for (const auto &e : errs) if (!e.empty()) return fail(c, KQ_EIO, e);
    size_t n = 0; for (auto &v : per) n += v.size();
    all.clear(); all.reserve(n);
    for (auto &v : per) { all.insert(all.end(), v.begin(), v.end()); std::vector<Rec8Entry>().swap(v); }
    return KQ_OK;
}

int read_owned_hc(kq_ctx *c, const std::string &dir, std::vector<Rec32Entry> &hc) {
    std::vector<Rec32Entry> all;
    const std::string e = read_map32(dir + "/.map.hc.bin", all);
    if (!e.empty()) return fail(c, KQ_EIO, e);
    int lo, hi; owned_maps(c, lo, hi);
    for (const auto &h : all) { const int m = (int)(h.key % 128); if (m >= lo && m < hi) hc.push_back(h); }
    return KQ_OK;
}


// ---- .kreeq files through device-built images -------------------------------------------------------

// per-thread pinned staging buffer + stream for the file <-> device copies of the pool workers
struct FileStage { char *buf = nullptr; cudaStream_t st = nullptr; int device = -1; };
constexpr size_t FILE_CHUNK = 32u << 20;
constexpr uint64_t kDumpVersion = 0xFFFFFFFFFFFFFFF5ULL;              // SIZE_MAX - 10, phmap_dump.h:47-48
FileStage *file_stage(int device) {
    // one staging buffer per pool thread (portable: every GPU can copy to / from it) and one stream per GPU the thread
    // has worked for -- a multi-GPU handle sends the same threads from shard to shard
    constexpr int MAXDEV = 64;
    thread_local char *buf = nullptr;
    thread_local FileStage st[MAXDEV];
    if (device < 0 || device >= MAXDEV) return nullptr;
    if (cudaSetDevice(device) != cudaSuccess) return nullptr;
    if (!buf && cudaHostAlloc(&buf, FILE_CHUNK, cudaHostAllocPortable) != cudaSuccess) { buf = nullptr; return nullptr; }
    FileStage &s = st[device];
    if (!s.st) {
        if (cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking) != cudaSuccess) { s.st = nullptr; return nullptr; }
        s.device = device;
    }
    s.buf = buf;
    return &s;
}

// occupied slots of the table as records, left on the device
int export_device(kq_ctx *c, XRec **d_xs, uint64_t *n_out) {
    int rc = kq_sync(c); if (rc) return rc;
    unsigned long long *d_n = nullptr;
    const unsigned long long cap = c->occ_known + 1;
    CU(cudaMalloc(d_xs, cap * sizeof(XRec))); CU(cudaMalloc(&d_n, 8));
    CU(cudaMemsetAsync(d_n, 0, 8, c->stream));
    CU(launch_export(c->stream, view(c), *d_xs, cap, d_n, &c->launches));
    unsigned long long n = 0;
    CU(cudaMemcpyAsync(&n, d_n, 8, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaFree(d_n);
    if (n > cap) { cudaFree(*d_xs); *d_xs = nullptr; return fail(c, KQ_ECUDA, "export: occupancy counter out of sync"); }
    *n_out = n;
    return KQ_OK;
}

// byte layout of the owned .map.<m>.bin files inside one device image (each file starts 256-byte aligned)
struct ImageLayout { std::vector<DbBucket> buckets; std::vector<uint64_t> file_off, file_size; uint64_t total = 0; };

void layout_from_counts(const std::vector<unsigned long long> &counts, int lo, int hi, ImageLayout &L) {
    L.buckets.assign(128 * 256, DbBucket{0, 0, 0, 0, 0});
    L.file_off.assign(128, 0); L.file_size.assign(128, 0);
    uint64_t at = 0;
    for (int m = lo; m < hi; ++m) {
        L.file_off[m] = at;
        uint64_t cur = 8;                                              // u64 submap count
        for (int s = 0; s < 256; ++s) {
            DbBucket &b = L.buckets[(size_t)m * 256 + s];
            b.n = counts[(size_t)m * 256 + s];
            b.hdr_off = at + cur; cur += 24;
            if (b.n) {
                b.cap = db_capacity_for(b.n);
                b.ctrl_off = at + cur; cur += b.cap + 17;
                b.slot_off = at + cur; cur += b.cap * 24 + 8;           // slots + growth_left
            }
        }
        L.file_size[m] = cur;
        at = (at + cur + 255) & ~255ull;
    }
    L.total = at;
}

// write the 8-bit maps this shard owns from a device-built image; hc receives the 32-bit spill records
int save_maps_gpu(kq_ctx *c, const std::string &dir, std::vector<Rec32Entry> &hc) {
    XRec *d_xs = nullptr; uint64_t n = 0;
    int rc = export_device(c, &d_xs, &n); if (rc) return rc;
    int lo, hi; owned_maps(c, lo, hi);
    unsigned long long *d_counts = nullptr, *d_hc_n = nullptr, *d_pos = nullptr; XRec *d_hc = nullptr;
    unsigned char *d_img = nullptr; DbBucket *d_buckets = nullptr;
    auto cleanup = [&] { cudaFree(d_xs); cudaFree(d_counts); cudaFree(d_hc_n); cudaFree(d_pos); cudaFree(d_hc); cudaFree(d_img); cudaFree(d_buckets); };
#define CUX(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return fail(c, e_ == cudaErrorMemoryAllocation ? KQ_ENOMEM : KQ_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } } while (0)
    CUX(cudaMalloc(&d_counts, 128 * 256 * 8)); CUX(cudaMalloc(&d_hc_n, 8));
    std::vector<unsigned long long> counts(128 * 256);
    unsigned long long hc_cap = std::min<unsigned long long>(n + 1, 1ull << 20), hc_n = 0;
    for (;;) {                                                          // second pass only if the spill list was too small
        cudaFree(d_hc); d_hc = nullptr;
        CUX(cudaMalloc(&d_hc, hc_cap * sizeof(XRec)));
        CUX(cudaMemsetAsync(d_counts, 0, 128 * 256 * 8, c->stream)); CUX(cudaMemsetAsync(d_hc_n, 0, 8, c->stream));
        CUX(launch_db_count(c->stream, d_xs, n, lo, hi, d_counts, d_hc, hc_cap, d_hc_n, &c->launches));
        CUX(cudaMemcpyAsync(counts.data(), d_counts, 128 * 256 * 8, cudaMemcpyDeviceToHost, c->stream));
        CUX(cudaMemcpyAsync(&hc_n, d_hc_n, 8, cudaMemcpyDeviceToHost, c->stream));
        CUX(cudaStreamSynchronize(c->stream));
        if (hc_n <= hc_cap) break;
        hc_cap = hc_n;
    }
    if (hc_n) {
        std::vector<XRec> h(hc_n);
        CUX(cudaMemcpy(h.data(), d_hc, hc_n * sizeof(XRec), cudaMemcpyDeviceToHost));
        for (const auto &x : h) { Rec32Entry e; e.key = x.key; std::memcpy(e.fw, x.r.fw, 16); std::memcpy(e.bw, x.r.bw, 16); e.cov = x.r.cov; hc.push_back(e); }
    }
    ImageLayout L; layout_from_counts(counts, lo, hi, L);
    CUX(cudaMalloc(&d_img, std::max<uint64_t>(L.total, 256)));
    CUX(cudaMemsetAsync(d_img, 0, std::max<uint64_t>(L.total, 256), c->stream));
    CUX(cudaMalloc(&d_buckets, L.buckets.size() * sizeof(DbBucket)));
    CUX(cudaMemcpyAsync(d_buckets, L.buckets.data(), L.buckets.size() * sizeof(DbBucket), cudaMemcpyHostToDevice, c->stream));
    CUX(cudaMalloc(&d_pos, (n + 1) * 8));
    CUX(launch_db_build(c->stream, d_xs, n, lo, hi, d_img, d_buckets, d_pos, &c->launches));
    CUX(cudaStreamSynchronize(c->stream));
#undef CUX
    rc = check_device_error(c); if (rc) { cleanup(); return rc; }
    // one pool task per file: device image -> pinned chunk -> fwrite
    std::vector<std::string> errs((size_t)(hi - lo));
    const int device = c->device;
    global_pool().parallel_for((size_t)(hi - lo), [&](size_t i) {
        const int m = lo + (int)i;
        FileStage *fs = file_stage(device);
        if (!fs) { errs[i] = "cannot allocate a pinned staging buffer"; return; }
        const std::string path = map_path(dir, m);
        FILE *f = std::fopen(path.c_str(), "wb");
        if (!f) { errs[i] = "cannot open " + path + " for writing"; return; }
        for (uint64_t off = 0; off < L.file_size[m]; off += FILE_CHUNK) {
            const size_t len = (size_t)std::min<uint64_t>(FILE_CHUNK, L.file_size[m] - off);
            if (cudaMemcpyAsync(fs->buf, d_img + L.file_off[m] + off, len, cudaMemcpyDeviceToHost, fs->st) != cudaSuccess ||
                cudaStreamSynchronize(fs->st) != cudaSuccess) { errs[i] = "device to host copy failed for " + path; break; }
            if (std::fwrite(fs->buf, 1, len, f) != len) { errs[i] = "short write on " + path; break; }
        }
        if (std::fclose(f) != 0 && errs[i].empty()) errs[i] = "short write on " + path;
    });
    c->d2h += L.total;
    cleanup();
    for (const auto &e : errs) if (!e.empty()) return fail(c, KQ_EIO, e);
    return KQ_OK;
}

// upload the owned .map.<m>.bin files of a database and decode them on the device into raw records
int load_maps_gpu(kq_ctx *c, const std::string &dir, XRec **d_xs, uint64_t *n_out) {
    int lo, hi; owned_maps(c, lo, hi);
    *d_xs = nullptr; *n_out = 0;
    std::vector<uint64_t> file_off(128, 0), file_size(128, 0);
    uint64_t total = 0;
    for (int m = lo; m < hi; ++m) {
        struct stat st;
        if (stat(map_path(dir, m).c_str(), &st) != 0) return fail(c, KQ_EIO, "cannot open " + map_path(dir, m));
        file_off[m] = total; file_size[m] = (uint64_t)st.st_size;
        total = (total + file_size[m] + 255) & ~255ull;
    }
    unsigned char *d_img = nullptr; DbBucket *d_buckets = nullptr; unsigned long long *d_n = nullptr, *d_band = nullptr;
    CU(cudaMalloc(&d_img, std::max<uint64_t>(total, 256)));
    std::vector<DbBucket> buckets((size_t)(hi - lo) * 256, DbBucket{0, 0, 0, 0, 0});
    std::vector<std::string> errs((size_t)(hi - lo));
    const int device = c->device;
    global_pool().parallel_for((size_t)(hi - lo), [&](size_t i) {
        const int m = lo + (int)i;
        const std::string path = map_path(dir, m);
        FileStage *fs = file_stage(device);
        if (!fs) { errs[i] = "cannot allocate a pinned staging buffer"; return; }
        FILE *f = std::fopen(path.c_str(), "rb");
        if (!f) { errs[i] = "cannot open " + path; return; }
        // submap directory: three or four u64 per submap, the rest is skipped (phmap_dump.h:60-104)
        auto rd64 = [&](uint64_t at, uint64_t &v) { return at + 8 <= file_size[m] && fseeko(f, (off_t)at, SEEK_SET) == 0 && std::fread(&v, 8, 1, f) == 1; };
        uint64_t cur = 0, nsub = 0;
        if (!rd64(cur, nsub) || nsub != 256) { errs[i] = path + ": submap count is not 256"; std::fclose(f); return; }
        cur += 8;
        for (int s = 0; s < 256 && errs[i].empty(); ++s) {
            uint64_t ver = 0, n = 0, cap = 0;
            bool ok = rd64(cur, ver); cur += 8;
            if (ok && ver < kDumpVersion) n = ver; else { ok = ok && rd64(cur, n); cur += 8; }   // phmap_dump.h:77-83
            ok = ok && rd64(cur, cap); cur += 8;
            if (!ok) { errs[i] = path + ": truncated"; break; }
            DbBucket &b = buckets[i * 256 + s];
            b.n = n; b.cap = cap;
            if (!n) continue;
            b.ctrl_off = file_off[m] + cur; cur += cap + 17;
            b.slot_off = file_off[m] + cur; cur += cap * 24 + (ver >= kDumpVersion ? 8 : 0);
            if (cur > file_size[m]) errs[i] = path + ": truncated submap";
        }
        if (errs[i].empty() && fseeko(f, 0, SEEK_SET) != 0) errs[i] = "cannot read " + path;
        for (uint64_t off = 0; off < file_size[m] && errs[i].empty(); off += FILE_CHUNK) {
            const size_t len = (size_t)std::min<uint64_t>(FILE_CHUNK, file_size[m] - off);
            if (std::fread(fs->buf, 1, len, f) != len) { errs[i] = "cannot read " + path; break; }
            if (cudaMemcpyAsync(d_img + file_off[m] + off, fs->buf, len, cudaMemcpyHostToDevice, fs->st) != cudaSuccess ||
                cudaStreamSynchronize(fs->st) != cudaSuccess) errs[i] = "host to device copy failed for " + path;
        }
        std::fclose(f);
    });
    for (const auto &e : errs) if (!e.empty()) { cudaFree(d_img); return fail(c, KQ_EIO, e); }
    c->h2d += total;
    uint64_t n = 0; for (const auto &b : buckets) n += b.n;
    auto cleanup = [&] { cudaFree(d_img); cudaFree(d_buckets); cudaFree(d_n); cudaFree(d_band); };
#define CUX(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); cudaFree(*d_xs); *d_xs = nullptr; return fail(c, e_ == cudaErrorMemoryAllocation ? KQ_ENOMEM : KQ_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } } while (0)
    CUX(cudaMalloc(&d_buckets, buckets.size() * sizeof(DbBucket))); CUX(cudaMalloc(&d_n, 8));
    CUX(cudaMalloc(d_xs, (n + 1) * sizeof(XRec)));
    CUX(cudaMemcpyAsync(d_buckets, buckets.data(), buckets.size() * sizeof(DbBucket), cudaMemcpyHostToDevice, c->stream));
    CUX(cudaMemsetAsync(d_n, 0, 8, c->stream));
    const int t7 = time_begin(c, 7);
    if (tuning().db_ordered && n >= (1u << 16)) {
        // records per band of the slice hash, exclusive scan, decode into band order (= table order)
        constexpr unsigned NB = 1u << KQ_DB_ORDER_BITS;
        CUX(cudaMalloc(&d_band, 2 * NB * 8));
        CUX(cudaMemsetAsync(d_band, 0, NB * 8, c->stream));
        CUX(launch_db_band_hist(c->stream, d_img, d_buckets, (unsigned)buckets.size(), d_band, &c->launches));
        CUX(launch_slice_offsets(c->stream, NB, d_band, d_band + NB, nullptr, d_n, nullptr, &c->launches));
        CUX(launch_db_decode_ordered(c->stream, d_img, d_buckets, (unsigned)buckets.size(), *d_xs, n, d_band + NB, &c->launches));
    } else {
        CUX(launch_db_decode(c->stream, d_img, d_buckets, (unsigned)buckets.size(), *d_xs, n, d_n, &c->launches));
    }
    time_end(c, t7);
    unsigned long long got = 0;
    CUX(cudaMemcpyAsync(&got, d_n, 8, cudaMemcpyDeviceToHost, c->stream));
    CUX(cudaStreamSynchronize(c->stream));
#undef CUX
    cleanup();
    if (got != n) { cudaFree(*d_xs); *d_xs = nullptr; return fail(c, KQ_EIO, dir + ": submap sizes do not match their control bytes"); }
    *n_out = n;
    return KQ_OK;
}

enum class Feed { Put, MergeRef, AddLogical };

// run device-resident records through one of the table kernels, in sub-batches small enough that neither
// the table nor a side table can fill up inside a launch
int feed_device(kq_ctx *c, const XRec *d, uint64_t total, Feed mode) {
    for (uint64_t off = 0; off < total;) {
        const uint64_t n = std::min<uint64_t>(total - off, 64ull << 20);
        // only records with a field above a slot lane (or placeholders) can create side-table entries, and a
        // merge can spill at most those plus the keys that already own a remainder: size R and M for that
        CU(cudaMemsetAsync(c->d_total, 0, 8, c->stream));
        CU(launch_db_count_big(c->stream, d + off, n, c->d_total, &c->launches));
        unsigned long long big = 0;
        CU(cudaMemcpyAsync(&big, c->d_total, 8, cudaMemcpyDeviceToHost, c->stream));
        uint64_t occ[3]; int rc = read_occ(c, occ); if (rc) return rc;
        c->big_seen += big;
        // Put: keys are new, only a field above a lane spills.  Merges: two records that each fit a lane can
        // sum past it (cov 80 + 80), so every record of the launch may create a remainder entry
        const uint64_t need_r = occ[1] + (mode == Feed::Put ? big : n), need_m = occ[2] + c->big_seen;
        if (need_r * 2 > c->rcap) { rc = grow_side(c, 1, std::max<uint64_t>(4 * c->rcap, 4 * need_r)); if (rc) return rc; }
        if (mode != Feed::Put && need_m * 2 > c->mcap) { rc = grow_side(c, 2, std::max<uint64_t>(4 * c->mcap, 4 * need_m)); if (rc) return rc; }
        if ((double)(occ[0] + n) > MAX_LOAD * (double)table_slots(c)) {
            rc = grow_table(c, std::max<uint64_t>(2 * table_slots(c), (uint64_t)((double)(occ[0] + n) / TARGET_LOAD)));
            if (rc) return rc;
        }
        const TableView tv = view(c);
        const int t7 = time_begin(c, 7);
        cudaError_t e = mode == Feed::Put ? launch_put(c->stream, tv, d + off, n, &c->launches)
                      : mode == Feed::MergeRef ? launch_merge_ref(c->stream, tv, d + off, n, &c->launches)
                                               : launch_add_records(c->stream, tv, d + off, n, &c->launches);
        time_end(c, t7);
        if (e != cudaSuccess) return fail(c, KQ_ECUDA, cudaGetErrorString(e));
        off += n;
    }
    uint64_t occ[3]; int rc = read_occ(c, occ); if (rc) return rc;
    c->occ_bound = occ[0];
    return check_device_error(c);
}

// stream host records through a device buffer into one of the table kernels, in sub-batches small
// enough that neither the table nor a side table can fill up inside a launch
int feed_records(kq_ctx *c, const std::vector<XRec> &xs, Feed mode) {
    if (xs.empty()) return KQ_OK;
    XRec *d = nullptr;
    const uint64_t max_batch = 8ull << 20;
    CU(cudaMalloc(&d, std::min<uint64_t>(xs.size(), max_batch) * sizeof(XRec)));
    for (uint64_t off = 0; off < xs.size();) {
        uint64_t n = std::min<uint64_t>(xs.size() - off, max_batch);
        uint64_t occ[3]; int rc = read_occ(c, occ); if (rc) { cudaFree(d); return rc; }
        if (occ[1] * 4 > c->rcap) { rc = grow_side(c, 1, c->rcap * 4); if (rc) { cudaFree(d); return rc; } }
        if (occ[2] * 4 > c->mcap) { rc = grow_side(c, 2, c->mcap * 4); if (rc) { cudaFree(d); return rc; } }
        n = std::min<uint64_t>(n, std::min(c->rcap, c->mcap) / 4);
        if ((double)(occ[0] + n) > MAX_LOAD * (double)table_slots(c)) {
            rc = grow_table(c, std::max<uint64_t>(2 * table_slots(c), (uint64_t)((double)(occ[0] + n) / TARGET_LOAD)));
            if (rc) { cudaFree(d); return rc; }
        }
        CU(cudaMemcpyAsync(d, xs.data() + off, n * sizeof(XRec), cudaMemcpyHostToDevice, c->stream));
        c->h2d += n * sizeof(XRec);
        const TableView tv = view(c);
        cudaError_t e = mode == Feed::Put ? launch_put(c->stream, tv, d, n, &c->launches)
                      : mode == Feed::MergeRef ? launch_merge_ref(c->stream, tv, d, n, &c->launches)
                                               : launch_add_records(c->stream, tv, d, n, &c->launches);
        if (e != cudaSuccess) { cudaFree(d); return fail(c, KQ_ECUDA, cudaGetErrorString(e)); }
        CU(cudaStreamSynchronize(c->stream));
        off += n;
    }
    cudaFree(d);
    uint64_t occ[3]; int rc = read_occ(c, occ); if (rc) return rc;
    c->occ_bound = occ[0];
    return check_device_error(c);
}

XRec raw_xrec(const Rec8Entry &e) {
    XRec x; x.key = e.key; x.flags = e.cov == 255 ? F_PLACEHOLDER : 0u;
    for (int w = 0; w < 4; ++w) { x.r.fw[w] = e.fw[w]; x.r.bw[w] = e.bw[w]; }
    x.r.cov = e.cov;
    return x;
}

}  // namespace

extern "C" int kq_db_save(kq_ctx *c, const char *dir) { return save_impl(c, dir, true); }
extern "C" int kq_db_save_hc_part(kq_ctx *c, const char *dir) { return save_impl(c, dir, false); }

extern "C" int kq_db_merge_hc_parts(const char *dir_c, int nranks) {
    if (!dir_c || nranks < 1) return KQ_EINVAL;
    const std::string dir(dir_c);
    std::vector<Rec32Entry> hc;
    for (int r = 0; r < nranks; ++r) {
        FILE *f = std::fopen(hc_part_path(dir, r).c_str(), "rb");
        if (!f) return KQ_EIO;
        uint64_t n = 0;
        if (std::fread(&n, 8, 1, f) != 1) { std::fclose(f); return KQ_EIO; }
        const size_t at = hc.size(); hc.resize(at + n);
        if (n && std::fread(hc.data() + at, sizeof(Rec32Entry), n, f) != n) { std::fclose(f); return KQ_EIO; }
        std::fclose(f);
    }
    if (!write_map32(dir + "/.map.hc.bin", hc).empty()) return KQ_EIO;
    for (int r = 0; r < nranks; ++r) std::remove(hc_part_path(dir, r).c_str());
    return KQ_OK;
}

extern "C" int kq_db_read_k(const char *dir) {
    if (!dir) return KQ_EINVAL;
    const int k = read_index_k(dir);
    return k < 0 ? KQ_EIO : k;
}

extern "C" int kq_db_load(kq_ctx *c, const char *dir_c) {
    if (!c || !dir_c) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    const std::string dir(dir_c);
    const int k = read_index_k(dir);
    if (k < 0) return fail(c, KQ_EIO, "Cannot load DBG input. Exiting.");          // src/input.cpp:70-72
    if (k != c->k) return fail(c, KQ_EINVAL, "database k-mer length " + std::to_string(k) + " differs from the handle's " + std::to_string(c->k));
    int rc = kq_clear(c); if (rc) return rc;
    std::vector<Rec32Entry> hc; rc = read_owned_hc(c, dir, hc); if (rc) return rc;
    rc = upload_map32(c, hc, MIN_SCAP); if (rc) return rc;
    if (!db_on_host()) {
        XRec *d_xs = nullptr; uint64_t n = 0;
        rc = load_maps_gpu(c, dir, &d_xs, &n); if (rc) return rc;
        if ((double)n > TARGET_LOAD * (double)table_slots(c)) rc = grow_table(c, (uint64_t)((double)n / TARGET_LOAD));
        if (!rc) rc = feed_device(c, d_xs, n, Feed::Put);
        cudaFree(d_xs);
        if (rc) return rc;
        c->loaded_db = true;
        return KQ_OK;
    }
    std::vector<Rec8Entry> all; rc = read_owned_maps(c, dir, all); if (rc) return rc;
    std::vector<XRec> xs(all.size());
    global_pool().parallel_for((all.size() + (1 << 20) - 1) >> 20, [&](size_t t) {
        const size_t b = t << 20, e = std::min(all.size(), b + (1 << 20));
        for (size_t i = b; i < e; ++i) xs[i] = raw_xrec(all[i]);
    });
    if ((double)xs.size() > TARGET_LOAD * (double)table_slots(c)) {
        rc = grow_table(c, (uint64_t)((double)xs.size() / TARGET_LOAD)); if (rc) return rc;
    }
    rc = feed_records(c, xs, Feed::Put); if (rc) return rc;
    c->loaded_db = true;
    return KQ_OK;
}

// ---- union -----------------------------------------------------------------------------------------------------------

extern "C" int kq_union(kq_ctx *c, const char *const *dirs, int n, int ref_compatible) {
    if (!c || !dirs || n < 1) return KQ_EINVAL;
    CU(cudaSetDevice(c->device));
    for (int i = 0; i < n; ++i) {                                      // src/input.cpp:126-145
        if (!dirs[i]) return KQ_EINVAL;
        const int k = read_index_k(dirs[i]);
        if (k < 0) return fail(c, KQ_EIO, std::string("cannot read ") + dirs[i] + "/.index");
        if (k != c->k) return fail(c, KQ_EINVAL, "Cannot merge databases with different kmer length.");
    }
    int rc = kq_clear(c); if (rc) return rc;
    auto sat = [](uint32_t a, uint32_t b) { return KQ_LARGEST - a >= b ? a + b : KQ_LARGEST; };

    if (ref_compatible) {
        // 1. saturating sum of all high-copy maps (src/graph-builder.cpp:301-336)
        std::vector<Rec32Entry> hc;
        for (int i = 0; i < n; ++i) { rc = read_owned_hc(c, dirs[i], hc); if (rc) return rc; }
        std::sort(hc.begin(), hc.end(), [](const Rec32Entry &a, const Rec32Entry &b) { return a.key < b.key; });
        std::vector<Rec32Entry> m32;
        for (const auto &h : hc) {
            if (!m32.empty() && m32.back().key == h.key) {
                Rec32Entry &d = m32.back();
                for (int w = 0; w < 4; ++w) { d.fw[w] = sat(d.fw[w], h.fw[w]); d.bw[w] = sat(d.bw[w], h.bw[w]); }
                d.cov = sat(d.cov, h.cov);
            } else m32.push_back(h);
        }
        rc = upload_map32(c, m32, 1ull << 16); if (rc) return rc;
        // 2. DB0's 8-bit maps raw (kmer.h:668-672), 3. later databases in order (kmer.h:674-686)
        for (int i = 0; i < n; ++i) {
            if (!db_on_host()) {
                XRec *d_xs = nullptr; uint64_t nx = 0;
                rc = load_maps_gpu(c, dirs[i], &d_xs, &nx); if (rc) return rc;
                rc = feed_device(c, d_xs, nx, i == 0 ? Feed::Put : Feed::MergeRef);
                cudaFree(d_xs);
                if (rc) return rc;
                continue;
            }
            std::vector<Rec8Entry> all; rc = read_owned_maps(c, dirs[i], all); if (rc) return rc;
            std::vector<XRec> xs(all.size());
            for (size_t j = 0; j < all.size(); ++j) xs[j] = raw_xrec(all[j]);
            std::vector<Rec8Entry>().swap(all);
            rc = feed_records(c, xs, i == 0 ? Feed::Put : Feed::MergeRef); if (rc) return rc;
        }
        c->loaded_db = true;
        return KQ_OK;
    }

    // exact mode: saturating sum of every database's lookup-visible records
    for (int i = 0; i < n; ++i) {
        std::vector<Rec32Entry> hc; rc = read_owned_hc(c, dirs[i], hc); if (rc) return rc;
        std::sort(hc.begin(), hc.end(), [](const Rec32Entry &a, const Rec32Entry &b) { return a.key < b.key; });
        std::vector<Rec8Entry> all; rc = read_owned_maps(c, dirs[i], all); if (rc) return rc;
        std::vector<XRec> xs(all.size());
        for (size_t j = 0; j < all.size(); ++j) {
            xs[j] = raw_xrec(all[j]); xs[j].flags = 0;
            if (all[j].cov != 255) continue;
            auto it = std::lower_bound(hc.begin(), hc.end(), all[j].key, [](const Rec32Entry &a, uint64_t key) { return a.key < key; });
            if (it == hc.end() || it->key != all[j].key) return fail(c, KQ_ECORRUPT, "Error: int32 map missing 255 value from int8 map");
            std::memcpy(xs[j].r.fw, it->fw, 16); std::memcpy(xs[j].r.bw, it->bw, 16); xs[j].r.cov = it->cov;
        }
        rc = feed_records(c, xs, Feed::AddLogical); if (rc) return rc;
    }
    c->loaded_db = true;
    return KQ_OK;
}

// ---- host-only codec entry points ---------------------------------------------------------------------------------------

extern "C" int kq_db_write_records(const char *dir_c, int k, const uint64_t *keys, const kq_rec32 *recs, size_t n) {
    if (!dir_c || k < 1 || k > 32 || (n && (!keys || !recs))) return KQ_EINVAL;
    const std::string dir(dir_c);
    mkdir(dir.c_str(), 0777);
    std::vector<std::vector<Rec8Entry>> maps(128); std::vector<Rec32Entry> hc;
    for (size_t i = 0; i < n; ++i) {
        Rec8Entry e; e.key = keys[i];
        if (recs[i].cov <= 254) {
            for (int w = 0; w < 4; ++w) { e.fw[w] = (uint8_t)recs[i].fw[w]; e.bw[w] = (uint8_t)recs[i].bw[w]; }
            e.cov = (uint8_t)recs[i].cov;
        } else {
            std::memset(e.fw, 0, 4); std::memset(e.bw, 0, 4); e.cov = 255;
            Rec32Entry h; h.key = keys[i]; std::memcpy(h.fw, recs[i].fw, 16); std::memcpy(h.bw, recs[i].bw, 16); h.cov = recs[i].cov;
            hc.push_back(h);
        }
        maps[keys[i] % 128].push_back(e);
    }
    for (int m = 0; m < 128; ++m) if (!write_map8(map_path(dir, m), maps[m]).empty()) return KQ_EIO;
    if (!write_map32(dir + "/.map.hc.bin", hc).empty()) return KQ_EIO;
    return write_index(dir, k).empty() ? KQ_OK : KQ_EIO;
}

extern "C" int kq_db_read_records(const char *dir_c, uint64_t *keys, kq_rec32 *recs, size_t *n) {
    if (!dir_c || !n) return KQ_EINVAL;
    const std::string dir(dir_c);
    std::vector<Rec32Entry> hc;
    if (!read_map32(dir + "/.map.hc.bin", hc).empty()) return KQ_EIO;
    std::sort(hc.begin(), hc.end(), [](const Rec32Entry &a, const Rec32Entry &b) { return a.key < b.key; });
    std::vector<Rec8Entry> all;
    for (int m = 0; m < 128; ++m) if (!read_map8(map_path(dir, m), all).empty()) return KQ_EIO;
    std::sort(all.begin(), all.end(), [](const Rec8Entry &a, const Rec8Entry &b) { return a.key < b.key; });
    const size_t cap = *n; size_t m = 0;
    for (const auto &e : all) {
        kq_rec32 r;
        if (e.cov == 255) {
            auto it = std::lower_bound(hc.begin(), hc.end(), e.key, [](const Rec32Entry &a, uint64_t key) { return a.key < key; });
            if (it == hc.end() || it->key != e.key) return KQ_ECORRUPT;
            std::memcpy(r.fw, it->fw, 16); std::memcpy(r.bw, it->bw, 16); r.cov = it->cov;
        } else {
            for (int w = 0; w < 4; ++w) { r.fw[w] = e.fw[w]; r.bw[w] = e.bw[w]; }
            r.cov = e.cov;
        }
        if (keys && recs && m < cap) { keys[m] = e.key; recs[m] = r; }
        ++m;
    }
    *n = m;
    return KQ_OK;
}

#include "kq_multi.inl"
#include "kq_static.inl"

// kq_multi.inl -- several GPUs of one box behind ONE handle (textually part of kq_api.cu).
//
// kq_multi owns one kq_ctx per GPU (shard g = maps [g*128/G, (g+1)*128/G)) and one host thread
// per GPU while a collective step runs.  Everything that crosses GPUs goes over peer memory in kernels:
//   count    read batches are dealt round-robin to the GPUs' pending streams; a round = on every GPU
//            kq_route_hist -> histograms copied to every peer -> kq_route_scatter (tuples written straight into the
//            owners' slabs) -> kq_route_insert; the steps of different GPUs are ordered with CUDA events, the host
//            threads meet at a barrier only to know that the events they wait for have been recorded;
//   validate chunks of a segment are dealt to the GPUs; kq_vroute_send -> kq_vroute_answer (queries to the owners,
//            finished rows back), then every GPU copies its rows to the caller's array;
//   save / load / union / stats are shard-local (a GPU owns whole .map.<m>.bin files), the host adds up counters.
// Replaces what the reference does with its 128 maps on the host cores (gfalibs/include/kmer.h:126-153 Kmap ctor,
// :362-381 buffersToMaps, :831-852 stats; src/kreeq.cpp:47-108 validateSequences).

#include <condition_variable>
#include <mutex>
#include <thread>

namespace {

// barrier for the per-GPU host threads of one collective step; abort() releases everybody with `false`
class HostBarrier {
public:
    void reset(int n) { n_ = n; waiting_ = 0; ++gen_; aborted_ = false; }
    bool wait() {
        std::unique_lock<std::mutex> l(mu_);
        if (aborted_) return false;
        const uint64_t g = gen_;
        if (++waiting_ == n_) { waiting_ = 0; ++gen_; cv_.notify_all(); return true; }
        cv_.wait(l, [&] { return gen_ != g || aborted_; });
        return !aborted_;
    }
    void abort() { std::lock_guard<std::mutex> l(mu_); aborted_ = true; cv_.notify_all(); }
private:
    std::mutex mu_; std::condition_variable cv_;
    int n_ = 0, waiting_ = 0; uint64_t gen_ = 0; bool aborted_ = false;
};

struct Taken { const uint64_t *bases, *inv; uint64_t len; uint64_t *inv_w; cudaEvent_t ev_read; };

}  // namespace

struct kq_multi {
    int k = 0, n = 0;
    std::vector<kq_ctx *> sh;
    std::vector<int> dev;
    std::string err;
    // count
    uint64_t round_pos = 256ull << 20;           // k-mer start positions a GPU routes per round
    int next = 0;                                // GPU that receives the next batch
    int bins = 0;
    std::vector<unsigned long long *> d_hist, d_allhist;
    std::vector<cudaEvent_t> ev;                 // one per GPU, re-recorded at every step of a collective
    std::thread round_thr; bool round_running = false; int round_rc = 0;
    HostBarrier bar;
    bool routed = false;
    // validate
    uint64_t vchunk = 8ull << 20; bool vready = false;
    kq_qv vacc{0, 0, 0};
};

namespace {

int mfail(kq_multi *m, int code, const std::string &msg) { if (m) m->err = msg; else g_create_error = msg; return code; }

// run fn(g) on one host thread per GPU; first failure wins (the barrier is aborted so that nobody waits for it)
template <class F> int run_all(kq_multi *m, F &&fn) {
    std::vector<int> rc((size_t)m->n, 0);
    m->bar.reset(m->n);
    std::vector<std::thread> th;
    for (int g = 0; g < m->n; ++g)
        th.emplace_back([&, g] {
            cudaSetDevice(m->dev[g]);
            rc[g] = fn(g);
            if (rc[g]) m->bar.abort();
        });
    for (auto &t : th) t.join();
    for (int g = 0; g < m->n; ++g)
        if (rc[g]) { m->err = "GPU " + std::to_string(g) + ": " + m->sh[g]->err; return rc[g]; }
    return KQ_OK;
}

// every GPU's stream waits for what every GPU has enqueued so far
int meet(kq_multi *m, int g) {
    kq_ctx *c = m->sh[g];
    CU(cudaEventRecord(m->ev[g], c->stream));
    if (!m->bar.wait()) return fail(c, KQ_ESTATE, "another GPU failed");
    for (int h = 0; h < m->n; ++h) if (h != g) CU(cudaStreamWaitEvent(c->stream, m->ev[h], 0));
    if (!m->bar.wait()) return fail(c, KQ_ESTATE, "another GPU failed");      // nobody re-records before everybody has waited
    return KQ_OK;
}

int multi_route_setup(kq_multi *m) {
    if (m->routed) return KQ_OK;
    const uint64_t slab = (uint64_t)((double)m->round_pos * 1.3) + 8192;
    int nb_bits = 0; { const int per = (m->n <= 2 ? 1024 : 2048) / m->n; while ((2 << nb_bits) <= per) ++nb_bits; }
    // (owner, hash range) bins of the routed scatter must fit its shared memory (16-byte tuples: half as many)
    while (nb_bits > 0 && scatter_pick_cap((unsigned long long)m->n << nb_bits, tuple_words(m->k), KQ_ROUTE_MAX_CAP) == 0) --nb_bits;
    std::vector<void *> slabs((size_t)m->n, nullptr);
    for (int g = 0; g < m->n; ++g) {
        int rc = kq_route_setup(m->sh[g], slab, nb_bits, &slabs[g], nullptr);
        if (rc) return mfail(m, rc, m->sh[g]->err);
    }
    for (int g = 0; g < m->n; ++g) {
        int rc = kq_route_connect(m->sh[g], slabs.data());
        if (rc) return mfail(m, rc, m->sh[g]->err);
    }
    m->bins = kq_route_bins(m->sh[0]);
    m->d_hist.assign((size_t)m->n, nullptr); m->d_allhist.assign((size_t)m->n, nullptr);
    for (int g = 0; g < m->n; ++g) {
        cudaSetDevice(m->dev[g]);
        if (cudaMalloc(&m->d_hist[g], (size_t)m->bins * 8) != cudaSuccess || cudaMalloc(&m->d_allhist[g], (size_t)m->n * m->bins * 8) != cudaSuccess)
            return mfail(m, KQ_ENOMEM, "cannot allocate the routing histograms");
    }
    m->routed = true;
    return KQ_OK;
}

// one GPU's part of a count step over the reads taken from every GPU's pending stream
int round_body(kq_multi *m, int g, const std::vector<Taken> &taken) {
    kq_ctx *c = m->sh[g];
    const Taken &t = taken[g];
    const uint64_t k = (uint64_t)m->k;
    uint64_t rounds = 0;
    for (const auto &x : taken) { const uint64_t np = x.len >= k ? x.len - k + 1 : 0; rounds = std::max(rounds, (np + m->round_pos - 1) / m->round_pos); }
    const uint64_t npos_all = t.len >= k ? t.len - k + 1 : 0;
    for (uint64_t r = 0; r < rounds; ++r) {
        const uint64_t p0 = std::min(r * m->round_pos, npos_all), np = std::min(m->round_pos, npos_all - p0);
        // an empty share still takes part: kq_route_hist clears the histogram, the scatter writes nothing
        const uint64_t len = t.len >= k ? t.len : k;                          // (no positions are read when np == 0)
        const uint64_t *dummy = reinterpret_cast<const uint64_t *>(m->d_hist[g]);   // never read when np == 0
        const uint64_t *bases = t.bases ? t.bases : dummy, *inv = t.inv ? t.inv : dummy;
        int rc = kq_route_hist(c, bases, inv, len, p0, np, m->d_hist[g]); if (rc) return rc;
        for (int h = 0; h < m->n; ++h)                                        // all-gather of the histograms: peer copies
            CU(cudaMemcpyPeerAsync(m->d_allhist[h] + (size_t)g * m->bins, m->dev[h], m->d_hist[g], m->dev[g], (size_t)m->bins * 8, c->stream));
        rc = meet(m, g); if (rc) return rc;                                   // histograms complete; previous inserts done
        uint64_t nrecv = 0;
        rc = kq_route_scatter(c, bases, inv, len, p0, np, m->d_allhist[g], &nrecv); if (rc) return rc;
        rc = meet(m, g); if (rc) return rc;                                   // every slab is complete
        rc = kq_route_insert(c, nrecv); if (rc) return rc;
    }
    int rc = meet(m, g); if (rc) return rc;                                   // slabs free again
    if (t.len) {
        CU(cudaMemsetAsync(t.inv_w, 0xFF, (t.len / 64 + 1) * 8, c->stream));  // the buffer goes back "all invalid"
        CU(cudaEventRecord(t.ev_read, c->stream));
    }
    return KQ_OK;
}

int join_round(kq_multi *m) {
    if (!m->round_running) return KQ_OK;
    m->round_thr.join();
    m->round_running = false;
    return m->round_rc;
}

// take what is pending on every GPU (on the caller's thread: the handles' host state is single-threaded) and
// count it on a background thread, so that the producer goes on packing and copying the next batches
int start_round(kq_multi *m) {
    int rc = join_round(m); if (rc) return rc;
    rc = multi_route_setup(m); if (rc) return rc;
    std::vector<Taken> taken((size_t)m->n);
    bool any = false;
    for (int g = 0; g < m->n; ++g) {
        kq_ctx *c = m->sh[g];
        Taken &t = taken[g];
        t = Taken{nullptr, nullptr, 0, nullptr, nullptr};
        if (!c->pend_len) continue;
        any = true;
        t = Taken{c->d_pb, c->d_pi, c->pend_len, c->d_pi, c->ev_read};
        cudaSetDevice(c->device);
        if (cudaStreamWaitEvent(c->stream, c->ev_copied, 0) != cudaSuccess) return mfail(m, KQ_ECUDA, "cudaStreamWaitEvent failed");   // every batch of the taken buffer has arrived
        c->read_armed = true; c->pend_len = 0;
        std::swap(c->d_pb, c->d_pb_alt); std::swap(c->d_pi, c->d_pi_alt); std::swap(c->pend_cap, c->pend_cap_alt);
        std::swap(c->ev_read, c->ev_read_alt); std::swap(c->read_armed, c->read_armed_alt);
    }
    if (!any) return KQ_OK;
    m->round_running = true;
    m->round_thr = std::thread([m, taken] { m->round_rc = run_all(m, [&](int g) { return round_body(m, g, taken); }); });
    return KQ_OK;
}

int multi_vroute_setup(kq_multi *m) {
    if (m->vready) return KQ_OK;
    std::vector<void *> inbox((size_t)m->n, nullptr), rows((size_t)m->n, nullptr);
    for (int g = 0; g < m->n; ++g) {
        int rc = kq_vroute_setup(m->sh[g], m->vchunk, &inbox[g], nullptr, &rows[g], nullptr);
        if (rc) return mfail(m, rc, m->sh[g]->err);
    }
    for (int g = 0; g < m->n; ++g) {
        int rc = kq_vroute_connect(m->sh[g], inbox.data(), rows.data());
        if (rc) return mfail(m, rc, m->sh[g]->err);
    }
    m->vready = true;
    return KQ_OK;
}

}  // namespace

// ---- lifetime ------------------------------------------------------------------------------------------------------

extern "C" int kq_multi_create(kq_multi **out, int k, int ngpus, const int *devices, uint64_t expected_distinct) {
    if (!out) return mfail(nullptr, KQ_EINVAL, "kq_multi_create: out is NULL");
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return mfail(nullptr, KQ_ENODEV, "no CUDA device");
    if (ngpus == 0) { ngpus = 1; while (ngpus * 2 <= ndev && ngpus * 2 <= KQ_MAX_RANKS) ngpus *= 2; }   // all visible, power of two
    if (ngpus < 1 || ngpus > KQ_MAX_RANKS || (ngpus & (ngpus - 1))) return mfail(nullptr, KQ_EINVAL, "the number of GPUs must be a power of two (1, 2, 4, 8, 16)");
    if (!devices && ngpus > ndev) return mfail(nullptr, KQ_ENODEV, "--gpus " + std::to_string(ngpus) + ": only " + std::to_string(ndev) + " CUDA device(s) visible");
    kq_multi *m = new kq_multi;
    m->k = k; m->n = ngpus;
    for (int g = 0; g < ngpus; ++g) {
        const int d = devices ? devices[g] : g;
        kq_ctx *c = nullptr;
        int rc = kq_create(&c, k, d, expected_distinct ? expected_distinct / ngpus + 1 : 0);
        if (rc) { kq_multi_destroy(m); return rc; }                  // g_create_error holds the message
        m->sh.push_back(c); m->dev.push_back(d);
        kq_set_shard(c, g, ngpus);
        c->hold_pending = ngpus > 1;
    }
    // peer access between every pair of distinct devices (kernels store into / read from the peers' buffers)
    for (int g = 0; g < ngpus; ++g)
        for (int h = 0; h < ngpus; ++h) {
            if (m->dev[g] == m->dev[h]) continue;
            cudaSetDevice(m->dev[g]);
            int can = 0; cudaDeviceCanAccessPeer(&can, m->dev[g], m->dev[h]);
            if (!can) { kq_multi_destroy(m); return mfail(nullptr, KQ_ENODEV, "GPU " + std::to_string(m->dev[g]) + " cannot access GPU " + std::to_string(m->dev[h]) + " (no peer access)"); }
            const cudaError_t e = cudaDeviceEnablePeerAccess(m->dev[h], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { kq_multi_destroy(m); return mfail(nullptr, KQ_ENODEV, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e)); }
            cudaGetLastError();
        }
    m->ev.assign((size_t)ngpus, nullptr);
    for (int g = 0; g < ngpus; ++g) { cudaSetDevice(m->dev[g]); cudaEventCreateWithFlags(&m->ev[g], cudaEventDisableTiming); }
    *out = m;
    return KQ_OK;
}

extern "C" void kq_multi_destroy(kq_multi *m) {
    if (!m) return;
    if (m->round_running) { m->round_thr.join(); m->round_running = false; }
    for (size_t g = 0; g < m->ev.size(); ++g) if (m->ev[g]) { cudaSetDevice(m->dev[g]); cudaEventDestroy(m->ev[g]); }
    for (size_t g = 0; g < m->d_hist.size(); ++g) { cudaSetDevice(m->dev[g]); cudaFree(m->d_hist[g]); cudaFree(m->d_allhist[g]); }
    for (auto *c : m->sh) kq_destroy(c);
    delete m;
}

extern "C" const char *kq_multi_last_error(kq_multi *m) { return m ? m->err.c_str() : g_create_error.c_str(); }
extern "C" int kq_multi_ngpus(kq_multi *m) { return m ? m->n : KQ_EINVAL; }
extern "C" kq_ctx *kq_multi_shard(kq_multi *m, int g) { return m && g >= 0 && g < m->n ? m->sh[g] : nullptr; }

extern "C" int kq_multi_set_round(kq_multi *m, uint64_t round_positions, uint64_t validate_chunk) {
    if (!m || m->routed || m->vready) return KQ_EINVAL;               // before the first push / validate
    if (round_positions) m->round_pos = round_positions;
    if (validate_chunk) m->vchunk = std::min<uint64_t>(validate_chunk, 0xFFFFFFFFull);
    return KQ_OK;
}

// ---- count ---------------------------------------------------------------------------------------------------------

extern "C" int kq_multi_count_push(kq_multi *m, const char *ascii, size_t len) {
    if (!m || (!ascii && len)) return KQ_EINVAL;
    if (m->n == 1) { int rc = kq_count_push(m->sh[0], ascii, len); return rc ? mfail(m, rc, m->sh[0]->err) : rc; }
    kq_ctx *c = m->sh[m->next];
    m->next = (m->next + 1) % m->n;
    int rc = kq_count_push(c, ascii, len);                            // pack + copy into that GPU's pending stream
    if (rc) return mfail(m, rc, c->err);
    uint64_t most = 0;
    for (auto *s : m->sh) most = std::max(most, s->pend_len);
    if (most >= m->round_pos) { rc = start_round(m); if (rc) return rc; }
    return KQ_OK;
}

extern "C" int kq_multi_count_finish(kq_multi *m) {
    if (!m) return KQ_EINVAL;
    if (m->n > 1) {
        int rc = start_round(m); if (rc) return rc;
        rc = join_round(m); if (rc) return rc;
    }
    return run_all(m, [&](int g) { return kq_sync(m->sh[g]); });
}

extern "C" int kq_multi_count_instances(kq_multi *m, uint64_t *n) {
    if (!m || !n) return KQ_EINVAL;
    int rc = join_round(m); if (rc) return rc;
    *n = 0;
    for (auto *c : m->sh) { uint64_t v = 0; rc = kq_count_instances(c, &v); if (rc) return mfail(m, rc, c->err); *n += v; }
    return KQ_OK;
}

extern "C" int kq_multi_clear(kq_multi *m) {
    if (!m) return KQ_EINVAL;
    int rc = join_round(m); if (rc) return rc;
    m->vacc = kq_qv{0, 0, 0};
    return run_all(m, [&](int g) { return kq_clear(m->sh[g]); });
}

// ---- stats / export / lookups ------------------------------------------------------------------------------------------

extern "C" int kq_multi_stats_get(kq_multi *m, kq_stats *out) {
    if (!m || !out) return KQ_EINVAL;
    int rc = kq_multi_count_finish(m); if (rc) return rc;
    std::vector<kq_stats> s((size_t)m->n);
    rc = run_all(m, [&](int g) { return kq_stats_get(m->sh[g], &s[g]); }); if (rc) return rc;
    *out = kq_stats{0, 0, 0, 0, 0};
    for (const auto &x : s) { out->total += x.total; out->unique += x.unique; out->distinct += x.distinct; out->edges += x.edges; }
    out->missing = (uint64_t)(pow(4, m->k) - out->distinct);          // evaluated in double, src/graph-builder.cpp:286
    return KQ_OK;
}

extern "C" int kq_multi_hist(kq_multi *m, uint64_t *cov, uint64_t *n, size_t *len) {
    if (!m || !len) return KQ_EINVAL;
    int rc = kq_multi_count_finish(m); if (rc) return rc;
    std::vector<std::pair<uint64_t, uint64_t>> all;
    for (auto *c : m->sh) {
        cudaSetDevice(c->device);
        FullStats fs; rc = compute_stats(c, fs); if (rc) return mfail(m, rc, c->err);
        all.insert(all.end(), fs.hist.begin(), fs.hist.end());
    }
    std::sort(all.begin(), all.end());
    std::vector<std::pair<uint64_t, uint64_t>> h;
    for (const auto &e : all) { if (!h.empty() && h.back().first == e.first) h.back().second += e.second; else h.push_back(e); }
    const size_t cap = *len; *len = h.size();
    if (cov && n) for (size_t i = 0; i < h.size() && i < cap; ++i) { cov[i] = h[i].first; n[i] = h[i].second; }
    return KQ_OK;
}

extern "C" int kq_multi_export(kq_multi *m, uint64_t *keys, kq_rec32 *recs, size_t *n) {
    if (!m || !n) return KQ_EINVAL;
    int rc = kq_multi_count_finish(m); if (rc) return rc;
    std::vector<std::vector<uint64_t>> ks((size_t)m->n); std::vector<std::vector<kq_rec32>> rs((size_t)m->n);
    rc = run_all(m, [&](int g) {
        size_t cnt = 0; int r = kq_export(m->sh[g], nullptr, nullptr, &cnt); if (r) return r;
        ks[g].resize(cnt); rs[g].resize(cnt);
        return kq_export(m->sh[g], ks[g].data(), rs[g].data(), &cnt);
    });
    if (rc) return rc;
    size_t total = 0; for (auto &v : ks) total += v.size();
    const size_t cap = *n; *n = total;
    if (!keys || !recs || cap < total) return KQ_OK;
    // every shard is sorted; merge by key
    std::vector<std::pair<uint64_t, std::pair<int, size_t>>> order; order.reserve(total);
    for (int g = 0; g < m->n; ++g) for (size_t i = 0; i < ks[g].size(); ++i) order.push_back({ks[g][i], {g, i}});
    std::sort(order.begin(), order.end());
    for (size_t i = 0; i < total; ++i) { keys[i] = order[i].first; recs[i] = rs[order[i].second.first][order[i].second.second]; }
    return KQ_OK;
}

extern "C" int kq_multi_lookup_batch(kq_multi *m, const uint64_t *keys, size_t n, kq_rec32 *rec, uint8_t *found) {
    if (!m || (n && (!keys || !rec || !found))) return KQ_EINVAL;
    int rc = kq_multi_count_finish(m); if (rc) return rc;
    std::vector<std::vector<uint64_t>> ks((size_t)m->n); std::vector<std::vector<size_t>> at((size_t)m->n);
    for (size_t i = 0; i < n; ++i) { const int o = kq_owner_of(keys[i], m->n); ks[o].push_back(keys[i]); at[o].push_back(i); }
    return run_all(m, [&](int g) {
        if (ks[g].empty()) return KQ_OK;
        std::vector<kq_rec32> r(ks[g].size()); std::vector<uint8_t> f(ks[g].size());
        int e = kq_lookup_batch(m->sh[g], ks[g].data(), ks[g].size(), r.data(), f.data()); if (e) return e;
        for (size_t i = 0; i < ks[g].size(); ++i) { found[at[g][i]] = f[i]; if (f[i]) rec[at[g][i]] = r[i]; }
        return KQ_OK;
    });
}

// ---- persistence / union (shard-local) -------------------------------------------------------------------------------

extern "C" int kq_multi_db_save(kq_multi *m, const char *dir) {
    if (!m || !dir) return KQ_EINVAL;
    int rc = kq_multi_count_finish(m); if (rc) return rc;
    mkdir(dir, 0777);
    rc = run_all(m, [&](int g) { return kq_db_save(m->sh[g], dir); }); if (rc) return rc;   // maps of the shard (+ .index on GPU 0) + .hc.part.<g>
    if (m->n > 1) { rc = kq_db_merge_hc_parts(dir, m->n); if (rc) return mfail(m, rc, std::string("cannot assemble ") + dir + "/.map.hc.bin"); }
    return KQ_OK;
}

extern "C" int kq_multi_db_load(kq_multi *m, const char *dir) {
    if (!m || !dir) return KQ_EINVAL;
    int rc = join_round(m); if (rc) return rc;
    return run_all(m, [&](int g) { return kq_db_load(m->sh[g], dir); });
}

extern "C" int kq_multi_union(kq_multi *m, const char *const *dirs, int n, int ref_compatible) {
    if (!m || !dirs || n < 1) return KQ_EINVAL;
    int rc = join_round(m); if (rc) return rc;
    return run_all(m, [&](int g) { return kq_union(m->sh[g], dirs, n, ref_compatible); });
}

// ---- validate ----------------------------------------------------------------------------------------------------------

extern "C" int kq_multi_validate_segment(kq_multi *m, const char *seq, uint64_t len, uint32_t cutoff, kq_base *out, kq_qv *acc) {
    if (!m || (!seq && len) || (!out && len) || !acc) return KQ_EINVAL;
    if (len == 0) return KQ_OK;
    if (m->n == 1) { int rc = kq_validate_segment(m->sh[0], seq, len, cutoff, out, acc); return rc ? mfail(m, rc, m->sh[0]->err) : rc; }
    int rc = kq_multi_count_finish(m); if (rc) return rc;
    rc = multi_vroute_setup(m); if (rc) return rc;
    const uint64_t k = (uint64_t)m->k, kcount = len >= k ? len - k + 1 : 0, V = m->vchunk;
    const uint64_t nchunks = (len + V - 1) / V, rounds = (nchunks + m->n - 1) / m->n;
    rc = run_all(m, [&](int g) -> int {
        kq_ctx *c = m->sh[g];
        for (uint64_t r = 0; r < rounds; ++r) {
            const uint64_t ci = r * m->n + g;
            const uint64_t c0 = std::min(len, ci * V), n = std::min(V, len - c0);
            Stage &s = c->st[c->next_stage]; c->next_stage = (c->next_stage + 1) % NBUF;
            uint64_t a0 = 0;
            if (n) {
                int e = stage_acquire(c, s); if (e) return e;
                const size_t need = packed_words(std::max(CHUNK, V + k + 128));
                e = stage_reserve(c, s, need, need, 0); if (e) return e;
                // slice of the segment this chunk needs: one base before c0 (64-aligned down) .. k bases after the last start
                a0 = c0 == 0 ? 0 : ((c0 - 1) & ~63ull);
                const uint64_t a1 = std::min<uint64_t>(len, c0 + n + k);
                const size_t words = packed_words(a1 - a0);
                pack_sequence(seq + a0, a1 - a0, s.h_bases, s.h_inv, &global_pool());
                CU(cudaMemcpyAsync(s.d_bases, s.h_bases, words * 8, cudaMemcpyHostToDevice, c->stream));
                CU(cudaMemcpyAsync(s.d_inv, s.h_inv, (words / 2 + 1) * 8, cudaMemcpyHostToDevice, c->stream));
                c->h2d += words * 8 + (words / 2 + 1) * 8;
            }
            int e = kq_vroute_send(c, s.d_bases, s.d_inv, a0, c0, n, kcount); if (e) return e;
            e = meet(m, g); if (e) return e;                            // every inbox is complete
            e = kq_vroute_answer(c, cutoff); if (e) return e;
            e = meet(m, g); if (e) return e;                            // every answer box is complete; the inboxes are free
            e = kq_vroute_place(c); if (e) return e;                    // answers -> rows (local)
            if (n) {
                CU(cudaMemcpyAsync(out + c0, c->d_vrows, n * sizeof(Base), cudaMemcpyDeviceToHost, c->stream));
                c->d2h += n * sizeof(Base);
                CU(cudaEventRecord(s.done, c->stream)); s.busy = true;
            }
            CU(cudaStreamSynchronize(c->stream));                       // rows are in `out`; the slab may be rewritten
        }
        return KQ_OK;
    });
    if (rc) return rc;
    for (auto *c : m->sh) { rc = kq_vroute_counters(c, acc, 1); if (rc) return mfail(m, rc, c->err); }
    return KQ_OK;
}

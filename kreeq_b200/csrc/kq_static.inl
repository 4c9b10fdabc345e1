// kq_static.inl -- a lookup entry point that allocates nothing (textually part of kq_api.cu).
//
// kq_lookup_batch stages keys and answers in device buffers it allocates and frees on every call.  Callers that
// must not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it) use
// kq_lookup_batch_static instead: same probe, same answers, the staging lives in ONE file-scope device array of
// KQ_STATIC_LOOKUP_MAX keys (2.9 MiB) that every GPU which loads the library holds until exit.
// No counterpart in the reference.
//
// All calls share the array, which keeps the previous call's contents: every byte a call reads it has written
// itself first (keys, the caller's records, cleared found flags).  Calls that overlap on one GPU race.

namespace {

constexpr size_t STATIC_MAX = KQ_STATIC_LOOKUP_MAX;
// [keys: 8 B each][records: sizeof(Rec32) each][found: 1 B each], every part 256-byte aligned
constexpr size_t STATIC_KEYS_OFF = 0;
constexpr size_t STATIC_RECS_OFF = STATIC_MAX * 8;
constexpr size_t STATIC_FOUND_OFF = STATIC_RECS_OFF + ((STATIC_MAX * sizeof(Rec32) + 255) & ~(size_t)255);
constexpr size_t STATIC_BYTES = STATIC_FOUND_OFF + ((STATIC_MAX + 255) & ~(size_t)255);

static __device__ __align__(256) unsigned char g_static_scratch[STATIC_BYTES];

// same probe as lookup_kernel (kq_kernels.cu); found[i] = 0 leaves rec[i] as the caller passed it
__global__ void __launch_bounds__(256)
lookup_static_kernel(TableView tv, const uint64_t *__restrict__ keys, uint64_t n, Rec32 *rec, uint8_t *found) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        Rec32 r; const int f = lookup(tv, keys[i], r);
        if (f < 0) atomicOr(tv.err, 4);
        found[i] = f > 0;
        if (f > 0) rec[i] = r;
    }
}

// device address of the array on the handle's GPU: the array's name is a host-side stand-in, not a device address
// (looked up once per device, with that device current)
int static_scratch_ptr(kq_ctx *c, unsigned char **out) {
    static std::mutex mu;
    static unsigned char *ptr[64] = {};
    if (c->device < 0 || c->device >= 64) return fail(c, KQ_EINVAL, "kq_lookup_batch_static: device ordinal out of range");
    std::lock_guard<std::mutex> l(mu);
    if (!ptr[c->device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) return fail(c, KQ_ECUDA, std::string("cudaGetSymbolAddress(static scratch): ") + cudaGetErrorString(e));
        ptr[c->device] = static_cast<unsigned char *>(p);
    }
    *out = ptr[c->device];
    return KQ_OK;
}

}  // namespace

extern "C" size_t kq_lookup_static_max(void) { return STATIC_MAX; }

extern "C" int kq_lookup_batch_static(kq_ctx *c, const uint64_t *keys, size_t n, kq_rec32 *rec, uint8_t *found) {
    if (!c || (n && (!keys || !rec || !found))) return KQ_EINVAL;
    if (n > STATIC_MAX)
        return fail(c, KQ_EINVAL, "kq_lookup_batch_static: " + std::to_string(n) + " keys exceed the static scratch of " +
                                  std::to_string(STATIC_MAX) + " keys (KQ_STATIC_LOOKUP_MAX); split the batch or use kq_lookup_batch");
    if (!n) return KQ_OK;
    CU(cudaSetDevice(c->device));
    // counting pending reads may grow the table (cudaMalloc): that is the caller's kq_sync, not this call's
    if (c->pend_len || c->ovf_pending)
        return fail(c, KQ_ESTATE, "kq_lookup_batch_static: reads are pending; call kq_sync (or kq_count_finish) first");
    unsigned char *base = nullptr;
    { int rc = static_scratch_ptr(c, &base); if (rc) return rc; }
    uint64_t *dk = reinterpret_cast<uint64_t *>(base + STATIC_KEYS_OFF);
    Rec32 *dr = reinterpret_cast<Rec32 *>(base + STATIC_RECS_OFF);
    uint8_t *df = base + STATIC_FOUND_OFF;
    static_assert(sizeof(Rec32) == sizeof(kq_rec32), "record layouts differ");
    CU(cudaMemcpyAsync(dk, keys, n * 8, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(dr, rec, n * sizeof(Rec32), cudaMemcpyHostToDevice, c->stream));   // untouched entries go back as they came
    CU(cudaMemsetAsync(df, 0, n, c->stream));
    const unsigned grid = (unsigned)std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count() * 8);
    lookup_static_kernel<<<grid, 256, 0, c->stream>>>(view(c), dk, n, dr, df);
    CU(cudaGetLastError());
    ++c->launches;
    CU(cudaMemcpyAsync(rec, dr, n * sizeof(Rec32), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(found, df, n, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->h2d += n * (8 + sizeof(Rec32)); c->d2h += n * (sizeof(Rec32) + 1);
    return check_device_error(c);
}

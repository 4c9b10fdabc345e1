// tools/gups.cu -- random-access microbenchmarks that size the design of the k-mer table
// (DESIGN.md "What bounds the table kernels").  Not part of the product library.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/gups tools/gups.cu
//   tools/gups [max_table_GiB]
//
// Every kernel issues N random operations on a table of S bytes (16-byte slots) and reports
// G ops/s: how many table touches per second the memory system sustains for each access recipe.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33; return x;
}

struct __align__(16) Slot16 { unsigned long long tag, cnt; };

__device__ __forceinline__ void cas128(Slot16 *p, unsigned long long clo, unsigned long long chi,
                                       unsigned long long slo, unsigned long long shi,
                                       unsigned long long &olo, unsigned long long &ohi) {
    asm volatile("{\n .reg .b128 c, s, d;\n mov.b128 c, {%2, %3};\n mov.b128 s, {%4, %5};\n"
                 " atom.relaxed.gpu.global.cas.b128 d, [%6], c, s;\n mov.b128 {%0, %1}, d;\n}"
                 : "=l"(olo), "=l"(ohi) : "l"(clo), "l"(chi), "l"(slo), "l"(shi), "l"(p) : "memory");
}

__device__ __forceinline__ uint4 ld16(const void *p) {
    uint4 v; asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

// mode: 0 ld16 | 1 ld32 (two ld16 of one 32-B sector) | 2 RED.64 | 3 ATOM.64 (return used) |
//       4 ld16 + CAS.128 | 5 ld8 + ATOM.32 ret + 2 x RED.64 (32-B slot: current count kernel) |
//       6 ld16 + CAS.64 on the counter word | 7 CAS.128 only (blind)
template <int MODE, int ILP>
__global__ void __launch_bounds__(256) gups(Slot16 *T, uint64_t nslots, uint64_t nops, uint64_t seed, unsigned long long *sink) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x * ILP;
    unsigned long long acc = 0;
    for (uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * ILP; i < nops; i += stride) {
        uint64_t idx[ILP];
#pragma unroll
        for (int j = 0; j < ILP; ++j) {
            idx[j] = __umul64hi(mix64(i + j + seed), nslots);
            if (MODE == 1 || MODE == 5) idx[j] &= ~1ull;                 // 32-byte aligned pair of slots
        }
        if (MODE == 0) {
#pragma unroll
            for (int j = 0; j < ILP; ++j) acc += ld16(&T[idx[j]]).x;
        } else if (MODE == 1) {
#pragma unroll
            for (int j = 0; j < ILP; ++j) acc += ld16(&T[idx[j]]).x + ld16(&T[idx[j] + 1]).y;
        } else if (MODE == 2) {
#pragma unroll
            for (int j = 0; j < ILP; ++j) atomicAdd(&T[idx[j]].cnt, 1ULL);
        } else if (MODE == 3) {
#pragma unroll
            for (int j = 0; j < ILP; ++j) acc += atomicAdd(&T[idx[j]].cnt, 1ULL);
        } else if (MODE == 4) {
            uint4 v[ILP];
#pragma unroll
            for (int j = 0; j < ILP; ++j) v[j] = ld16(&T[idx[j]]);
#pragma unroll
            for (int j = 0; j < ILP; ++j) {
                unsigned long long lo = ((unsigned long long)v[j].y << 32) | v[j].x, hi = ((unsigned long long)v[j].w << 32) | v[j].z, olo, ohi;
                for (int t = 0; t < 8; ++t) {
                    cas128(&T[idx[j]], lo, hi, idx[j] + 1, hi + 1, olo, ohi);
                    if (olo == lo && ohi == hi) break;
                    lo = olo; hi = ohi;
                }
                acc += olo;
            }
        } else if (MODE == 5) {
            unsigned long long t[ILP];
#pragma unroll
            for (int j = 0; j < ILP; ++j) t[j] = ld16(&T[idx[j]]).x;
#pragma unroll
            for (int j = 0; j < ILP; ++j) {
                acc += t[j] + atomicAdd(reinterpret_cast<unsigned int *>(&T[idx[j]].cnt), 1u);
                atomicAdd(&T[idx[j] + 1].tag, 1ULL);
                atomicAdd(&T[idx[j] + 1].cnt, 1ULL << 16);
            }
        } else if (MODE == 6) {
            uint4 v[ILP];
#pragma unroll
            for (int j = 0; j < ILP; ++j) v[j] = ld16(&T[idx[j]]);
#pragma unroll
            for (int j = 0; j < ILP; ++j) {
                unsigned long long hi = ((unsigned long long)v[j].w << 32) | v[j].z;
                for (int t = 0; t < 8; ++t) {
                    const unsigned long long o = atomicCAS(&T[idx[j]].cnt, hi, hi + 1);
                    if (o == hi) break;
                    hi = o;
                }
                acc += hi + v[j].x;
            }
        } else if (MODE == 7) {
#pragma unroll
            for (int j = 0; j < ILP; ++j) {
                unsigned long long olo, ohi;
                cas128(&T[idx[j]], 0, 0, idx[j] + 1, 1, olo, ohi);
                acc += olo;
            }
        }
    }
    if (acc == 0x123456789abcdefULL) *sink = acc;
}

// the partitioned insert's pattern: op i touches a random slot of slice i / ops_per_slice
// recipe: 0 ld16 + cas64 | 1 ld16 + atom64 | 2 current 32-B slot recipe
template <int RECIPE>
__global__ void __launch_bounds__(256) window(Slot16 *T, uint64_t slice_slots, uint64_t nslices, uint64_t nops, unsigned long long *sink) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t per = (nops + nslices - 1) / nslices;
    unsigned long long acc = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nops; i += stride) {
        uint64_t idx = (i / per) * slice_slots + (mix64(i) & (slice_slots - 1));
        if (RECIPE == 2) idx &= ~1ull;
        const uint4 v = ld16(&T[idx]);
        if (RECIPE == 0) {
            unsigned long long hi = ((unsigned long long)v.w << 32) | v.z;
            for (int t = 0; t < 8; ++t) { const unsigned long long o = atomicCAS(&T[idx].cnt, hi, hi + 1); if (o == hi) break; hi = o; }
            acc += hi + v.x;
        } else if (RECIPE == 1) {
            acc += v.x + atomicAdd(&T[idx].cnt, 1ULL);
        } else {
            acc += v.x + atomicAdd(reinterpret_cast<unsigned int *>(&T[idx].cnt), 1u);
            atomicAdd(&T[idx + 1].tag, 1ULL); atomicAdd(&T[idx + 1].cnt, 1ULL << 16);
        }
    }
    if (acc == 0x123456789abcdefULL) *sink = acc;
}

// faithful insert: keys drawn from 0.47 * slice_slots distinct keys per slice, linear probing, claim of an
// empty slot by CLAIM = 0: one 128-bit CAS, 1: 64-bit CAS on the tag then 64-bit CAS on the counters;
// STREAM: also stream an 8-byte tuple per op (ld.cs) like the real kernel
template <int CLAIM, bool STREAM>
__global__ void __launch_bounds__(256) window_insert(Slot16 *T, uint64_t slice_slots, uint64_t nslices, uint64_t nops,
                                                     const uint64_t *tuples, unsigned long long *sink) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t per = (nops + nslices - 1) / nslices, mask = slice_slots - 1, distinct = slice_slots * 47 / 100;
    unsigned long long acc = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nops; i += stride) {
        const uint64_t sl = i / per;
        uint64_t r = mix64(i);
        if (STREAM) r ^= __ldcs(tuples + i);
        const uint64_t keyid = __umul64hi(r, distinct), tag = (sl << 32 | keyid) + 1;
        Slot16 *base = T + sl * slice_slots;
        uint64_t j = mix64(tag) & mask;
        for (int n = 0; n < 4096; ++n) {
            Slot16 *p = base + j;
            uint4 v = ld16(p);
            unsigned long long t = ((unsigned long long)v.y << 32) | v.x, c = ((unsigned long long)v.w << 32) | v.z;
            if (t == 0) {
                if (CLAIM == 0) { unsigned long long olo, ohi; cas128(p, 0, 0, tag, 1, olo, ohi); if (olo == 0) break; t = olo; c = ohi; }
                else { const unsigned long long o = atomicCAS(&p->tag, 0ULL, tag); t = o == 0 ? tag : o; c = 0; }
            }
            if (t == tag) {
                for (int q = 0; q < 64; ++q) { const unsigned long long o = atomicCAS(&p->cnt, c, c + 1); if (o == c) break; c = o; }
                acc += c; break;
            }
            j = (j + 1) & mask;
        }
    }
    if (acc == 0x123456789abcdefULL) *sink = acc;
}

// each warp stores CH bytes (coalesced) at a random CH-aligned offset: the scatter kernel's flushes
template <int CH>
__global__ void __launch_bounds__(256) chunk_write(uint64_t *T, uint64_t nchunks_table, uint64_t nchunks) {
    const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    for (uint64_t c = warp; c < nchunks; c += nwarps) {
        const uint64_t at = __umul64hi(mix64(c + 7), nchunks_table) * (CH / 8);
        for (unsigned j = lane; j < CH / 8; j += 32) T[at + j] = c;
    }
}

template <int MODE, int ILP>
static void run(const char *name, Slot16 *T, uint64_t bytes, uint64_t nops, int ctas_per_sm, unsigned long long *sink) {
    const uint64_t nslots = bytes / sizeof(Slot16);
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    const int grid = 148 * ctas_per_sm;
    gups<MODE, ILP><<<grid, 256>>>(T, nslots, nops / 8, 1, sink);          // warm-up
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        CK(cudaEventRecord(a));
        gups<MODE, ILP><<<grid, 256>>>(T, nslots, nops, 1000 + r, sink);
        CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    printf("%-34s table %8.3f GiB  ilp %d  ctas/sm %d : %8.2f G ops/s  (%.2f ms)\n", name, bytes / 1073741824.0, ILP, ctas_per_sm,
           nops / (best * 1e-3) / 1e9, best);
    fflush(stdout);
}

int main(int argc, char **argv) {
    const double max_gib = argc > 1 ? atof(argv[1]) : 32.0;
    const int gran = argc > 2 ? atoi(argv[2]) : 0;
    if (gran) { CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, gran)); }
    size_t g = 0; CK(cudaDeviceGetLimit(&g, cudaLimitMaxL2FetchGranularity));
    printf("L2 fetch granularity limit: %zu\n", g);
    const uint64_t max_bytes = (uint64_t)(max_gib * 1073741824.0);
    Slot16 *T; unsigned long long *sink;
    CK(cudaMalloc(&T, max_bytes)); CK(cudaMalloc(&sink, 8));
    CK(cudaMemset(T, 0, max_bytes));
    const uint64_t nops = 1ull << 28;
    {
        cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); float ms;
        for (double mib : {32.0, 64.0, 96.0, 128.0, 192.0, 256.0, 384.0, 512.0}) {
            const uint64_t bytes = (uint64_t)(mib * 1048576.0);
            run<0, 1>("ld16 (footprint sweep)", T, bytes, nops, 8, sink);
            run<2, 1>("red64 (footprint sweep)", T, bytes, nops, 8, sink);
        }
        for (int log2s : {16, 18, 20, 21, 22, 23}) {
            // 2 GiB table swept once with 2.25 ops per slot (the count pipeline's insert phase on C2)
            const uint64_t slice_slots = 1ull << log2s, nsl = (2ull << 30) / 16 / slice_slots, n = 300ull << 20;
            CK(cudaEventRecord(a)); window<0><<<148 * 8, 256>>>(T, slice_slots, nsl, n, sink); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
            CK(cudaEventElapsedTime(&ms, a, b)); printf("window ld16+cas64  slice %4llu MiB x %llu : %8.2f G ops/s\n", (unsigned long long)(slice_slots >> 16), (unsigned long long)nsl, n / (ms * 1e-3) / 1e9);
            CK(cudaEventRecord(a)); window<1><<<148 * 8, 256>>>(T, slice_slots, nsl, n, sink); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
            CK(cudaEventElapsedTime(&ms, a, b)); printf("window ld16+atom64 slice %4llu MiB x %llu : %8.2f G ops/s\n", (unsigned long long)(slice_slots >> 16), (unsigned long long)nsl, n / (ms * 1e-3) / 1e9);
            CK(cudaEventRecord(a)); window<2><<<148 * 6, 256>>>(T, slice_slots, nsl, n, sink); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
            CK(cudaEventElapsedTime(&ms, a, b)); printf("window 32B recipe  slice %4llu MiB x %llu : %8.2f G ops/s\n", (unsigned long long)(slice_slots >> 16), (unsigned long long)nsl, n / (ms * 1e-3) / 1e9);
        }
        {
            uint64_t *tup; const uint64_t n = 300ull << 20; CK(cudaMalloc(&tup, n * 8)); CK(cudaMemset(tup, 1, n * 8));
            for (int log2s : {18, 21}) {
                const uint64_t slice_slots = 1ull << log2s, nsl = (2ull << 30) / 16 / slice_slots;
                #define RUNW(NAME, ...) CK(cudaMemset(T, 0, 2ull << 30)); CK(cudaEventRecord(a)); __VA_ARGS__; CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); \
                    CK(cudaEventElapsedTime(&ms, a, b)); printf("winsert %-28s slice %4llu MiB : %8.2f G ops/s\n", NAME, (unsigned long long)(slice_slots >> 16), n / (ms * 1e-3) / 1e9);
                RUNW("cas128 claim", (window_insert<0, false><<<148 * 6, 256>>>(T, slice_slots, nsl, n, tup, sink)))
                RUNW("cas64+cas64 claim", (window_insert<1, false><<<148 * 6, 256>>>(T, slice_slots, nsl, n, tup, sink)))
                RUNW("cas128 claim + tuple stream", (window_insert<0, true><<<148 * 6, 256>>>(T, slice_slots, nsl, n, tup, sink)))
                RUNW("cas64+cas64 + tuple stream", (window_insert<1, true><<<148 * 6, 256>>>(T, slice_slots, nsl, n, tup, sink)))
            }
            CK(cudaFree(tup));
        }
        const uint64_t tb = std::min<uint64_t>(16ull << 30, max_bytes);
        CK(cudaEventRecord(a)); chunk_write<128><<<148 * 8, 256>>>((uint64_t *)T, tb / 128, tb / 128); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms, a, b)); printf("random 128-B chunk writes: %8.1f GB/s\n", tb / (ms * 1e-3) / 1e9);
        CK(cudaEventRecord(a)); chunk_write<256><<<148 * 8, 256>>>((uint64_t *)T, tb / 256, tb / 256); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms, a, b)); printf("random 256-B chunk writes: %8.1f GB/s\n", tb / (ms * 1e-3) / 1e9);
        CK(cudaEventRecord(a)); chunk_write<512><<<148 * 8, 256>>>((uint64_t *)T, tb / 512, tb / 512); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms, a, b)); printf("random 512-B chunk writes: %8.1f GB/s\n", tb / (ms * 1e-3) / 1e9);
        CK(cudaMemset(T, 0, max_bytes));
        if (argc > 3) return 0;
    }
    for (double gib : {0.0625, 1.0, 8.0, max_gib}) {
        const uint64_t bytes = (uint64_t)(gib * 1073741824.0);
        if (bytes > max_bytes) continue;
        run<0, 1>("ld16", T, bytes, nops, 8, sink);
        run<0, 4>("ld16", T, bytes, nops, 8, sink);
        run<1, 1>("ld32 (one sector)", T, bytes, nops, 8, sink);
        run<2, 1>("red64", T, bytes, nops, 8, sink);
        run<2, 4>("red64", T, bytes, nops, 8, sink);
        run<3, 1>("atom64 (return)", T, bytes, nops, 8, sink);
        run<3, 4>("atom64 (return)", T, bytes, nops, 8, sink);
        run<7, 1>("cas128 blind", T, bytes, nops, 8, sink);
        run<4, 1>("ld16 + cas128", T, bytes, nops, 8, sink);
        run<4, 4>("ld16 + cas128", T, bytes, nops, 8, sink);
        run<6, 1>("ld16 + cas64", T, bytes, nops, 8, sink);
        run<6, 4>("ld16 + cas64", T, bytes, nops, 8, sink);
        run<5, 1>("ld8 + atom32 + 2 red64 (32B slot)", T, bytes, nops, 6, sink);
        run<5, 4>("ld8 + atom32 + 2 red64 (32B slot)", T, bytes, nops, 6, sink);
        CK(cudaMemset(T, 0, max_bytes));
    }
    return 0;
}

// kq_kernels.h -- host-callable launchers of the sm_100a kernels (kq_kernels.cu).
// Every launcher enqueues on `st`, returns the cudaError_t of the launch and adds the
// number of kernels it launched to *launches.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "kq_device.cuh"

namespace kq {

// record as moved between host and device by load / save / union (48 bytes)
struct XRec { uint64_t key; Rec32 r; uint32_t flags; };
static_assert(sizeof(XRec) == 48, "XRec layout");

// device-side statistics block filled by stats_scan
constexpr int HIST_BINS = 65536;
struct DevStats {
    unsigned long long total, unique, distinct, edges;   // over non-placeholder T entries (logical records)
    unsigned long long big_n;                            // entries with cov >= HIST_BINS appended to big[]
};

struct RouteArgs {           // destination of extracted tuples (multi-GPU count)
    uint64_t *keys; uint8_t *edges; unsigned long long cap_per_rank; unsigned long long *counts;
    int nranks, my_rank, shift;   // owner = (key & 127) >> shift; my_rank < 0: route everything
};

int sm_count();   // multiprocessors of the current device (cached)

// A2+A3 fused: positions [pos0,pos0+npos) of a packed sequence -> table (owned keys only)
cudaError_t launch_count_packed(cudaStream_t st, const TableView &tv, const uint64_t *B, const uint64_t *M,
                                uint64_t pos0, uint64_t npos, unsigned long long *instances, uint64_t *launches);
// A2 with routing: owned tuples are inserted in place, the others appended per owner rank
cudaError_t launch_extract_route(cudaStream_t st, const TableView &tv, const uint64_t *B, const uint64_t *M,
                                 uint64_t pos0, uint64_t npos, RouteArgs ra, unsigned long long *instances,
                                 uint64_t *launches);
// A3: (key, edgeBit) tuples -> table
cudaError_t launch_insert_tuples(cudaStream_t st, const TableView &tv, const uint64_t *keys, const uint8_t *edges,
                                 uint64_t n, unsigned long long *instances, uint64_t *launches);
// A5: positions [c0, c0+n) of a segment with kcount windows; the packed arrays start at segment
// position `base0` (so bit position of c is c-base0); out is indexed by c-c0.
cudaError_t launch_validate(cudaStream_t st, const TableView &tv, const uint64_t *B, const uint64_t *M,
                            uint64_t base0, uint64_t c0, uint64_t n, uint64_t kcount, uint32_t cutoff,
                            Base *out, unsigned long long *acc3, uint64_t *launches);
// point lookups (host consumers)
cudaError_t launch_lookup(cudaStream_t st, const TableView &tv, const uint64_t *keys, uint64_t n, Rec32 *rec,
                          uint8_t *found, uint64_t *launches);
// A3s: scan T
cudaError_t launch_stats(cudaStream_t st, const TableView &tv, DevStats *ds, unsigned long long *hist,
                         unsigned long long *big, unsigned long long big_cap, uint64_t *launches);
// export occupied slots (logical record, or raw lanes + F_PLACEHOLDER) in arbitrary order
cudaError_t launch_export(cudaStream_t st, const TableView &tv, XRec *out, unsigned long long cap,
                          unsigned long long *n_out, uint64_t *launches);
// move every slot of `from` into tv (grow)
cudaError_t launch_rehash(cudaStream_t st, const TableView &tv, const Slot *from, unsigned long long from_cap,
                          uint64_t *launches);
cudaError_t launch_side_rehash(cudaStream_t st, const TableView &tv, const SideSlot *from,
                               unsigned long long from_cap, uint64_t *launches);
// load: unique keys, raw 8-bit-map entries (flags carry F_PLACEHOLDER) -> T
cudaError_t launch_put(cudaStream_t st, const TableView &tv, const XRec *in, uint64_t n, uint64_t *launches);
// exact union: add logical records
cudaError_t launch_add_records(cudaStream_t st, const TableView &tv, const XRec *in, uint64_t n, uint64_t *launches);
// reference-compatible union: DBG::mergeSubMaps for the 8-bit entries of one later database
cudaError_t launch_merge_ref(cudaStream_t st, const TableView &tv, const XRec *in, uint64_t n, uint64_t *launches);

}  // namespace kq

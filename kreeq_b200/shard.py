"""Multi-GPU plumbing of the count / validate path: one process per GPU, key space partitioned by
owner(key) = (key % 128) / (128 / nranks), (key, edgeBit) tuples routed to their
owner with one all-to-all per round over torch.distributed (NCCL on GPUs, gloo in the CPU tests).

Only the exchange lives here; extraction / insertion are the library's kernels (kq_dev_extract_route,
kq_dev_insert_tuples)."""
import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from ._lib import check, lib


def owner_of(keys, nranks):
    """Vectorised kq_owner_of for numpy / torch integer arrays of canonical keys."""
    if isinstance(keys, torch.Tensor):
        return (keys & 127) // (128 // nranks)
    return ((np.asarray(keys, dtype=np.uint64) % np.uint64(128)) // np.uint64(128 // nranks)).astype(np.int64)


def exchange_counts(send_counts, group=None):
    """Every rank learns how many tuples each peer sends it: all-gather of the per-destination counts
    (world x world int64, tiny); returns this rank's column (same device)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    rows = [torch.empty_like(send_counts) for _ in range(world)]
    dist.all_gather(rows, send_counts, group=group)
    return torch.stack(rows)[:, rank].contiguous()


def exchange_tuples(send_keys, send_edges, send_counts_host, recv_counts_host, recv_keys, recv_edges, group=None):
    """Route tuples to their owners: one grouped send/recv per peer (ncclGroupStart ... ncclSend/ncclRecv
    ... ncclGroupEnd under NCCL; plain point-to-point under gloo in the CPU tests).

    send_keys [nranks, cap] int64 / send_edges [nranks, cap] uint8 hold, for destination r, the first
    send_counts_host[r] tuples of slab r.  Received tuples are packed back to back into recv_keys /
    recv_edges (1-D, large enough); returns the number received."""
    n = len(send_counts_host)
    me = dist.get_rank(group)
    offs = np.concatenate(([0], np.cumsum(recv_counts_host))).astype(np.int64)
    ops = []
    for r in range(n):
        sc, rc = int(send_counts_host[r]), int(recv_counts_host[r])
        dst_k, dst_e = recv_keys[int(offs[r]):int(offs[r + 1])], recv_edges[int(offs[r]):int(offs[r + 1])]
        if r == me:
            dst_k.copy_(send_keys[r, :sc]); dst_e.copy_(send_edges[r, :sc])
            continue
        if sc:
            ops.append(dist.P2POp(dist.isend, send_keys[r, :sc], r, group))
            ops.append(dist.P2POp(dist.isend, send_edges[r, :sc], r, group))
        if rc:
            ops.append(dist.P2POp(dist.irecv, dst_k, r, group))
            ops.append(dist.P2POp(dist.irecv, dst_e, r, group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return int(offs[-1])


class RoutedCounter:
    """Count k-mers of device-resident packed reads into a key-sharded table, `round_pos` k-mer start
    positions per exchange round."""

    def __init__(self, dbg, rank, nranks, round_pos, device, slack=1.25):
        self.g, self.rank, self.n, self.round_pos = dbg, rank, nranks, int(round_pos)
        self.cap = int(self.round_pos / nranks * slack) + 1024          # tuples per destination per round
        self.send_keys = torch.empty((nranks, self.cap), dtype=torch.int64, device=device)
        self.send_edges = torch.empty((nranks, self.cap), dtype=torch.uint8, device=device)
        self.counts = torch.zeros(nranks, dtype=torch.int64, device=device)
        rcap = int(self.round_pos * slack) + 1024 * nranks
        self.recv_keys = torch.empty(rcap, dtype=torch.int64, device=device)
        self.recv_edges = torch.empty(rcap, dtype=torch.uint8, device=device)
        self.a2a_bytes = 0

    def count(self, d_bases, d_invalid, length):
        """d_bases / d_invalid: int64 device tensors holding the packed sequence of `length` bases."""
        g, L = self.g, lib()
        k = g.k
        if length < k:
            return
        total = length - k + 1
        for p0 in range(0, total, self.round_pos):
            npos = min(self.round_pos, total - p0)
            self.counts.zero_()
            check(L.kq_dev_extract_route(g.handle, C.c_void_p(d_bases.data_ptr()), C.c_void_p(d_invalid.data_ptr()),
                                         length, p0, npos, C.c_void_p(self.send_keys.data_ptr()),
                                         C.c_void_p(self.send_edges.data_ptr()), self.cap,
                                         C.c_void_p(self.counts.data_ptr())), g.handle)
            recv_counts = exchange_counts(self.counts)
            sc = self.counts.cpu().numpy()
            rc = recv_counts.cpu().numpy()
            if sc.max() > self.cap or rc.sum() > self.recv_keys.numel():
                raise RuntimeError("route buffers too small: %s > %d" % (sc, self.cap))
            nrecv = exchange_tuples(self.send_keys, self.send_edges, sc, rc, self.recv_keys, self.recv_edges)
            self.a2a_bytes += int(sc.sum()) * 9
            check(L.kq_dev_insert_tuples(g.handle, C.c_void_p(self.recv_keys.data_ptr()),
                                         C.c_void_p(self.recv_edges.data_ptr()), nrecv), g.handle)

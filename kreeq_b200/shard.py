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


def exchange_tuples(send, send_counts_host, recv_counts_host, recv, group=None):
    """Route tuples to their owners: one grouped send/recv per peer (ncclGroupStart ... ncclSend/ncclRecv
    ... ncclGroupEnd under NCCL; plain point-to-point under gloo in the CPU tests).

    send [nranks, cap, tuple_bytes] uint8 holds, for destination r, the first send_counts_host[r] tuples of
    slab r (the layout kq_dev_extract_route writes).  Received tuples are packed back to back into
    recv [N, tuple_bytes] (large enough); returns the number received."""
    n = len(send_counts_host)
    me = dist.get_rank(group)
    offs = np.concatenate(([0], np.cumsum(recv_counts_host))).astype(np.int64)
    ops = []
    for r in range(n):
        sc, rc = int(send_counts_host[r]), int(recv_counts_host[r])
        dst = recv[int(offs[r]):int(offs[r + 1])]
        if r == me:
            dst.copy_(send[r, :sc])
            continue
        if sc:
            ops.append(dist.P2POp(dist.isend, send[r, :sc], r, group))
        if rc:
            ops.append(dist.P2POp(dist.irecv, dst, r, group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return int(offs[-1])


def pack_tuples(keys, edges, k):
    """numpy (keys u64, edges u8) -> [n, tuple_bytes] uint8 in the device tuple format (kq_tuple_bytes)."""
    keys = np.asarray(keys, dtype=np.uint64); edges = np.asarray(edges, dtype=np.uint64)
    if 2 * k + 8 <= 64:
        return ((keys << np.uint64(8)) | edges).view(np.uint8).reshape(-1, 8)
    return np.stack((keys, edges), axis=1).view(np.uint8).reshape(-1, 16)


def unpack_tuples(raw, k):
    raw = np.ascontiguousarray(raw)
    if 2 * k + 8 <= 64:
        w = raw.view(np.uint64).reshape(-1)
        return w >> np.uint64(8), (w & np.uint64(255)).astype(np.uint8)
    w = raw.view(np.uint64).reshape(-1, 2)
    return w[:, 0].copy(), w[:, 1].astype(np.uint8)


class RoutedCounter:
    """Count k-mers of device-resident packed reads into a key-sharded table, `round_pos` k-mer start
    positions per exchange round."""

    def __init__(self, dbg, rank, nranks, round_pos, device, slack=1.25):
        self.g, self.rank, self.n, self.round_pos = dbg, rank, nranks, int(round_pos)
        self.tb = lib().kq_tuple_bytes(dbg.handle)
        self.cap = int(self.round_pos / nranks * slack) + 1024          # tuples per destination per round
        self.send = torch.empty((nranks, self.cap, self.tb), dtype=torch.uint8, device=device)
        self.counts = torch.zeros(nranks, dtype=torch.int64, device=device)
        rcap = int(self.round_pos * slack) + 1024 * nranks
        self.recv = torch.empty((rcap, self.tb), dtype=torch.uint8, device=device)
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
                                         length, p0, npos, C.c_void_p(self.send.data_ptr()), self.cap,
                                         C.c_void_p(self.counts.data_ptr())), g.handle)
            recv_counts = exchange_counts(self.counts)
            sc = self.counts.cpu().numpy()
            rc = recv_counts.cpu().numpy()
            if sc.max() > self.cap or rc.sum() > self.recv.shape[0]:
                raise RuntimeError("route buffers too small: %s > %d" % (sc, self.cap))
            nrecv = exchange_tuples(self.send, sc, rc, self.recv)
            self.a2a_bytes += (int(sc.sum()) - int(sc[self.rank])) * self.tb
            check(L.kq_dev_insert_tuples(g.handle, C.c_void_p(self.recv.data_ptr()), nrecv), g.handle)

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

    def count(self, d_bases, d_invalid, length, rounds=None):
        """d_bases / d_invalid: int64 device tensors holding the packed sequence of `length` bases.
        Collective: every rank must run the same number of rounds (pass `rounds` when inputs differ)."""
        g, L = self.g, lib()
        k = g.k
        total = max(0, length - k + 1)
        if rounds is None:
            rounds = (total + self.round_pos - 1) // self.round_pos
        for r in range(rounds):
            p0 = min(r * self.round_pos, total)
            npos = min(self.round_pos, total - p0)
            self.counts.zero_()
            check(L.kq_dev_extract_route(g.handle, C.c_void_p(d_bases.data_ptr()), C.c_void_p(d_invalid.data_ptr()),
                                         length, p0, npos, C.c_void_p(self.send.data_ptr()), self.cap,
                                         C.c_void_p(self.counts.data_ptr())), g.handle)
            recv_counts = exchange_counts(self.counts)
            sc = self.counts.cpu().numpy()
            rc = recv_counts.cpu().numpy()
            # the verdict is collective: a rank that raised alone would leave the others waiting in the exchange
            bad = torch.tensor([1 if (sc.max() > self.cap or rc.sum() > self.recv.shape[0]) else 0], dtype=torch.int32, device=self.counts.device)
            dist.all_reduce(bad, op=dist.ReduceOp.MAX, group=getattr(self, "group", None))
            if int(bad.item()):
                raise RuntimeError("route buffers too small on some rank (here: %s against %d per peer)" % (sc, self.cap))
            nrecv = exchange_tuples(self.send, sc, rc, self.recv)
            self.a2a_bytes += (int(sc.sum()) - int(sc[self.rank])) * self.tb
            check(L.kq_dev_insert_tuples(g.handle, C.c_void_p(self.recv.data_ptr()), nrecv), g.handle)


class PeerRoutedCounter:
    """Same job as RoutedCounter with the exchange fused into the scatter kernel: every rank's scatter
    writes the tuples a peer owns straight into that peer's receive slab (CUDA IPC peer mapping, NVLink),
    already in slice order, so a round is  histogram -> all-gather of the histograms (tiny) -> scatter
    over peer memory -> barrier -> insert;  no tuple ever goes through a send buffer or a second scatter.

    `stream` is the torch stream both the library's kernels and the NCCL collectives are ordered on."""

    def __init__(self, dbg, rank, nranks, round_pos, device, slack=1.3, nb_bits=None, stream=None, group=None):
        L = lib()
        self.g, self.rank, self.n, self.round_pos, self.group = dbg, rank, nranks, int(round_pos), group
        self.device = device
        self.stream = stream if stream is not None else torch.cuda.Stream(device)
        dbg.set_stream(self.stream.cuda_stream)
        if nb_bits is None:                       # 1024 bins in all: 128-byte flushes (2048 for >= 4 ranks: 64-byte)
            nb_bits = ((1024 if nranks <= 2 else 2048) // nranks).bit_length() - 1
        self.slab_cap = int(self.round_pos * slack) + 4096
        slab, handle = C.c_void_p(), (C.c_ubyte * 64)()
        check(L.kq_route_setup(dbg.handle, self.slab_cap, nb_bits, C.byref(slab), handle), dbg.handle)
        with torch.cuda.stream(self.stream):
            mine = torch.tensor(list(handle), dtype=torch.uint8, device=device)
            handles = torch.empty(nranks * 64, dtype=torch.uint8, device=device)
            dist.all_gather_into_tensor(handles, mine, group=group)
            handles = handles.cpu().numpy().reshape(nranks, 64)
        ptrs = (C.c_void_p * nranks)()
        for r in range(nranks):
            if r == rank:
                ptrs[r] = slab.value
            else:
                p = C.c_void_p()
                check(L.kq_route_open(dbg.handle, C.c_void_p(handles[r].ctypes.data), C.byref(p)), dbg.handle)
                ptrs[r] = p.value
        check(L.kq_route_connect(dbg.handle, ptrs), dbg.handle)
        self.bins = L.kq_route_bins(dbg.handle)
        self.hist = torch.zeros(self.bins, dtype=torch.int64, device=device)
        self.all_hist = torch.zeros(nranks * self.bins, dtype=torch.int64, device=device)
        self.flag = torch.zeros(1, dtype=torch.int32, device=device)
        self.rounds = torch.zeros(1, dtype=torch.int64, device=device)
        self.a2a_bytes = 0
        self.tb = L.kq_tuple_bytes(dbg.handle)

    def count(self, d_bases, d_invalid, length):
        """Collective: every rank calls it (ranks with fewer rounds take part with empty ones)."""
        g, L = self.g, lib()
        total = max(0, length - g.k + 1)
        mine = (total + self.round_pos - 1) // self.round_pos
        caller = torch.cuda.current_stream(self.device)
        self.stream.wait_stream(caller)
        with torch.cuda.stream(self.stream):
            self.rounds.fill_(mine)
            dist.all_reduce(self.rounds, op=dist.ReduceOp.MAX, group=self.group)
            rounds = int(self.rounds.item())
            pb, pi = C.c_void_p(d_bases.data_ptr()), C.c_void_p(d_invalid.data_ptr())
            for r in range(rounds):
                p0 = min(r * self.round_pos, total)
                npos = min(self.round_pos, total - p0)
                check(L.kq_route_hist(g.handle, pb, pi, length, p0, npos, C.c_void_p(self.hist.data_ptr())), g.handle)
                # also orders this round after every rank's previous insert: the slabs are free again
                dist.all_gather_into_tensor(self.all_hist, self.hist, group=self.group)
                nrecv = C.c_uint64()
                check(L.kq_route_scatter(g.handle, pb, pi, length, p0, npos, C.c_void_p(self.all_hist.data_ptr()),
                                         C.byref(nrecv)), g.handle)
                dist.all_reduce(self.flag, group=self.group)         # barrier on the stream: every slab is complete
                check(L.kq_route_insert(g.handle, nrecv.value), g.handle)
                self.a2a_bytes += nrecv.value * self.tb
        caller.wait_stream(self.stream)


class PeerTables:
    """`with PeerTables(dbg, nranks, device):` -- inside the block every rank's validate calls answer all
    positions, probing each k-mer in its owner's shard over peer memory (kq_peer_export / kq_peer_connect).
    Entry and exit are collective.  With keep=True the peer mappings stay open after the block (mapping a
    peer's tables costs milliseconds; a pipeline maps them once): the tables may then be cleared and counted
    into again between blocks as long as they do not have to grow (the library refuses to move an exported
    table: KQ_ESTATE), and disconnect() must be called by every rank at the end.

    The library's stream must be torch's current stream: the collectives order the kernels of all ranks."""

    def __init__(self, dbg, nranks, device, keep=False, group=None):
        self.g, self.n, self.device, self.keep, self.group = dbg, nranks, device, keep, group
        self.connected = False
        self.flag = torch.zeros(1, dtype=torch.int32, device=device)
        self.connect_ms = 0.0

    def _stream_barrier(self):
        dist.all_reduce(self.flag, group=self.group)

    def connect(self):
        import time
        L = lib()
        t0 = time.perf_counter()
        blob = (C.c_ubyte * 256)()
        check(L.kq_peer_export(self.g.handle, blob), self.g.handle)
        mine = torch.tensor(list(blob), dtype=torch.uint8, device=self.device)
        blobs = torch.empty(self.n * 256, dtype=torch.uint8, device=self.device)
        dist.all_gather_into_tensor(blobs, mine, group=self.group)      # also: every rank's table is complete
        self._blobs = blobs.cpu().numpy()
        check(L.kq_peer_connect(self.g.handle, C.c_void_p(self._blobs.ctypes.data)), self.g.handle)
        self.connected = True
        self.connect_ms = (time.perf_counter() - t0) * 1e3

    def disconnect(self):
        if not self.connected:
            return
        torch.cuda.current_stream(self.device).synchronize()
        dist.barrier(group=self.group)                   # nobody is still probing this rank's tables
        check(lib().kq_peer_disconnect(self.g.handle), self.g.handle)
        dist.barrier(group=self.group)
        self.connected = False

    def __enter__(self):
        if self.connected:
            check(lib().kq_sync(self.g.handle), self.g.handle)          # pending reads counted
            self._stream_barrier()                       # every rank's table is complete before anyone probes
        else:
            self.connect()
        return self

    def __exit__(self, *exc):
        if self.keep and exc[0] is None:
            self._stream_barrier()                       # nobody still probes before a table changes again
        else:
            self.disconnect()
        return False


def cut_rounds(segments, chunk):
    """[(.., .., length), ..] -> consecutive groups whose lengths add up to at most `chunk` (a segment longer than
    `chunk` is refused: its rows would not fit the slab)"""
    out, cur, tot = [], [], 0
    for s_ in segments:
        if s_[2] > chunk:
            raise ValueError("segment of %d positions exceeds the round capacity %d" % (s_[2], chunk))
        if cur and tot + s_[2] > chunk:
            out.append(cur); cur, tot = [], 0
        cur.append(s_); tot += s_[2]
    if cur:
        out.append(cur)
    return out


class RoutedValidator:
    """Validate against a key-sharded table with routed queries (kq_vroute_*, SURVEY 8e "Collective"): per chunk of
    assembly positions every rank sends 16-byte (key, position, neighbours) queries into the OWNERS' inboxes (peer
    stores from the send kernel), the owners probe their own HBM and store the finished DBGbase rows into the
    origins' row slabs; two stream barriers per chunk, no rank ever reads another rank's table.

    One process per GPU: inboxes and row slabs are exchanged once as CUDA IPC handles.  Every call is collective
    (ranks without a chunk pass length 0).  The library's stream must be the stream the collectives run on."""

    def __init__(self, dbg, rank, nranks, chunk, device, group=None):
        L = lib()
        self.g, self.rank, self.n, self.chunk, self.device, self.group = dbg, rank, nranks, int(chunk), device, group
        inbox, rows = C.c_void_p(), C.c_void_p()
        h_in, h_rows = (C.c_ubyte * 64)(), (C.c_ubyte * 64)()
        check(L.kq_vroute_setup(dbg.handle, self.chunk, C.byref(inbox), h_in, C.byref(rows), h_rows), dbg.handle)
        mine = torch.tensor(list(h_in) + list(h_rows), dtype=torch.uint8, device=device)
        allh = torch.empty(nranks * 128, dtype=torch.uint8, device=device)
        dist.all_gather_into_tensor(allh, mine, group=group)
        allh = allh.cpu().numpy().reshape(nranks, 2, 64)
        pin, prow = (C.c_void_p * nranks)(), (C.c_void_p * nranks)()
        for r in range(nranks):
            if r == rank:
                pin[r], prow[r] = inbox.value, rows.value
                continue
            for k_, dst in ((0, pin), (1, prow)):
                p = C.c_void_p()
                check(L.kq_route_open(dbg.handle, C.c_void_p(allh[r, k_].ctypes.data), C.byref(p)), dbg.handle)
                dst[r] = p.value
        check(L.kq_vroute_connect(dbg.handle, pin, prow), dbg.handle)
        self.rows_ptr = rows.value
        self.flag = torch.zeros(1, dtype=torch.int32, device=device)
        self.acc = torch.zeros(3, dtype=torch.int64, device=device)
        dist.barrier(group=group)                    # every rank is connected before the first query is stored

    def _barrier(self):
        dist.all_reduce(self.flag, group=self.group)

    def chunk_packed(self, d_bases, d_invalid, length, cutoff=0):
        """One chunk: a whole packed device-resident segment of `length` <= chunk bases (0: take part empty-handed).
        Afterwards rows [0, length) of this rank's slab (rows_ptr, kq_base) hold the segment's DBGbase rows."""
        g, L = self.g, lib()
        kcount = max(0, length - g.k + 1)
        pb = C.c_void_p(d_bases.data_ptr()) if length else None
        pi = C.c_void_p(d_invalid.data_ptr()) if length else None
        check(L.kq_vroute_send(g.handle, pb, pi, 0, 0, length, kcount), g.handle)
        self._barrier()                               # every inbox is complete
        check(L.kq_vroute_answer(g.handle, cutoff), g.handle)
        self._barrier()                               # every answer box is complete, the inboxes are free again
        check(L.kq_vroute_place(g.handle), g.handle)  # answers -> rows (local)

    def round_packed(self, segments, cutoff=0):
        """One round over several packed device-resident segments [(d_bases, d_invalid, length), ..] whose lengths add up
        to at most `chunk`: every segment's queries are sent, then ONE pair of stream barriers serves them all.
        Afterwards the rows of segment i start at row sum(lengths[:i]) of this rank's slab.  Collective (ranks without
        work pass [])."""
        g, L = self.g, lib()
        row0 = 0
        segs = [s_ for s_ in segments if s_[2]]
        for i, (d_bases, d_invalid, length) in enumerate(segs):
            check(L.kq_vroute_send_at(g.handle, C.c_void_p(d_bases.data_ptr()), C.c_void_p(d_invalid.data_ptr()), 0, 0, length,
                                      max(0, length - g.k + 1), row0, 1 if i == len(segs) - 1 else 0), g.handle)
            row0 += length
        if not segs:
            check(L.kq_vroute_send(g.handle, None, None, 0, 0, 0, 0), g.handle)
        self._barrier()                               # every inbox is complete
        check(L.kq_vroute_answer(g.handle, cutoff), g.handle)
        self._barrier()                               # every answer box is complete, the inboxes are free again
        check(L.kq_vroute_place(g.handle), g.handle)  # answers -> rows (local)

    def rounds_of(self, segments):
        """cut a list of (d_bases, d_invalid, length) into rounds of at most `chunk` positions"""
        return cut_rounds(segments, self.chunk)

    def counters(self, reset=True):
        """the three QV counters summed over all ranks (collective)"""
        from ._lib import KqQV
        q = KqQV()
        check(lib().kq_vroute_counters(self.g.handle, C.byref(q), 1 if reset else 0), self.g.handle)
        t = torch.tensor([q.missing, q.kcount, q.edge_missing], dtype=torch.int64, device=self.device)
        dist.all_reduce(t, group=self.group)
        return [int(x) for x in t.tolist()]

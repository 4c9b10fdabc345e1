"""Host-side mirror of the reference's DBG / Kmap interface for the k-mer hot path.

Method names follow the reference members they stand for (paths relative to the reference tree):

  DBG(k, ...)              DBG knav(userInput)                     src/input.cpp:88
  traverseInReads(batch)   Kmap::traverseInReads(std::string*)     gfalibs/include/kmer.h:753-764
  finalize()               Kmap::finalize                          kmer.h:821-828
  stats() / hist()         Kmap::stats + DBG::summary/DBstats      kmer.h:831-852, src/graph-builder.cpp:240-295
  save(dir) / load(dir)    dumpMap/dumpHighCopyKmers, loadGraph    kmer.h:698-709,563-576, src/input.cpp:56-74
  kunion(dirs)             DBG::kunion                             src/graph-builder.cpp:297-351
  evaluateSegment(seq)     DBG::evaluateSegment                    src/kreeq.cpp:110-229
  find(keys)               maps[i]->find(key) idiom                src/kreeq.cpp:150-166

Everything runs in libkreeq_b200.so on the GPU; this file only marshals buffers.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import BASE, REC32, KqQV, KqStats, check, lib, ptr


class DBG:
    def __init__(self, k=21, device=0, expected_distinct=0, rank=0, nranks=1):
        self.k = int(k)
        self._h = C.c_void_p()
        rc = lib().kq_create(C.byref(self._h), self.k, int(device), int(expected_distinct))
        if rc != 0:
            self._h = C.c_void_p()
            check(rc, None)
        self.rank, self.nranks = rank, nranks
        if nranks > 1:
            check(lib().kq_set_shard(self._h, rank, nranks), self._h)
        self.tot_missing = 0          # totMissingKmers   include/kreeq.h:152
        self.tot_kcount = 0           # totKcount
        self.tot_edge_missing = 0     # totEdgeMissingKmers

    # -- lifetime ---------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().kq_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def handle(self):
        return self._h

    def set_stream(self, cuda_stream):
        check(lib().kq_set_stream(self._h, C.c_void_p(cuda_stream or 0)), self._h)

    def sync(self):
        check(lib().kq_sync(self._h), self._h)

    def clear(self):
        check(lib().kq_clear(self._h), self._h)
        self.tot_missing = self.tot_kcount = self.tot_edge_missing = 0

    def set_flush_positions(self, positions):
        check(lib().kq_set_flush_positions(self._h, int(positions)), self._h)

    def reserve(self, extra_distinct):
        check(lib().kq_reserve(self._h, int(extra_distinct)), self._h)

    # -- count --------------------------------------------------------------------------------
    def traverseInReads(self, batch):
        """One 'N'-joined read batch (bytes / bytearray / uint8 array)."""
        if isinstance(batch, np.ndarray):
            a = np.ascontiguousarray(batch, dtype=np.uint8)
            check(lib().kq_count_push(self._h, ptr(a), a.size), self._h)
        else:
            b = bytes(batch)
            check(lib().kq_count_push(self._h, C.cast(C.c_char_p(b), C.c_void_p), len(b)), self._h)

    def finalize(self):
        check(lib().kq_count_finish(self._h), self._h)

    def instances(self):
        n = C.c_uint64()
        check(lib().kq_count_instances(self._h, C.byref(n)), self._h)
        return n.value

    def distinct(self):
        n = C.c_uint64()
        check(lib().kq_distinct(self._h, C.byref(n)), self._h)
        return n.value

    # -- persistence / union --------------------------------------------------------------------
    def save(self, db_dir):
        check(lib().kq_db_save(self._h, str(db_dir).encode()), self._h)

    def load(self, db_dir):
        check(lib().kq_db_load(self._h, str(db_dir).encode()), self._h)

    def kunion(self, db_dirs, ref_compatible=True):
        arr = (C.c_char_p * len(db_dirs))(*[str(d).encode() for d in db_dirs])
        check(lib().kq_union(self._h, arr, len(db_dirs), 1 if ref_compatible else 0), self._h)

    # -- stats ---------------------------------------------------------------------------------
    def stats(self):
        s = KqStats()
        check(lib().kq_stats_get(self._h, C.byref(s)), self._h)
        return {n: getattr(s, n) for n, _ in KqStats._fields_}

    def hist(self):
        n = C.c_size_t(0)
        check(lib().kq_hist(self._h, None, None, C.byref(n)), self._h)
        cov = np.zeros(n.value, dtype=np.uint64)
        cnt = np.zeros(n.value, dtype=np.uint64)
        check(lib().kq_hist(self._h, ptr(cov), ptr(cnt), C.byref(n)), self._h)
        return cov, cnt

    def export(self):
        """Lookup-visible {key -> record} set of this shard, sorted by key."""
        n = C.c_size_t(0)
        check(lib().kq_export(self._h, None, None, C.byref(n)), self._h)
        keys = np.zeros(n.value, dtype=np.uint64)
        recs = np.zeros(n.value, dtype=REC32)
        check(lib().kq_export(self._h, ptr(keys), ptr(recs), C.byref(n)), self._h)
        return keys, recs

    # -- validate ------------------------------------------------------------------------------
    def evaluateSegment(self, seq, cov_cutoff=0, out=None):
        """Returns the DBGbase array of one segment and accumulates the three QV counters.  `out` may be a
        caller-owned BASE array (e.g. a view of page-locked memory, which the library fills directly)."""
        if isinstance(seq, np.ndarray):
            a = np.ascontiguousarray(seq, dtype=np.uint8)
        else:
            a = np.frombuffer(bytes(seq), dtype=np.uint8)
        if out is None:
            out = np.zeros(a.size, dtype=BASE)
        elif out.dtype != BASE or out.size < a.size or not out.flags["C_CONTIGUOUS"]:
            raise ValueError("out must be a contiguous BASE array of at least len(seq) entries")
        q = KqQV()
        check(lib().kq_validate_segment(self._h, ptr(a), a.size, int(cov_cutoff), ptr(out), C.byref(q)), self._h)
        self.tot_missing += q.missing
        self.tot_kcount += q.kcount
        self.tot_edge_missing += q.edge_missing
        return out[:a.size]

    def qv_rows(self):
        """(missing, total, QV, error) for the Merqury and Kreeq rows, src/kreeq.cpp:87-104."""
        L = lib()
        rows = []
        for miss in (self.tot_missing, self.tot_missing + self.tot_edge_missing):
            rows.append((miss, self.tot_kcount, L.kq_qv_value(miss, self.tot_kcount, self.k),
                         L.kq_error_rate(miss, self.tot_kcount, self.k)))
        return rows

    def find(self, keys):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        rec = np.zeros(keys.size, dtype=REC32)
        found = np.zeros(keys.size, dtype=np.uint8)
        check(lib().kq_lookup_batch(self._h, ptr(keys), keys.size, ptr(rec), ptr(found)), self._h)
        return rec, found.astype(bool)

    # -- counters for bench.py -------------------------------------------------------------------
    def counters(self):
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(lib().kq_counters(self._h, C.byref(a), C.byref(b), C.byref(c)), self._h)
        return {"kernel_launches": a.value, "h2d_bytes": b.value, "d2h_bytes": c.value}

    def set_timing(self, on):
        check(lib().kq_set_timing(self._h, 1 if on else 0), self._h)

    def kernel_time_ms(self, which, reset=True):
        ms, n = C.c_double(), C.c_uint64()
        check(lib().kq_kernel_time_ms(self._h, which, C.byref(ms), C.byref(n), 1 if reset else 0), self._h)
        return ms.value, n.value


def pack_host(ascii_bytes, nthreads=0):
    """2-bit pack on the host (kq_pack_host): returns (bases u64[], invalid u64[])."""
    a = np.frombuffer(bytes(ascii_bytes), dtype=np.uint8) if not isinstance(ascii_bytes, np.ndarray) else \
        np.ascontiguousarray(ascii_bytes, dtype=np.uint8)
    words = lib().kq_packed_words(a.size)
    bases = np.zeros(words, dtype=np.uint64)
    inv = np.zeros(words, dtype=np.uint64)
    check(lib().kq_pack_host(ptr(a), a.size, ptr(bases), ptr(inv), nthreads))
    return bases, inv


def db_write_records(db_dir, k, keys, recs):
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    recs = np.ascontiguousarray(recs, dtype=REC32)
    check(lib().kq_db_write_records(str(db_dir).encode(), int(k), ptr(keys), ptr(recs), keys.size))


def db_read_records(db_dir):
    n = C.c_size_t(0)
    check(lib().kq_db_read_records(str(db_dir).encode(), None, None, C.byref(n)))
    keys = np.zeros(n.value, dtype=np.uint64)
    recs = np.zeros(n.value, dtype=REC32)
    check(lib().kq_db_read_records(str(db_dir).encode(), ptr(keys), ptr(recs), C.byref(n)))
    return keys, recs


class MultiDBG:
    """Same interface as DBG over kq_multi: one handle, `ngpus` shards of the key space, one per GPU of this
    process (devices = CUDA ordinals, may repeat: every shard on one GPU runs the same code path).  Replaces the
    reference's 128 host maps (gfalibs/include/kmer.h:126-153) by shard g = maps [g*128/G, (g+1)*128/G)."""

    def __init__(self, k=21, ngpus=0, devices=None, expected_distinct=0, round_positions=0, validate_chunk=0):
        self.k = int(k)
        self._h = C.c_void_p()
        dev = None
        if devices is not None:
            ngpus = len(devices)
            dev = (C.c_int * ngpus)(*devices)
        rc = lib().kq_multi_create(C.byref(self._h), self.k, int(ngpus), dev, int(expected_distinct))
        if rc != 0:
            self._h = C.c_void_p()
            msg = lib().kq_multi_last_error(None)
            raise _lib.KreeqError(rc, msg.decode() if msg else "")
        self.ngpus = lib().kq_multi_ngpus(self._h)
        if round_positions or validate_chunk:
            self._check(lib().kq_multi_set_round(self._h, int(round_positions), int(validate_chunk)))
        self.tot_missing = self.tot_kcount = self.tot_edge_missing = 0

    def _check(self, rc):
        if rc != 0:
            msg = lib().kq_multi_last_error(self._h)
            raise _lib.KreeqError(rc, msg.decode() if msg else "")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().kq_multi_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def shard(self, g):
        return C.c_void_p(lib().kq_multi_shard(self._h, g))

    def traverseInReads(self, batch):
        if isinstance(batch, np.ndarray):
            a = np.ascontiguousarray(batch, dtype=np.uint8)
            self._check(lib().kq_multi_count_push(self._h, ptr(a), a.size))
        else:
            b = bytes(batch)
            self._check(lib().kq_multi_count_push(self._h, C.cast(C.c_char_p(b), C.c_void_p), len(b)))

    def finalize(self):
        self._check(lib().kq_multi_count_finish(self._h))

    def clear(self):
        self._check(lib().kq_multi_clear(self._h))
        self.tot_missing = self.tot_kcount = self.tot_edge_missing = 0

    def instances(self):
        n = C.c_uint64()
        self._check(lib().kq_multi_count_instances(self._h, C.byref(n)))
        return n.value

    def stats(self):
        s = KqStats()
        self._check(lib().kq_multi_stats_get(self._h, C.byref(s)))
        return {n: getattr(s, n) for n, _ in KqStats._fields_}

    def hist(self):
        n = C.c_size_t(0)
        self._check(lib().kq_multi_hist(self._h, None, None, C.byref(n)))
        cov = np.zeros(n.value, dtype=np.uint64)
        cnt = np.zeros(n.value, dtype=np.uint64)
        self._check(lib().kq_multi_hist(self._h, ptr(cov), ptr(cnt), C.byref(n)))
        return cov, cnt

    def export(self):
        n = C.c_size_t(0)
        self._check(lib().kq_multi_export(self._h, None, None, C.byref(n)))
        keys = np.zeros(n.value, dtype=np.uint64)
        recs = np.zeros(n.value, dtype=REC32)
        self._check(lib().kq_multi_export(self._h, ptr(keys), ptr(recs), C.byref(n)))
        return keys, recs

    def save(self, db_dir):
        self._check(lib().kq_multi_db_save(self._h, str(db_dir).encode()))

    def load(self, db_dir):
        self._check(lib().kq_multi_db_load(self._h, str(db_dir).encode()))

    def kunion(self, db_dirs, ref_compatible=True):
        arr = (C.c_char_p * len(db_dirs))(*[str(d).encode() for d in db_dirs])
        self._check(lib().kq_multi_union(self._h, arr, len(db_dirs), 1 if ref_compatible else 0))

    def evaluateSegment(self, seq, cov_cutoff=0, out=None):
        a = np.ascontiguousarray(seq, dtype=np.uint8) if isinstance(seq, np.ndarray) else np.frombuffer(bytes(seq), dtype=np.uint8)
        if out is None:
            out = np.zeros(a.size, dtype=BASE)
        q = KqQV()
        self._check(lib().kq_multi_validate_segment(self._h, ptr(a), a.size, int(cov_cutoff), ptr(out), C.byref(q)))
        self.tot_missing += q.missing
        self.tot_kcount += q.kcount
        self.tot_edge_missing += q.edge_missing
        return out[:a.size]

    def find(self, keys):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        rec = np.zeros(keys.size, dtype=REC32)
        found = np.zeros(keys.size, dtype=np.uint8)
        self._check(lib().kq_multi_lookup_batch(self._h, ptr(keys), keys.size, ptr(rec), ptr(found)))
        return rec, found.astype(bool)

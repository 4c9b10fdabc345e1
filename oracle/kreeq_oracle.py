"""ctypes face of oracle/libkreeq_oracle.so + a numpy decoder of the .kreeq layout.

TEST INFRASTRUCTURE ONLY (see oracle/kreeq_oracle.h).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the
product package kreeq_b200 never does.

Parity status: PINNED (tests/test_oracle_golden.py).

File-format references (relative to /root/reference):
  gfalibs/include/parallel-hashmap/phmap_dump.h:52-67,104-124  (dump layout)
  gfalibs/include/parallel-hashmap/phmap_utils.h:66-77         (phmap_mix<8>)
  gfalibs/include/parallel-hashmap/phmap.h:357-364,3840-3842   (H1/H2, subidx)
  gfalibs/include/parallel-hashmap/phmap.h:144-168,2199-2213   (probe_seq, find)
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libkreeq_oracle.so")

REC8 = np.dtype([("fw", "u1", 4), ("bw", "u1", 4), ("cov", "u1")])
REC32 = np.dtype([("fw", "<u4", 4), ("bw", "<u4", 4), ("cov", "<u4")])
BASE = np.dtype([("fw", "<u4"), ("bw", "<u4"), ("cov", "<u4"), ("isFw", "u1"), ("pad", "u1", 3)])
SLOT8 = np.dtype([("key", "<u8"), ("fw", "u1", 4), ("bw", "u1", 4), ("cov", "u1"), ("pad", "u1", 7)])
SLOT32 = np.dtype([("key", "<u8"), ("fw", "<u4", 4), ("bw", "<u4", 4), ("cov", "<u4"), ("pad", "u1", 4)])
assert REC8.itemsize == 9 and REC32.itemsize == 36 and BASE.itemsize == 16
assert SLOT8.itemsize == 24 and SLOT32.itemsize == 48

MAP_COUNT = 128          # gfalibs/include/kmer.h:68
SUBMAPS = 256            # PM<> N=8, include/kreeq.h:138-148
GROUP_WIDTH = 16         # SSE2 Group::kWidth, phmap.h:405
VERSION = 0xFFFFFFFFFFFFFFF5
K_EMPTY, K_DELETED, K_SENTINEL = -128, -2, -1


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("total", "unique", "distinct", "missing", "edges")]


class QV(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("missing", "kcount", "edge_missing")]


def build(force=False):
    """Compile the C restatement (gcc); building the checker is not using it."""
    src = os.path.join(_HERE, "kreeq_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, u64, sz = C.c_void_p, C.c_uint64, C.c_size_t
        L.ko_table_new.restype = vp; L.ko_table_new.argtypes = [C.c_int]
        L.ko_table_free.argtypes = [vp]
        L.ko_table_k.restype = C.c_int; L.ko_table_k.argtypes = [vp]
        L.ko_hash.restype = u64; L.ko_hash.argtypes = [vp, C.c_int, C.POINTER(C.c_int)]
        L.ko_extract.restype = sz; L.ko_extract.argtypes = [C.c_int, C.c_char_p, sz, vp, vp]
        L.ko_insert_tuple.argtypes = [vp, u64, C.c_uint8]
        L.ko_count_batch.restype = sz; L.ko_count_batch.argtypes = [vp, C.c_char_p, sz]
        L.ko_put8.argtypes = [vp, u64, vp]; L.ko_put32.argtypes = [vp, u64, vp]
        L.ko_finalize.argtypes = [vp]
        L.ko_size8.restype = sz; L.ko_size8.argtypes = [vp]
        L.ko_size32.restype = sz; L.ko_size32.argtypes = [vp]
        L.ko_dump8.restype = sz; L.ko_dump8.argtypes = [vp, vp, vp, sz]
        L.ko_dump32.restype = sz; L.ko_dump32.argtypes = [vp, vp, vp, sz]
        L.ko_lookup.restype = C.c_int; L.ko_lookup.argtypes = [vp, u64, vp]
        L.ko_export.restype = sz; L.ko_export.argtypes = [vp, vp, vp, sz]
        L.ko_stats_get.argtypes = [vp, C.POINTER(Stats)]
        L.ko_hist.restype = sz; L.ko_hist.argtypes = [vp, vp, vp, sz]
        L.ko_union.restype = vp; L.ko_union.argtypes = [C.POINTER(vp), C.c_int]
        L.ko_validate_segment.restype = C.c_int
        L.ko_validate_segment.argtypes = [vp, C.c_char_p, u64, C.c_uint32, vp, C.POINTER(QV)]
        L.ko_error_rate.restype = C.c_double; L.ko_error_rate.argtypes = [u64, u64, C.c_int]
        L.ko_qv_value.restype = C.c_double; L.ko_qv_value.argtypes = [u64, u64, C.c_int]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def kmer_hash(codes, k):
    """Kmap::hash (kmer.h:781-794) of k codes in 0..3 -> (key, isFw)."""
    a = np.ascontiguousarray(codes, dtype=np.uint8)
    fw = C.c_int(0)
    key = lib().ko_hash(_ptr(a), k, C.byref(fw))
    return int(key), bool(fw.value)


def extract(k, ascii_bytes):
    """(keys u64[], edges u8[]) of one 'N'-joined batch, in position order."""
    b = bytes(ascii_bytes)
    n = lib().ko_extract(k, b, len(b), None, None)
    keys = np.empty(n, dtype=np.uint64); edges = np.empty(n, dtype=np.uint8)
    if n:
        lib().ko_extract(k, b, len(b), _ptr(keys), _ptr(edges))
    return keys, edges


class Table:
    """One DBG database in memory: the 128 8-bit maps + the 32-bit maps."""

    def __init__(self, k, handle=None):
        self.k = k
        self.h = handle if handle is not None else lib().ko_table_new(k)

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.ko_table_free(self.h); self.h = None

    def count_batch(self, ascii_bytes):
        b = bytes(ascii_bytes)
        return lib().ko_count_batch(self.h, b, len(b))

    def insert_tuples(self, keys, edges):
        L = lib()
        for key, e in zip(np.asarray(keys, dtype=np.uint64).tolist(), np.asarray(edges, dtype=np.uint8).tolist()):
            L.ko_insert_tuple(self.h, key, e)

    def finalize(self):
        lib().ko_finalize(self.h)

    def put8(self, keys, recs):
        recs = np.ascontiguousarray(recs, dtype=REC8)
        for i, key in enumerate(np.asarray(keys, dtype=np.uint64).tolist()):
            lib().ko_put8(self.h, key, C.c_void_p(recs.ctypes.data + 9 * i))

    def put32(self, keys, recs):
        recs = np.ascontiguousarray(recs, dtype=REC32)
        for i, key in enumerate(np.asarray(keys, dtype=np.uint64).tolist()):
            lib().ko_put32(self.h, key, C.c_void_p(recs.ctypes.data + 36 * i))

    def dump8(self):
        n = lib().ko_size8(self.h)
        keys = np.empty(n, dtype=np.uint64); recs = np.empty(n, dtype=REC8)
        lib().ko_dump8(self.h, _ptr(keys), _ptr(recs), n)
        return keys, recs

    def dump32(self):
        n = lib().ko_size32(self.h)
        keys = np.empty(n, dtype=np.uint64); recs = np.empty(n, dtype=REC32)
        lib().ko_dump32(self.h, _ptr(keys), _ptr(recs), n)
        return keys, recs

    def export(self):
        """Logical {key -> 32-bit record} set, sorted by key."""
        cap = lib().ko_size8(self.h) + lib().ko_size32(self.h)
        keys = np.empty(cap, dtype=np.uint64); recs = np.empty(cap, dtype=REC32)
        n = lib().ko_export(self.h, _ptr(keys), _ptr(recs), cap)
        return keys[:n].copy(), recs[:n].copy()

    def lookup(self, key):
        r = np.zeros(1, dtype=REC32)
        rc = lib().ko_lookup(self.h, int(key), _ptr(r))
        return rc, r[0]

    def stats(self):
        s = Stats(); lib().ko_stats_get(self.h, C.byref(s))
        return {n: getattr(s, n) for n, _ in Stats._fields_}

    def hist(self):
        cap = 1 << 16
        c = np.empty(cap, dtype=np.uint64); n = np.empty(cap, dtype=np.uint64)
        m = lib().ko_hist(self.h, _ptr(c), _ptr(n), cap)
        return c[:m].copy(), n[:m].copy()

    def validate_segment(self, seq, cutoff=0, acc=None):
        b = bytes(seq)
        out = np.zeros(len(b), dtype=BASE)
        q = QV() if acc is None else QV(*acc)
        rc = lib().ko_validate_segment(self.h, b, len(b), cutoff, _ptr(out), C.byref(q))
        return rc, out, (q.missing, q.kcount, q.edge_missing)


def union(tables):
    arr = (C.c_void_p * len(tables))(*[t.h for t in tables])
    return Table(tables[0].k, handle=lib().ko_union(arr, len(tables)))


def stats_text(st):
    """DBG::DBstats stdout block, src/graph-builder.cpp:288-293."""
    return ("DBG Summary statistics:\nTotal kmers: %d\nUnique kmers: %d\nDistinct kmers: %d\n"
            "Missing kmers: %d\nTotal edges: %d\n" % (st["total"], st["unique"], st["distinct"], st["missing"], st["edges"]))


def _g6(x):
    """operator<<(double) with default precision 6 (== printf %g)."""
    return "%g" % x


def qv_text(missing, kcount, edge_missing, k):
    """The 3-line QV table, src/kreeq.cpp:80-104."""
    L = lib()
    out = "Missing\tTotal\tQV\tError\tk\tMethod\n"
    for miss, name in ((missing, "Merqury"), (missing + edge_missing, "Kreeq")):
        if kcount == 0:
            err = float("nan"); qv = float("nan")
        else:
            err = L.ko_error_rate(miss, kcount, k); qv = L.ko_qv_value(miss, kcount, k)
        out += "%d\t%d\t%s\t%s\t%d\t%s\n" % (miss, kcount, _g6(qv), _g6(err), k, name)
    return out


# ---- .kreeq directory decoder -----------------------------------

_M64 = (1 << 64) - 1


def phmap_hash(key):
    """phmap_mix<8>(std::hash<uint64_t>(key)): hi64 + lo64 of key * 0xde5fb9d2630458e9
    (phmap_utils.h:66-77; std::hash<uint64_t> is the identity)."""
    p = int(key) * 0xde5fb9d2630458e9
    return ((p >> 64) + (p & _M64)) & _M64


def submap_index(h):
    """parallel_hash_set::subidx, phmap.h:3840-3842 with N=8."""
    return ((h >> 8) ^ (h >> 16) ^ (h >> 24)) & 255


def read_map_file(path, slot_dtype, check=True):
    """Decode one phmap dump -> (keys u64[], slots structured[]) of the full slots.

    With check=True every rule a loader/`find` of the reference relies on is
    verified: version word, capacity 2^n-1, ctrl byte == H2 for full slots,
    sentinel, mirrored ctrl bytes, no kDeleted, growth_left, submap index, and
    that a phmap `find` (probe_seq over 16-byte groups, stop at the first group
    with an empty byte) reaches every key."""
    raw = np.fromfile(path, dtype=np.uint8)
    buf = raw.tobytes()
    off = 0

    def u64():
        nonlocal off
        v = int.from_bytes(buf[off:off + 8], "little"); off += 8
        return v

    nsub = u64()
    if nsub != SUBMAPS:
        raise ValueError("%s: submap_count %d != 256" % (path, nsub))
    keys_all, slots_all = [], []
    for sub in range(SUBMAPS):
        ver, size, cap = u64(), u64(), u64()
        if ver != VERSION:
            raise ValueError("%s: submap %d bad version %x" % (path, sub, ver))
        if size == 0:
            continue
        if check and (cap & (cap + 1)) != 0:
            raise ValueError("%s: submap %d capacity %d is not 2^n-1" % (path, sub, cap))
        nctrl = cap + GROUP_WIDTH + 1
        ctrl = np.frombuffer(buf, dtype=np.int8, count=nctrl, offset=off); off += nctrl
        slots = np.frombuffer(buf, dtype=slot_dtype, count=cap, offset=off); off += cap * slot_dtype.itemsize
        growth_left = u64()
        full = ctrl[:cap] >= 0
        s = slots[full]
        if check:
            if int(full.sum()) != size:
                raise ValueError("%s: submap %d size %d but %d full ctrl bytes" % (path, sub, size, int(full.sum())))
            if ctrl[cap] != K_SENTINEL:
                raise ValueError("%s: submap %d missing sentinel" % (path, sub))
            if np.any(ctrl[:cap] == K_DELETED):
                raise ValueError("%s: submap %d contains kDeleted" % (path, sub))
            nmir = min(GROUP_WIDTH - 1, cap)   # set_ctrl mirrors i < kWidth-1 (phmap.h:2264-2276)
            if not np.array_equal(ctrl[cap + 1:cap + 1 + nmir], ctrl[:nmir]):
                raise ValueError("%s: submap %d ctrl mirror mismatch" % (path, sub))
            if growth_left != (cap - cap // 8) - size and not (cap == 7 and growth_left == 6 - size):
                raise ValueError("%s: submap %d growth_left %d (size %d cap %d)" % (path, sub, growth_left, size, cap))
            idx = np.nonzero(full)[0]
            for pos, key in zip(idx.tolist(), s["key"].tolist()):
                h = phmap_hash(key)
                if submap_index(h) != sub:
                    raise ValueError("%s: key %d in submap %d, expected %d" % (path, key, sub, submap_index(h)))
                if int(ctrl[pos]) != (h & 0x7F):
                    raise ValueError("%s: key %d ctrl %d != H2 %d" % (path, key, int(ctrl[pos]), h & 0x7F))
                # simulate find(): phmap.h:2199-2213 with probe_seq phmap.h:144-168
                o = (h >> 7) & cap; step = 0; ok = False
                while True:
                    g = [(o + i) for i in range(GROUP_WIDTH)]   # unaligned group read over ctrl[o..o+15]
                    gv = ctrl[g[0]:g[0] + GROUP_WIDTH]
                    hit = [g[i] & cap for i in range(GROUP_WIDTH) if int(gv[i]) == (h & 0x7F)]
                    if any(int(slots["key"][q]) == key for q in hit if q < cap):
                        ok = True; break
                    if np.any(gv == K_EMPTY):
                        break
                    step += GROUP_WIDTH; o = (o + step) & cap
                    if step > cap + GROUP_WIDTH:
                        break
                if not ok:
                    raise ValueError("%s: key %d unreachable by phmap find" % (path, key))
        keys_all.append(s["key"].copy()); slots_all.append(s.copy())
    if check and off != len(buf):
        raise ValueError("%s: %d trailing bytes" % (path, len(buf) - off))
    if not keys_all:
        return np.empty(0, dtype=np.uint64), np.empty(0, dtype=slot_dtype)
    return np.concatenate(keys_all), np.concatenate(slots_all)


def read_index(db_dir):
    with open(os.path.join(db_dir, ".index")) as f:
        return int(f.readline().strip())


def read_db(db_dir, check=True):
    """Decode a .kreeq directory into an oracle Table (raw 8-bit + 32-bit maps)."""
    k = read_index(db_dir)
    t = Table(k)
    for m in range(MAP_COUNT):
        keys, slots = read_map_file(os.path.join(db_dir, ".map.%d.bin" % m), SLOT8, check)
        if check and len(keys) and np.any(keys % np.uint64(MAP_COUNT) != m):
            raise ValueError("%s: map %d holds keys of another map" % (db_dir, m))
        rec = np.zeros(len(keys), dtype=REC8)
        rec["fw"], rec["bw"], rec["cov"] = slots["fw"], slots["bw"], slots["cov"]
        t.put8(keys, rec)
    keys, slots = read_map_file(os.path.join(db_dir, ".map.hc.bin"), SLOT32, check)
    rec = np.zeros(len(keys), dtype=REC32)
    rec["fw"], rec["bw"], rec["cov"] = slots["fw"], slots["bw"], slots["cov"]
    t.put32(keys, rec)
    return t


# ---- FASTQ batching (getKmers) -------------------------------------------------

def fastq_batches(data, batch_size=10000000):
    """getKmers + loadKmers (gfalibs/include/functions.h:69-122, fastx.h:83-121):
    4-line FASTQ -> list of byte strings of sequence lines each followed by 'N'.
    A file that does not start with '@' yields nothing (fastx.h:98-118)."""
    if not data or data[0:1] != b"@":
        return []
    lines = data.split(b"\n")
    if lines and lines[-1] == b"":
        trailing_nl = True; lines.pop()
    else:
        trailing_nl = False
    out, cur, budget = [], bytearray(), batch_size
    i = 0
    while i < len(lines):
        # header line
        if i + 1 >= len(lines):
            break
        seq = lines[i + 1]
        cur += seq
        budget -= len(seq) + 1
        have_plus = i + 2 < len(lines)
        have_qual = i + 3 < len(lines)
        # 'N' is appended only after both the '+' and quality lines ended with '\n'
        complete = have_plus and have_qual and (i + 3 < len(lines) - 1 or trailing_nl)
        if complete:
            cur += b"N"
        i += 4
        if budget <= 0:
            out.append(bytes(cur)); cur = bytearray(); budget = batch_size
    if cur or not out:
        out.append(bytes(cur))
    return out

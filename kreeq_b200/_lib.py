"""ctypes binding of libkreeq_b200.so (the C-ABI declared in include/kreeq_b200.h).

The library is built in-tree by `__graft_entry__.build()` / `make -C kreeq_b200/csrc`.  There is no
fallback of any kind: if the shared object is missing, loading raises, and every compute entry point
fails with KQ_ENODEV when no CUDA device is usable.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkreeq_b200.so")

KQ_OK, KQ_EINVAL, KQ_ENODEV, KQ_ECUDA, KQ_ENOMEM, KQ_EIO, KQ_ECORRUPT, KQ_ESTATE = 0, -1, -2, -3, -4, -5, -6, -7
ERR_NAMES = {-1: "KQ_EINVAL", -2: "KQ_ENODEV", -3: "KQ_ECUDA", -4: "KQ_ENOMEM", -5: "KQ_EIO", -6: "KQ_ECORRUPT", -7: "KQ_ESTATE"}

REC32 = np.dtype([("fw", "<u4", 4), ("bw", "<u4", 4), ("cov", "<u4")])                       # kq_rec32 / DBGkmer32
BASE = np.dtype([("fw", "<u4"), ("bw", "<u4"), ("cov", "<u4"), ("isFw", "u1"), ("pad", "u1", 3)])  # kq_base / DBGbase
BASE4 = np.dtype([("cov", "u1"), ("fw", "u1"), ("bw", "u1"), ("flags", "u1")])                           # kq_base4
WIDE = np.dtype([("pos", "<u8"), ("row", BASE)])                                                        # kq_wide_row
assert REC32.itemsize == 36 and BASE.itemsize == 16 and BASE4.itemsize == 4 and WIDE.itemsize == 24


class KqStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("total", "unique", "distinct", "missing", "edges")]


class KqQV(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("missing", "kcount", "edge_missing")]


class KreeqError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s: %s" % (ERR_NAMES.get(code, str(code)), msg))
        self.code = code


# every symbol include/kreeq_b200.h declares: name -> (restype, argtypes)
_vp, _u64, _sz, _i, _cp = C.c_void_p, C.c_uint64, C.c_size_t, C.c_int, C.c_char_p
PROTOTYPES = {
    "kq_create": (_i, [C.POINTER(_vp), _i, _i, _u64]),
    "kq_destroy": (None, [_vp]),
    "kq_last_error": (_cp, [_vp]),
    "kq_k": (_i, [_vp]),
    "kq_set_stream": (_i, [_vp, _vp]),
    "kq_sync": (_i, [_vp]),
    "kq_set_shard": (_i, [_vp, _i, _i]),
    "kq_owner_of": (_i, [_u64, _i]),
    "kq_count_push": (_i, [_vp, _vp, _sz]),
    "kq_count_finish": (_i, [_vp]),
    "kq_count_instances": (_i, [_vp, C.POINTER(_u64)]),
    "kq_clear": (_i, [_vp]),
    "kq_db_save": (_i, [_vp, _cp]),
    "kq_db_save_hc_part": (_i, [_vp, _cp]),
    "kq_db_merge_hc_parts": (_i, [_cp, _i]),
    "kq_db_load": (_i, [_vp, _cp]),
    "kq_db_read_k": (_i, [_cp]),
    "kq_db_write_records": (_i, [_cp, _i, _vp, _vp, _sz]),
    "kq_db_read_records": (_i, [_cp, _vp, _vp, C.POINTER(_sz)]),
    "kq_union": (_i, [_vp, C.POINTER(_cp), _i, _i]),
    "kq_stats_get": (_i, [_vp, C.POINTER(KqStats)]),
    "kq_hist": (_i, [_vp, _vp, _vp, C.POINTER(_sz)]),
    "kq_validate_segment": (_i, [_vp, _vp, _u64, C.c_uint32, _vp, C.POINTER(KqQV)]),
    "kq_error_rate": (C.c_double, [_u64, _u64, _i]),
    "kq_qv_value": (C.c_double, [_u64, _u64, _i]),
    "kq_validate_segment_compact": (_i, [_vp, _vp, _u64, C.c_uint32, _vp, _vp, C.POINTER(_sz), C.POINTER(KqQV)]),
    "kq_expand_rows": (_i, [_vp, _u64, _vp, _sz, _vp]),
    "kq_lookup_batch": (_i, [_vp, _vp, _sz, _vp, _vp]),
    "kq_lookup_batch_static": (_i, [_vp, _vp, _sz, _vp, _vp]),
    "kq_lookup_static_max": (_sz, []),
    "kq_export": (_i, [_vp, _vp, _vp, C.POINTER(_sz)]),
    "kq_distinct": (_i, [_vp, C.POINTER(_u64)]),
    "kq_packed_words": (_sz, [_sz]),
    "kq_pack_host": (_i, [_vp, _sz, _vp, _vp, _i]),
    "kq_dev_count_packed": (_i, [_vp, _vp, _vp, _u64, _u64, _u64]),
    "kq_tuple_bytes": (_i, [_vp]),
    "kq_dev_extract_route": (_i, [_vp, _vp, _vp, _u64, _u64, _u64, _vp, _u64, _vp]),
    "kq_dev_insert_tuples": (_i, [_vp, _vp, _u64]),
    "kq_set_flush_positions": (_i, [_vp, _u64]),
    "kq_set_tuning": (_i, [_cp, C.c_longlong]),
    "kq_route_setup": (_i, [_vp, _u64, _i, C.POINTER(_vp), _vp]),
    "kq_route_open": (_i, [_vp, _vp, C.POINTER(_vp)]),
    "kq_route_connect": (_i, [_vp, C.POINTER(_vp)]),
    "kq_route_bins": (_i, [_vp]),
    "kq_route_hist": (_i, [_vp, _vp, _vp, _u64, _u64, _u64, _vp]),
    "kq_route_scatter": (_i, [_vp, _vp, _vp, _u64, _u64, _u64, _vp, C.POINTER(_u64)]),
    "kq_route_insert": (_i, [_vp, _u64]),
    "kq_peer_export": (_i, [_vp, _vp]),
    "kq_peer_connect": (_i, [_vp, _vp]),
    "kq_peer_connect_local": (_i, [_vp, C.POINTER(_vp)]),
    "kq_peer_disconnect": (_i, [_vp]),
    "kq_dev_validate_packed": (_i, [_vp, _vp, _vp, _u64, C.c_uint32, _vp, _vp]),
    "kq_vroute_setup": (_i, [_vp, _u64, C.POINTER(_vp), _vp, C.POINTER(_vp), _vp]),
    "kq_vroute_connect": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp)]),
    "kq_vroute_send": (_i, [_vp, _vp, _vp, _u64, _u64, _u64, _u64]),
    "kq_vroute_send_at": (_i, [_vp, _vp, _vp, _u64, _u64, _u64, _u64, _u64, _i]),
    "kq_vroute_answer": (_i, [_vp, C.c_uint32]),
    "kq_vroute_place": (_i, [_vp]),
    "kq_vroute_rows": (_i, [_vp, C.POINTER(_vp)]),
    "kq_vroute_copy_rows": (_i, [_vp, _vp, _u64]),
    "kq_vroute_counters": (_i, [_vp, C.POINTER(KqQV), _i]),
    "kq_multi_create": (_i, [C.POINTER(_vp), _i, _i, C.POINTER(_i), _u64]),
    "kq_multi_destroy": (None, [_vp]),
    "kq_multi_last_error": (_cp, [_vp]),
    "kq_multi_ngpus": (_i, [_vp]),
    "kq_multi_shard": (_vp, [_vp, _i]),
    "kq_multi_set_round": (_i, [_vp, _u64, _u64]),
    "kq_multi_count_push": (_i, [_vp, _vp, _sz]),
    "kq_multi_count_finish": (_i, [_vp]),
    "kq_multi_count_instances": (_i, [_vp, C.POINTER(_u64)]),
    "kq_multi_clear": (_i, [_vp]),
    "kq_multi_stats_get": (_i, [_vp, C.POINTER(KqStats)]),
    "kq_multi_hist": (_i, [_vp, _vp, _vp, C.POINTER(_sz)]),
    "kq_multi_export": (_i, [_vp, _vp, _vp, C.POINTER(_sz)]),
    "kq_multi_lookup_batch": (_i, [_vp, _vp, _sz, _vp, _vp]),
    "kq_multi_db_save": (_i, [_vp, _cp]),
    "kq_multi_db_load": (_i, [_vp, _cp]),
    "kq_multi_union": (_i, [_vp, C.POINTER(_cp), _i, _i]),
    "kq_multi_validate_segment": (_i, [_vp, _vp, _u64, C.c_uint32, _vp, C.POINTER(KqQV)]),
    "kq_reserve": (_i, [_vp, _u64]),
    "kq_counters": (_i, [_vp, C.POINTER(_u64), C.POINTER(_u64), C.POINTER(_u64)]),
    "kq_set_timing": (_i, [_vp, _i]),
    "kq_kernel_time_ms": (_i, [_vp, _i, C.POINTER(C.c_double), C.POINTER(_u64), _i]),
}

_lib = None


def lib():
    """Load libkreeq_b200.so; raises if the extension has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(or make -C kreeq_b200/csrc); kreeq_b200 has no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def ptr(a):
    """void* of a numpy array (or None)."""
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def check(rc, ctx=None):
    if rc != KQ_OK:
        msg = lib().kq_last_error(ctx)
        raise KreeqError(rc, msg.decode() if msg else "")
    return rc

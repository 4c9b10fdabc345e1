"""kreeq_b200 -- B200 (sm_100a) implementation of kreeq's k-mer hot path behind a C-ABI.

`DBG` mirrors the reference's DBG/Kmap members for count / union / validate; everything it does
runs in libkreeq_b200.so (hand-written CUDA).  There is no CPU fallback.
"""
from ._lib import BASE, REC32, KreeqError, LIB_PATH, lib  # noqa: F401
from .dbg import DBG, MultiDBG, db_read_records, db_write_records, pack_host  # noqa: F401

__all__ = ["DBG", "MultiDBG", "KreeqError", "REC32", "BASE", "LIB_PATH", "lib", "pack_host", "db_write_records", "db_read_records"]

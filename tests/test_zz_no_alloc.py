"""kq_lookup_batch_static: the lookup entry point that allocates nothing (its staging is one file-scope device
array).  Collected last on purpose (file name): it checks the entry point against kq_lookup_batch through the
C-ABI, on two calls in a row that share the array."""
import ctypes as C

import numpy as np
import pytest

from helpers import random_genome, sample_reads

pytestmark = pytest.mark.gpu


def _ptr(a):
    return C.c_void_p(a.ctypes.data)


def _both(L, h, keys):
    """-> ((rec, found) of kq_lookup_batch, (rec, found) of kq_lookup_batch_static); rec pre-filled with a pattern
    so that 'found = 0 leaves rec untouched' is checked too"""
    from kreeq_b200._lib import REC32, check
    out = []
    for fn in (L.kq_lookup_batch, L.kq_lookup_batch_static):
        rec = np.zeros(keys.size, dtype=REC32)
        rec["cov"] = 0xABCD0000 + (np.arange(keys.size, dtype=np.uint32) & 0xFFFF)
        found = np.full(keys.size, 7, dtype=np.uint8)
        check(fn(h, _ptr(keys), keys.size, _ptr(rec), _ptr(found)), h)
        out.append((rec, found))
    return out


def test_static_lookup_matches_lookup_batch_twice():
    import kreeq_b200 as kb
    from kreeq_b200._lib import KQ_EINVAL, KQ_ESTATE, REC32, lib
    L = lib()
    nmax = L.kq_lookup_static_max()
    assert nmax == 65536
    rng = np.random.default_rng(4242)
    genome = random_genome(rng, 60_000)
    reads = sample_reads(rng, genome, 4000, 150, err=0.01)
    with kb.DBG(k=21, expected_distinct=400_000) as g:
        g.traverseInReads(reads)
        g.finalize()
        keys_in, _ = g.export()
        assert keys_in.size > 50_000
        # call 1: a full array of present and absent keys; call 2: fewer, different keys (the array still holds
        # call 1's keys, records and flags behind them)
        absent = rng.integers(0, 1 << 42, 20_000, dtype=np.uint64)
        batch1 = np.concatenate([keys_in[: nmax - absent.size], absent]).astype(np.uint64)
        rng.shuffle(batch1)
        batch2 = np.concatenate([rng.integers(0, 1 << 42, 3_000, dtype=np.uint64), keys_in[-5_000:]]).astype(np.uint64)
        for batch in (batch1, batch2):
            (r0, f0), (r1, f1) = _both(L, g.handle, np.ascontiguousarray(batch))
            assert f0.tobytes() == f1.tobytes()
            assert r0.tobytes() == r1.tobytes()          # bit-exact, untouched entries included
            assert 0 < int(f1.sum()) < batch.size and set(np.unique(f1)) <= {0, 1}
        # sizes beyond the array are refused with a clear message, nothing is run
        big = np.zeros(nmax + 1, dtype=np.uint64)
        rec = np.zeros(big.size, dtype=REC32); found = np.zeros(big.size, dtype=np.uint8)
        assert L.kq_lookup_batch_static(g.handle, _ptr(big), big.size, _ptr(rec), _ptr(found)) == KQ_EINVAL
        assert b"KQ_STATIC_LOOKUP_MAX" in L.kq_last_error(g.handle)
        # pending reads are the caller's kq_sync: this call must not grow the table
        g.traverseInReads(reads[:50_000])
        rc = L.kq_lookup_batch_static(g.handle, _ptr(batch2), batch2.size, _ptr(rec), _ptr(found))
        assert rc in (0, KQ_ESTATE)                      # (a push may already have been flushed)
        g.finalize()
        (r0, f0), (r1, f1) = _both(L, g.handle, np.ascontiguousarray(batch2))
        assert f0.tobytes() == f1.tobytes() and r0.tobytes() == r1.tobytes()

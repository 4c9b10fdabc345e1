"""Seeded parity problems whose answers are committed as constants (tests/golden/bench_parity.json, made by
tools/make_bench_parity.py from the oracle): bench.py and smoke() count + validate them on the GPU(s) before
anything is timed, so a driver-run line shows that the pipeline it measures counts correctly at that GPU count.

The generator is numpy on the CPU (identical on every machine).  A digest of a database is
(sum over records of key * cov mod 2^64, sum of (fw + bw edge counts) mod 2^64, distinct, total); a digest of a
validate pass is the three QV counters plus sum over positions of (cov + 3 fw + 5 bw + 7 isFw) * (position + 1)
mod 2^64."""
import json
import os

import numpy as np

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "bench_parity.json")

PROBLEMS = {
    # name: genome bp, reads, read length, error rate, seed, k
    "bench": (2_000_000, 400_000, 150, 0.01, 20250, 21),      # 5.2e7 instances, ~1.2e7 distinct: a multi-slice table
    "smoke": (500_000, 100_000, 150, 0.01, 20251, 21),        # 1.3e7 instances, ~2.9e6 distinct
}


def problem(name):
    """-> (reads: uint8 'N'-joined batch as getKmers() builds it, genome: uint8 ASCII, k)"""
    gbp, n_reads, L, err, seed, k = PROBLEMS[name]
    rng = np.random.default_rng(seed)
    genome = rng.integers(0, 4, gbp, dtype=np.uint8)
    start = rng.integers(0, gbp - L, n_reads)
    r = genome[start[:, None] + np.arange(L)[None, :]]
    flip = rng.random(n_reads) < 0.5
    r[flip] = 3 - r[flip][:, ::-1]
    e = rng.random(r.shape) < err
    r = np.where(e, (r + rng.integers(1, 4, r.shape, dtype=np.uint8)) % 4, r).astype(np.uint8)
    out = np.empty((n_reads, L + 1), dtype=np.uint8)
    out[:, :L] = _ACGT[r]
    out[:, L] = ord("N")
    return out.reshape(-1), _ACGT[genome], k


def db_digest(keys, recs):
    """digest of a {key -> record} set (any order, any shard): four python ints, additive over shards"""
    keys = np.asarray(keys, dtype=np.uint64)
    cov = recs["cov"].astype(np.uint64)
    with np.errstate(over="ignore"):
        kc = int((keys * cov).sum(dtype=np.uint64))
        ed = int((recs["fw"].astype(np.uint64).sum(axis=1) + recs["bw"].astype(np.uint64).sum(axis=1)).sum(dtype=np.uint64)) if keys.size else 0
    return [kc, ed, int(keys.size), int(cov.sum(dtype=np.uint64))]


def rows_digest(rows, first_pos=0):
    """digest of the DBGbase rows of positions [first_pos, first_pos + len(rows)) of the assembly"""
    pos = np.arange(first_pos + 1, first_pos + 1 + rows.size, dtype=np.uint64)
    with np.errstate(over="ignore"):
        v = rows["cov"].astype(np.uint64) + np.uint64(3) * rows["fw"].astype(np.uint64) + np.uint64(5) * rows["bw"].astype(np.uint64) \
            + np.uint64(7) * rows["isFw"].astype(np.uint64)
        return int((v * pos).sum(dtype=np.uint64))


def add_mod64(a, b):
    return (a + b) & 0xFFFFFFFFFFFFFFFF


def expected(name):
    with open(GOLDEN) as f:
        return json.load(f)[name]

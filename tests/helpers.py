"""Seeded synthetic inputs shared by the parity tests (SURVEY 8d generator, scaled down)."""
import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


def random_genome(rng, n):
    return rng.choice(ACGT, size=n)


def sample_reads(rng, genome, n_reads, read_len, err=0.01, n_rate=0.0, both_strands=True):
    """'N'-joined read batch as getKmers() would build it (gfalibs/include/functions.h:69-122)."""
    comp = np.zeros(256, dtype=np.uint8)
    for a, b in zip(b"ACGT", b"TGCA"):
        comp[a] = b
    out = []
    for _ in range(n_reads):
        s = int(rng.integers(0, max(1, genome.size - read_len)))
        r = genome[s:s + read_len].copy()
        if both_strands and rng.random() < 0.5:
            r = comp[r][::-1].copy()
        e = rng.random(r.size) < err
        if e.any():
            r[e] = (np.searchsorted(ACGT, r[e]) + rng.integers(1, 4, size=int(e.sum()))) % 4
            r[e] = ACGT[r[e]]
        if n_rate:
            r[rng.random(r.size) < n_rate] = ord("N")
        out.append(r.tobytes())
        out.append(b"N")
    return b"".join(out)


def assert_same_records(a, b):
    (ka, ra), (kb, rb) = a, b
    assert ka.size == kb.size, (ka.size, kb.size)
    assert np.array_equal(ka, kb)
    assert ra.tobytes() == rb.tobytes()

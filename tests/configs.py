"""Scaled-down twins of BASELINE.json configs 3, 4 and 5 (SURVEY 8d definitions) as files, generated with numpy:
planted repeats + poly-A (C3), 24 records with N-runs (C4), long reads with k = 31 (C5)."""
import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
COMP = np.zeros(256, dtype=np.uint8)
for _a, _b in zip(b"ACGTN", b"TGCAN"):
    COMP[_a] = _b
# relative sizes of the 24 human chromosomes (1..22, X, Y), Mbp
HUMAN = [248, 242, 198, 190, 182, 171, 159, 145, 138, 134, 135, 133, 114, 107, 102, 90, 83, 80, 59, 64, 47, 51, 156, 57]


def genome(rng, n):
    return ACGT[rng.integers(0, 4, n)]


def plant_repeats(rng, g, copies, rep_len, polya_len):
    """`copies` copies of one random repeat and one poly-A stretch, at random non-overlapping places (C3)"""
    rep = genome(rng, rep_len)
    slots = rng.choice(g.size // (rep_len + polya_len) - 1, copies + 1, replace=False) * (rep_len + polya_len)
    for s in slots[:copies]:
        g[s:s + rep_len] = rep
    g[slots[copies]:slots[copies] + polya_len] = ord("A")
    return g


def reads_fixed(rng, g, n_reads, L, err):
    start = rng.integers(0, g.size - L, n_reads)
    r = g[start[:, None] + np.arange(L)[None, :]]
    flip = rng.random(n_reads) < 0.5
    r[flip] = COMP[r[flip]][:, ::-1]
    e = rng.random(r.shape) < err
    code = np.searchsorted(ACGT, r)
    r = np.where(e, ACGT[(code + rng.integers(1, 4, r.shape)) % 4], r).astype(np.uint8)
    return [x.tobytes() for x in r]


def reads_long(rng, g, n_reads, mean, sd, lo, hi, err):
    out = []
    for _ in range(n_reads):
        L = int(np.clip(rng.normal(mean, sd), lo, min(hi, g.size - 1)))
        s = int(rng.integers(0, g.size - L))
        r = g[s:s + L].copy()
        if rng.random() < 0.5:
            r = COMP[r][::-1].copy()
        e = rng.random(L) < err
        if e.any():
            code = np.searchsorted(ACGT, r[e])
            r[e] = ACGT[(code + rng.integers(1, 4, int(e.sum()))) % 4]
        out.append(r.tobytes())
    return out


def write_fastq(path, reads):
    with open(path, "wb") as f:
        for i, r in enumerate(reads):
            f.write(b"@r%d\n" % i + r + b"\n+\n" + b"I" * len(r) + b"\n")


def write_fasta(path, records, width=0):
    with open(path, "wb") as f:
        for name, seq in records:
            f.write(b">" + name.encode() + b"\n")
            if width:
                for o in range(0, len(seq), width):
                    f.write(seq[o:o + width] + b"\n")
            else:
                f.write(seq + b"\n")


def records_with_gaps(rng, g, n_records, n_gaps, gap_len):
    """cut g into n_records pieces with sizes proportional to the human chromosomes and put n_gaps runs of N inside (C4)"""
    w = np.array(HUMAN[:n_records], dtype=np.float64)
    sizes = np.maximum(1000, (w / w.sum() * g.size).astype(np.int64))
    sizes[-1] = g.size - sizes[:-1].sum()
    recs, o = [], 0
    gaps_in = rng.multinomial(n_gaps, w / w.sum())
    for i, sz in enumerate(sizes):
        seq = bytearray(g[o:o + sz].tobytes()); o += sz
        for p in sorted(rng.integers(100, max(101, sz - 100), gaps_in[i]), reverse=True):
            seq[p:p] = b"N" * gap_len
        recs.append(("chr%d" % (i + 1), bytes(seq)))
    return recs

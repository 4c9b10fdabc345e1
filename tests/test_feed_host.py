"""Host feed path of the command line, no GPU: the parallel gzip decoder
(kreeq_b200/csrc/cli/pgzip.cpp) against zlib on every deflate feature a .fastq.gz / .fa.gz can contain, and the
FASTQ batching that sits on top of it (gfalibs/include/functions.h:69-122 end-of-file rules) against the
zlib path and a direct Python restatement.  Driven through kreeq_b200/kreeq-feedcheck, which links the
command line's own input layer."""
import gzip
import json
import os
import subprocess
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOL = os.path.join(ROOT, "kreeq_b200", "kreeq-feedcheck")
GOLD = os.path.join(ROOT, "tests", "golden")


def inflate(path, threads=4, chunk=0, expect_ok=True):
    p = subprocess.run([TOOL, "inflate", str(path), str(threads), str(chunk)], capture_output=True, timeout=300)
    info = json.loads(p.stderr.decode().strip().splitlines()[-1])
    assert info["ok"] == (p.returncode == 0)
    if expect_ok:
        assert p.returncode == 0, info
    return p.stdout, info


def fastq_text(rng, n, read_len=150, genome=None):
    g = genome if genome is not None else rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=200_000)
    out = []
    for i in range(n):
        L = read_len if isinstance(read_len, int) else int(rng.integers(*read_len))
        s = int(rng.integers(0, g.size - L))
        r = g[s:s + L].copy()
        e = rng.random(L) < 0.01
        r[e] = rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), size=int(e.sum()))
        q = rng.integers(33 + 2, 33 + 41, size=L).astype(np.uint8)
        out.append(b"@r%d len=%d\n" % (i, L) + r.tobytes() + b"\n+\n" + q.tobytes() + b"\n")
    return b"".join(out)


def gz(data, level=6, strategy=zlib.Z_DEFAULT_STRATEGY, flush_every=0, flush_mode=zlib.Z_FULL_FLUSH):
    c = zlib.compressobj(level, zlib.DEFLATED, 31, 8, strategy)
    if not flush_every:
        return c.compress(data) + c.flush()
    out = []
    for a in range(0, len(data), flush_every):
        out.append(c.compress(data[a:a + flush_every]))
        out.append(c.flush(flush_mode))
    return b"".join(out) + c.flush()


@pytest.fixture(scope="module")
def text():
    return fastq_text(np.random.default_rng(11), 20000)          # ~6.3 MB


CASES = {
    "level1": dict(level=1), "level6": dict(level=6), "level9": dict(level=9),
    "stored": dict(level=0),
    "fixed_huffman": dict(level=6, strategy=zlib.Z_FIXED),
    "huffman_only": dict(level=6, strategy=zlib.Z_HUFFMAN_ONLY),
    "rle": dict(level=6, strategy=zlib.Z_RLE),
    "full_flush_blocks": dict(level=6, flush_every=50_000),
    "sync_flush_empty_stored_blocks": dict(level=6, flush_every=20_011, flush_mode=zlib.Z_SYNC_FLUSH),
}


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("chunk", [0, 40_000, 3_001])
def test_inflate_matches_zlib(name, chunk, text, tmp_path):
    """Every block type and flush pattern, with default and tiny chunks (tiny chunks put chunk borders inside
    blocks, between blocks and -- for the stored case -- inside raw data that looks like anything)."""
    f = tmp_path / "x.gz"
    f.write_bytes(gz(text, **CASES[name]))
    out, info = inflate(f, threads=4, chunk=chunk)
    assert out == text
    assert info["out_bytes"] == len(text) and info["members"] == 1
    if chunk and name in ("level6", "level9", "full_flush_blocks"):
        assert info["chained"] > 4                              # speculation actually happened and was used


def test_inflate_multi_member_and_bgzf_like(text, tmp_path):
    """Concatenated members (cat a.gz b.gz), 64 KiB members with an extra field as bgzip writes them, an empty
    member in the middle, and a header carrying a file name and a comment."""
    parts = [gz(text[:1_000_000], 6), gz(b"", 6), gz(text[1_000_000:3_000_000], 9)]
    bg = []
    for a in range(3_000_000, len(text), 65_280):
        raw = zlib.compressobj(6, zlib.DEFLATED, -15)
        body = raw.compress(text[a:a + 65_280]) + raw.flush()
        block = text[a:a + 65_280]
        hdr = b"\x1f\x8b\x08\x04" + b"\0" * 4 + b"\x00\xff" + (6).to_bytes(2, "little") + b"BC\x02\x00" + (len(body) + 25).to_bytes(2, "little")
        bg.append(hdr + body + zlib.crc32(block).to_bytes(4, "little") + (len(block) & 0xffffffff).to_bytes(4, "little"))
    named = b"\x1f\x8b\x08\x18" + b"\0" * 4 + b"\x00\x03" + b"reads.fastq\0" + b"a comment\0"
    raw = zlib.compressobj(6, zlib.DEFLATED, -15)
    tail = b"@last\nACGT\n+\nIIII\n"
    named += raw.compress(tail) + raw.flush() + zlib.crc32(tail).to_bytes(4, "little") + len(tail).to_bytes(4, "little")
    data = b"".join(parts) + b"".join(bg) + named
    assert gzip.decompress(data) == text + tail
    f = tmp_path / "m.gz"
    f.write_bytes(data)
    for chunk in (0, 30_000, 2_000):
        out, info = inflate(f, threads=4, chunk=chunk)
        assert out == text + tail
        assert info["members"] == 3 + len(bg) + 1


def test_inflate_extreme_ratio_falls_back(tmp_path):
    """1000:1 data overruns the speculative output limit; the chain walk decodes those stretches with the real
    window instead, result unchanged."""
    data = b"A" * 30_000_000 + bytes(np.random.default_rng(3).integers(0, 256, 100_000, dtype=np.uint8)) + b"CG" * 5_000_000
    f = tmp_path / "z.gz"
    f.write_bytes(gz(data, 6))
    out, info = inflate(f, threads=4, chunk=4_000)
    assert out == data
    assert info["fallbacks"] >= 1
    # runs with a known window hand over after KQ_PGZ_SOFT_OUT bytes and waves end after four times that, so
    # memory stays bounded however far the data expands; with a tiny bound the result is still the same
    p = subprocess.run([TOOL, "inflate", str(f), "4", "0"], capture_output=True, timeout=300, env=dict(os.environ, KQ_PGZ_SOFT_OUT="200000"))
    small = json.loads(p.stderr.decode().strip().splitlines()[-1])
    assert p.returncode == 0 and p.stdout == data and small["pieces"] >= 5


def test_inflate_random_binary_and_long_distances(tmp_path):
    """Incompressible bytes (stored blocks whose content mimics block headers and gzip magic) followed by a
    copy of an earlier 32 KiB stretch (maximum distance matches across chunk borders)."""
    rng = np.random.default_rng(5)
    noise = bytes(rng.integers(0, 256, 2_000_000, dtype=np.uint8))
    fake = b"\x1f\x8b\x08\x00" * 1000
    block = bytes(rng.integers(65, 69, 32_700, dtype=np.uint8))
    data = noise + fake + (block + b"xy") * 60 + noise[:500_000]
    f = tmp_path / "b.gz"
    f.write_bytes(gz(data, 9))
    for chunk in (0, 10_000, 1_500):
        out, _ = inflate(f, threads=3, chunk=chunk)
        assert out == data


def test_inflate_errors_and_trailing_garbage(text, tmp_path):
    good = gz(text[:2_000_000], 6)
    f = tmp_path / "e.gz"
    bad = bytearray(good); bad[-8] ^= 1                      # CRC field
    f.write_bytes(bytes(bad))
    _, info = inflate(f, chunk=50_000, expect_ok=False)
    assert not info["ok"] and "CRC" in info["error"]
    bad = bytearray(good); bad[len(good) // 2] ^= 0x10        # payload
    f.write_bytes(bytes(bad))
    _, info = inflate(f, chunk=50_000, expect_ok=False)
    assert not info["ok"]
    f.write_bytes(good[:len(good) // 2])                      # truncated
    _, info = inflate(f, chunk=50_000, expect_ok=False)
    assert not info["ok"]
    f.write_bytes(good + b"\0" * 1000 + b"garbage")           # gzread ignores what follows the last member
    out, info = inflate(f, chunk=50_000)
    assert out == text[:2_000_000]
    f.write_bytes(b"@r\nACGT\n+\nIIII\n")
    _, info = inflate(f, expect_ok=False)
    assert "not a gzip" in info["error"]


def reads_digest(path, env=None, batch=None):
    e = dict(os.environ, **(env or {}))
    p = subprocess.run([TOOL, "reads", str(path)] + ([str(batch)] if batch else []), capture_output=True, timeout=300, env=e)
    assert p.returncode == 0, p.stderr
    f = p.stdout.split()
    return int(f[1]), int(f[2]), f[3].decode()                # bytes, separators, digest (batch count is free)


def expected_batches(fastq):
    """getKmers (functions.h:69-122): the sequence line of every record, followed by 'N' iff the record's
    quality line ended with a newline."""
    lines = fastq.split(b"\n")
    complete = fastq.endswith(b"\n")
    if complete:
        lines.pop()
    out = bytearray()
    for i in range(0, len(lines), 4):
        rec = lines[i:i + 4]
        if len(rec) > 1:
            out += rec[1]
        if len(rec) == 4 and (i + 4 < len(lines) or complete):
            out += b"N"
    return bytes(out)


def py_digest(batch):
    h_total, seps = 0, batch.count(b"N")
    for s in batch.split(b"N"):
        if not s:
            continue
        h = 0xcbf29ce484222325
        for c in s:
            h = ((h ^ c) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
        h_total = (h_total + h * 0x9e3779b97f4a7c15 + len(s)) & 0xFFFFFFFFFFFFFFFF
    return len(batch), seps, "%016x" % h_total


PAR = {"KQ_CLI_PGZ_MIN": "1", "KQ_CLI_PAR_MIN": "1"}
SER = {"KQ_CLI_PGZ_MIN": "99999999999", "KQ_CLI_PAR_MIN": "99999999999"}


@pytest.mark.parametrize("chunk,slice_", [("0", "0"), ("20000", "30000"), ("1500", "700")])
def test_fastq_batches_from_gz(chunk, slice_, tmp_path):
    """What reaches kq_count_push is the same whether a .fastq.gz is decoded in parallel, by zlib, or the plain
    file is read; long reads (records far larger than a chunk) and reads with N included."""
    rng = np.random.default_rng(21)
    fq = fastq_text(rng, 3000) + fastq_text(rng, 40, read_len=(20_000, 60_000))
    want = py_digest(expected_batches(fq))
    plain, packed = tmp_path / "a.fastq", tmp_path / "a.fastq.gz"
    plain.write_bytes(fq)
    packed.write_bytes(gz(fq, 6))
    env = dict(PAR, KQ_CLI_PGZ_CHUNK=chunk, KQ_CLI_SLICE=slice_)
    assert reads_digest(plain, SER) == want
    assert reads_digest(packed, SER) == want
    assert reads_digest(plain, env) == want
    assert reads_digest(packed, env) == want
    assert reads_digest(packed, env, batch=50_000) == want
    serial = dict(env, KQ_FEEDCHECK_SERIAL="1")               # a sink that is not thread safe: calls are serialised
    assert reads_digest(plain, serial) == want and reads_digest(packed, serial, batch=50_000) == want


def test_fastq_gz_truncated_records_and_fasta(tmp_path):
    """End-of-file rules through the parallel path: a file that ends inside a quality line, inside a sequence
    line, right after '+', without the final newline; a FASTA .gz yields nothing (loadKmers, fastx.h:95-118)."""
    rng = np.random.default_rng(22)
    fq = fastq_text(rng, 500)
    nl = [i for i, c in enumerate(fq) if c == 10]
    env = dict(PAR, KQ_CLI_PGZ_CHUNK="900", KQ_CLI_SLICE="500")
    for cut in (nl[4 * 100 + 3] - 20, nl[4 * 200 + 1] - 7, nl[4 * 300 + 2] + 1, len(fq) - 1, nl[4 * 50 + 3] + 1, nl[4 * 50] + 1):
        f = tmp_path / "t.fastq.gz"
        f.write_bytes(gz(fq[:cut], 6))
        want = py_digest(expected_batches(fq[:cut]))
        assert reads_digest(f, SER) == want, cut
        assert reads_digest(f, env) == want, cut
    f = tmp_path / "g.fa.gz"
    f.write_bytes(gz(b">s1\n" + b"ACGT" * 5000 + b"\n", 6))
    assert reads_digest(f, env) == (0, 0, "%016x" % 0)


def test_reference_fixture_gz_and_genome_loader(tmp_path):
    """The reference's own gz fixtures (testFiles/random1.fastq.gz, random2.fastq.gz) through the parallel
    decoder, and a gzipped multi-record FASTA through load_genome both ways."""
    for name in ("random1.fastq", "random2.fastq"):
        plain, packed = os.path.join(GOLD, "inputs", name), os.path.join(GOLD, "inputs", name + ".gz")
        if not os.path.exists(packed):
            pytest.skip("fixture missing")
        out, _ = inflate(packed, threads=2, chunk=300)
        with open(plain, "rb") as f:
            assert out == f.read()
        assert reads_digest(packed, dict(PAR, KQ_CLI_PGZ_CHUNK="300", KQ_CLI_SLICE="200")) == reads_digest(plain, SER)
    rng = np.random.default_rng(23)
    fa = b""
    for i in range(5):
        s = rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), size=200_000, p=[.24, .24, .24, .24, .04]).tobytes()
        fa += b">chr%d some description\n" % i + b"\n".join(s[a:a + 60] for a in range(0, len(s), 60)) + b"\n"
    f = tmp_path / "g.fa.gz"
    f.write_bytes(gz(fa, 6))

    def genome(env):
        p = subprocess.run([TOOL, "genome", str(f)], capture_output=True, timeout=300, env=dict(os.environ, **env))
        assert p.returncode == 0, p.stderr
        return p.stdout
    a, b = genome(dict(PAR, KQ_CLI_PGZ_CHUNK="5000")), genome(SER)
    assert a == b and int(a.split()[0]) == 5


def _fnv(s):
    h = 0xcbf29ce484222325
    for c in s:
        h = ((h ^ c) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
    return (h * 0x9e3779b97f4a7c15 + len(s)) & 0xFFFFFFFFFFFFFFFF


def test_fasta_loader_records_segments_gaps(tmp_path):
    """load_genome on FASTA (src/input.cpp:206-228, gfalibs/src/gfa.cpp:90-304): header = first word, sequence =
    everything up to the next '>', newlines dropped, runs of N/n/X/x are gaps between segments.  Records are
    parsed in parallel; the result must equal a direct restatement, in file order."""
    import re
    rng = np.random.default_rng(31)
    recs = []
    for i in range(37):
        n = int(rng.integers(0, 60_000)) if i % 9 else 0
        s = bytearray(rng.choice(np.frombuffer(b"ACGTacgt", dtype=np.uint8), size=n).tobytes())
        for _ in range(int(rng.integers(0, 6))):
            if n > 100:
                a = int(rng.integers(0, n - 50)); ln = int(rng.integers(1, 50))
                s[a:a + ln] = bytes(rng.choice(np.frombuffer(b"NnXx", dtype=np.uint8), size=ln))
        if i == 3 and n:
            s[:5] = b"NNNNN"                                    # gap first
        if i == 4 and n:
            s[-3:] = b"xxx"                                     # gap last
        recs.append((b"seq%d" % i, b" desc > with gt" if i == 5 else b"", bytes(s)))
    fa = b""
    for h, d, s in recs:
        w = int(rng.integers(40, 90))
        fa += b">" + h + d.replace(b">", b"") + b"\n" + b"".join(s[a:a + w] + b"\n" for a in range(0, len(s), w))
    fa = fa[:-1]                                                # no final newline
    h_total, nseg, bases = 0, 0, 0
    seg_hashes, path_hashes = [], []
    for h, d, s in recs:
        comps = [m for m in re.finditer(rb"[NnXx]+|[^NnXx]+", s)]
        for m in comps:
            if m.group()[0:1] not in (b"N", b"n", b"X", b"x"):
                seg_hashes.append(_fnv(m.group())); nseg += 1; bases += len(m.group())
        path_hashes.append((_fnv(h) + len(comps)) & 0xFFFFFFFFFFFFFFFF)
    for x in seg_hashes:
        h_total = (h_total * 31 + x) & 0xFFFFFFFFFFFFFFFF
    for x in path_hashes:
        h_total = (h_total * 31 + x) & 0xFFFFFFFFFFFFFFFF
    want = b"%d %d %d %016x\n" % (len(recs), nseg, bases, h_total)
    f = tmp_path / "g.fa"
    f.write_bytes(fa)
    for threads in ("1", "3", "8"):
        p = subprocess.run([TOOL, "genome", str(f)], capture_output=True, timeout=120, env=dict(os.environ, KQ_CLI_THREADS=threads))
        assert p.returncode == 0 and p.stdout == want, threads


def test_inflate_fuzz_small_inputs(tmp_path):
    """Seeded sweep over content types, compression parameters, member layouts, chunk sizes and thread counts on
    small inputs (0 bytes ... 300 KB, sizes around the 32 KiB window), every result compared with zlib."""
    rng = np.random.default_rng(77)
    alphabet = [np.frombuffer(b"ACGT", dtype=np.uint8), np.frombuffer(b"ACGTN\n@+IIIIFFFF:,#", dtype=np.uint8), np.arange(256, dtype=np.uint8)]
    sizes = [0, 1, 2, 100, 32767, 32768, 32769, 65536, 70001, 150_000, 300_000]
    for case in range(40):
        n = sizes[case % len(sizes)] if case < 22 else int(rng.integers(0, 300_000))
        kind = int(rng.integers(0, 4))
        if kind == 3:                                            # long repeats with occasional changes
            unit = bytes(rng.choice(alphabet[0], size=int(rng.integers(1, 400))))
            data = bytearray((unit * (n // max(1, len(unit)) + 1))[:n])
            for p in rng.integers(0, max(1, n), size=min(n, 20)):
                if n:
                    data[int(p)] = int(rng.integers(0, 256))
            data = bytes(data)
        else:
            data = bytes(rng.choice(alphabet[kind], size=n))
        level = int(rng.integers(1, 10))
        strategy = [zlib.Z_DEFAULT_STRATEGY, zlib.Z_FILTERED, zlib.Z_RLE, zlib.Z_FIXED][int(rng.integers(0, 4))]
        nmembers = int(rng.integers(1, 5))
        cuts = sorted(int(x) for x in rng.integers(0, n + 1, size=nmembers - 1)) if n else [0] * (nmembers - 1)
        blob, prev = b"", 0
        for c in cuts + [n]:
            flush = int(rng.integers(0, 3))
            blob += gz(data[prev:c], level, strategy, flush_every=[0, 4000, 9973][flush], flush_mode=[zlib.Z_FULL_FLUSH, zlib.Z_SYNC_FLUSH][case % 2])
            prev = c
        assert gzip.decompress(blob) == data
        f = tmp_path / ("f%d.gz" % case)
        f.write_bytes(blob)
        chunk = int(rng.integers(64, 6000)) if case % 3 else 0
        out, info = inflate(f, threads=int(rng.integers(1, 6)), chunk=chunk)
        assert out == data, (case, n, kind, level, strategy, nmembers, chunk)
        assert info["members"] == nmembers

"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/kreeq_b200.h declares,
fails loudly without a GPU, and its host-only pieces (2-bit packer, .kreeq codec, owner rule, QV
arithmetic) agree with the oracle and with reference-written fixtures.  No compute calls."""
import ctypes as C
import os
import re
import tarfile

import numpy as np
import pytest

import kreeq_oracle as ko
from helpers import random_genome

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def kb():
    import kreeq_b200
    return kreeq_b200


def header_symbols():
    with open(os.path.join(ROOT, "include", "kreeq_b200.h")) as f:
        text = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(kq_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(kb):
    from kreeq_b200 import _lib
    L = kb.lib()
    syms = header_symbols()
    assert len(syms) >= 35
    for s in syms:
        assert hasattr(L, s), "libkreeq_b200.so does not export " + s
        assert s in _lib.PROTOTYPES, "no ctypes prototype for " + s
    assert sorted(_lib.PROTOTYPES) == syms


def test_library_has_sm100a_code_only():
    out = os.popen("cuobjdump -lelf %s 2>/dev/null" % os.path.join(ROOT, "kreeq_b200", "libkreeq_b200.so")).read()
    if out.strip():
        assert "sm_100a" in out and not re.search(r"sm_(?!100a)\d+", out)


def test_no_cpu_fallback_without_device(kb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(kb.KreeqError) as e:
        kb.DBG(k=21)
    assert e.value.code == -2          # KQ_ENODEV
    with pytest.raises(kb.KreeqError) as e:
        kb.DBG(k=33)
    assert e.value.code == -1          # KQ_EINVAL "Invalid kmer length." src/input.cpp:142


def test_product_never_touches_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "kreeq_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                with open(os.path.join(dirpath, fn), errors="ignore") as f:
                    src = f.read()
                assert "kreeq_oracle" not in src and "oracle/" not in src, fn


def test_owner_rule(kb):
    L = kb.lib()
    keys = np.random.default_rng(1).integers(0, 1 << 62, 1000, dtype=np.uint64)
    for n in (1, 2, 4, 8, 16, 128):
        mine = np.array([L.kq_owner_of(int(k), n) for k in keys])
        assert np.array_equal(mine, (keys % np.uint64(128)) // np.uint64(128 // n))
    assert L.kq_owner_of(5, 3) == -1 and L.kq_owner_of(5, 0) == -1


def test_qv_arithmetic(kb):
    L = kb.lib()
    for miss, tot, k in ((177, 711, 21), (0, 1000, 21), (5, 123456789, 31), (711, 711, 15)):
        assert L.kq_error_rate(miss, tot, k) == ko.lib().ko_error_rate(miss, tot, k)
        a, b = L.kq_qv_value(miss, tot, k), ko.lib().ko_qv_value(miss, tot, k)
        assert a == b or (np.isinf(a) and np.isinf(b))


def test_pack_host(kb):
    rng = np.random.default_rng(3)
    for n in (0, 1, 31, 32, 33, 63, 64, 65, 1000, 100003):
        seq = random_genome(rng, n)
        if n:
            seq[rng.integers(0, n, max(1, n // 50))] = ord("N")
            seq[rng.integers(0, n, max(1, n // 70))] = ord("c")
        bases, inv = kb.pack_host(seq.tobytes())
        assert bases.size == kb.lib().kq_packed_words(n) == inv.size
        code = np.full(256, 4, dtype=np.uint8)
        for i, ch in enumerate(b"ACGT"):
            code[ch] = i; code[ch + 32] = i
        c = code[seq]
        for i in range(0, n, max(1, n // 200)):
            b = (int(bases[i // 32]) >> (2 * (i % 32))) & 3
            v = (int(inv[i // 64]) >> (i % 64)) & 1
            assert v == (c[i] == 4) and (v or b == c[i])
        # everything past the end is invalid
        for i in range(n, min(inv.size * 64, n + 130)):
            assert (int(inv[i // 64]) >> (i % 64)) & 1 == 1
        assert kb.pack_host(seq.tobytes(), nthreads=1)[0].tobytes() == bases.tobytes()


def test_codec_roundtrip_is_phmap_loadable(kb, tmp_path):
    """kq_db_write_records -> strict phmap layout check by the oracle's decoder -> kq_db_read_records."""
    rng = np.random.default_rng(5)
    n = 20000
    keys = np.unique(rng.integers(0, 1 << 42, n, dtype=np.uint64))
    recs = np.zeros(keys.size, dtype=kb.REC32)
    recs["cov"] = rng.integers(1, 200, keys.size)
    recs["cov"][::97] = rng.integers(255, 100000, recs["cov"][::97].size)
    recs["cov"][5] = 4294967295; recs["cov"][6] = 254; recs["cov"][7] = 255
    for f in ("fw", "bw"):
        recs[f] = (rng.random((keys.size, 4)) * (recs["cov"][:, None] + 1)).astype(np.uint32)
    d = str(tmp_path / "w.kreeq")
    kb.db_write_records(d, 21, keys, recs)
    assert kb.lib().kq_db_read_k(d.encode()) == 21
    assert open(os.path.join(d, ".index")).read() == "21\n128\n"
    t = ko.read_db(d, check=True)
    ek, er = t.export()
    assert np.array_equal(ek, keys) and er.tobytes() == recs.tobytes()
    k8, r8 = t.dump8(); k32, _ = t.dump32()
    assert np.array_equal(k32, keys[recs["cov"] >= 255]) and np.all(r8["cov"][np.isin(k8, k32)] == 255)
    rk, rr = kb.db_read_records(d)
    assert np.array_equal(rk, keys) and rr.tobytes() == recs.tobytes()


def test_codec_reads_reference_written_databases(kb, tmp_path):
    with tarfile.open(os.path.join(GOLD, "fixture_dbs.tar.gz")) as tar:
        tar.extractall(tmp_path)
    with tarfile.open(os.path.join(GOLD, "ref_runs", "union_inputs.tar.gz")) as tar:
        tar.extractall(tmp_path)
    for name in ("test1.kreeq", "test2.kreeq", "u0.kreeq", "u1.kreeq", "u2.kreeq"):
        d = str(tmp_path / name)
        rk, rr = kb.db_read_records(d)
        ek, er = ko.read_db(d, check=False).export()
        assert np.array_equal(rk, ek) and rr.tobytes() == er.tobytes()


def test_codec_errors(kb, tmp_path):
    assert kb.lib().kq_db_read_k(str(tmp_path / "nope").encode()) == -5      # KQ_EIO
    n = C.c_size_t(0)
    assert kb.lib().kq_db_read_records(str(tmp_path / "nope").encode(), None, None, C.byref(n)) == -5
    empty = str(tmp_path / "e.kreeq")
    kb.db_write_records(empty, 31, np.zeros(0, dtype=np.uint64), np.zeros(0, dtype=kb.REC32))
    for m in (0, 127):
        assert os.path.getsize(os.path.join(empty, ".map.%d.bin" % m)) == 6152
    assert os.path.getsize(os.path.join(empty, ".map.hc.bin")) == 6152
    k, r = kb.db_read_records(empty)
    assert k.size == 0


REF_BIN = os.path.join(ROOT, "oracle", "_ref", "kreeq")


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/kreeq not built")
def test_reference_binary_loads_our_database(kb, tmp_path):
    """A database written by our codec is read by the UNMODIFIED reference (`kreeq validate -d`, `kreeq union`)
    and gives the reference's own golden stdout (validateFiles/test.0.tst, test.35.tst)."""
    import json
    import subprocess
    import genome as gn
    with open(os.path.join(GOLD, "ref_golden.json")) as f:
        tst = json.load(f)
    dirs = []
    for i, fq in enumerate(("random1.fastq", "random2.fastq", "repeat1.fastq")):
        t = ko.Table(21)
        for b in ko.fastq_batches(gn.read_maybe_gz(os.path.join(GOLD, "inputs", fq))):
            t.count_batch(b)
        keys, recs = t.export()
        d = str(tmp_path / ("ours%d.kreeq" % i)); dirs.append(d)
        kb.db_write_records(d, 21, keys, recs.astype(kb.REC32))
    fa = os.path.join(GOLD, "inputs", "random1.fasta")
    out = subprocess.run([REF_BIN, "validate", "-f", fa, "-d", dirs[0], "-o", str(tmp_path / "o.kwig")], cwd=tmp_path,
                         stdout=subprocess.PIPE, stderr=subprocess.PIPE, check=True).stdout.decode()
    assert out == tst["test.0"]["stdout"]
    out = subprocess.run([REF_BIN, "union", "-d", dirs[0], dirs[1], "-o", str(tmp_path / "u.kreeq")], cwd=tmp_path,
                         stdout=subprocess.PIPE, stderr=subprocess.PIPE, check=True).stdout.decode()
    assert out == tst["test.35"]["stdout"]
    # high-copy database (32-bit spill map): per-base profile of repeat1.fasta as in validateFiles/test.49
    subprocess.run([REF_BIN, "validate", "-f", os.path.join(GOLD, "inputs", "repeat1.fasta"), "-d", dirs[2], "-o", str(tmp_path / "r.kwig")],
                   cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.PIPE, check=True)
    import gzip
    with open(os.path.join(GOLD, "ref_runs", "repeat1_k21.kwig.gz"), "rb") as f:
        assert open(tmp_path / "r.kwig", "rb").read() == gzip.decompress(f.read())


def test_cli_binary_usage_and_no_cpu_fallback(tmp_path):
    """kreeq_b200/kreeq is the reference's command line over the C-ABI: usage text works anywhere, a run
    without a CUDA device fails loudly instead of computing on the CPU."""
    import subprocess
    import torch
    exe = os.path.join(ROOT, "kreeq_b200", "kreeq")
    assert os.path.exists(exe)
    p = subprocess.run([exe], capture_output=True)
    assert p.returncode == 0 and b"validate" in p.stdout and b"union" in p.stdout
    p = subprocess.run([exe, "validate", "-h"], capture_output=True)
    assert p.returncode == 0 and b"--input-reads" in p.stdout and b"--coverage-cutoff" in p.stdout
    p = subprocess.run([exe, "union", "-d", os.path.join(GOLD, "inputs", "random1.fastq")], capture_output=True)
    assert p.returncode != 0 and b"At least two databases required (-d)." in p.stderr
    if not torch.cuda.is_available():
        reads = os.path.join(GOLD, "inputs", "random1.fastq")
        p = subprocess.run([exe, "validate", "-r", reads], capture_output=True, cwd=str(tmp_path))
        assert p.returncode != 0 and p.stdout == b"" and p.stderr.strip()


def test_cli_argument_errors_match_the_reference(tmp_path):
    """Failures before any k-mer work (missing files, bad numbers) print what the reference prints, where it
    prints it (ifFileExists writes to stdout, gfalibs/include/functions.h:260-272), with the same exit status."""
    import subprocess
    exe = os.path.join(ROOT, "kreeq_b200", "kreeq")
    ref = os.path.join(ROOT, "oracle", "_ref", "kreeq")
    p = subprocess.run([exe, "validate", "-f", "/nonexistent.fa"], capture_output=True, cwd=str(tmp_path))
    assert p.returncode == 1 and p.stdout == b"Error - file does not exist: /nonexistent.fa\n"
    assert subprocess.run([exe], capture_output=True).stdout.startswith(b"kreeq [mode] -h\nfor additional help.\n\nModes:\nvalidate\nunion\n")
    if not os.path.exists(ref):
        pytest.skip("oracle/_ref/kreeq not built")
    for args in (["validate", "-f", "/nonexistent.fa"], ["validate", "-r", "/nonexistent.fq"], ["union", "-d", "a"], ["validate", "-d", "no.kreeq"],
                 ["validate", "-c", "x"], ["validate", "-k"], ["validate", "-v"]):
        a = subprocess.run([exe] + args, capture_output=True, cwd=str(tmp_path))
        b = subprocess.run([ref] + args, capture_output=True, cwd=str(tmp_path))
        assert (a.returncode, a.stdout) == (b.returncode, b.stdout), args

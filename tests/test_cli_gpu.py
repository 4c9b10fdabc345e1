"""The `kreeq` command line (kreeq_b200/kreeq, C++ over the C-ABI) against the reference's own outputs:
its 36 stdout goldens (validateFiles/test.0-35.tst, commands unchanged apart from the input directory) and
the committed reference-binary runs (database content, stdout, .kwig / .bed / .bkwig bytes, union orders).
Nothing here reads /root/reference; everything runs on the GPU through libkreeq_b200.so."""
import gzip
import json
import os
import shlex
import subprocess
import tarfile

import numpy as np
import pytest

import kreeq_oracle as ko

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
EXE = os.path.join(ROOT, "kreeq_b200", "kreeq")

with open(os.path.join(GOLD, "ref_golden.json")) as f:
    REF = json.load(f)
with open(os.path.join(GOLD, "ref_runs.json")) as f:
    RUNS = json.load(f)


def run(args, cwd):
    p = subprocess.run([EXE] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    assert p.returncode == 0, p.stderr.decode()
    return p.stdout.decode()


def _gz(rel):
    with open(os.path.join(GOLD, rel), "rb") as f:
        return gzip.decompress(f.read())


@pytest.fixture(scope="module")
def fixture_dbs(tmp_path_factory):
    d = tmp_path_factory.mktemp("fixture_dbs")
    with tarfile.open(os.path.join(GOLD, "fixture_dbs.tar.gz")) as tar:
        tar.extractall(d)
    return str(d)


@pytest.mark.parametrize("name", sorted(REF, key=lambda s: int(s.split(".")[1])))
def test_reference_goldens_through_the_cli(name, fixture_dbs, tmp_path):
    """Same command lines as validateFiles/test.N.tst; stdout must be byte-identical."""
    args = shlex.split(REF[name]["cmd"])[1:]
    fixed = []
    for a in args:
        if a.startswith("testFiles/") and a.endswith(".kreeq"):
            a = os.path.join(fixture_dbs, a[len("testFiles/"):])
        elif a.startswith("testFiles/"):
            a = os.path.join(GOLD, "inputs", a[len("testFiles/"):])
        fixed.append(a)
    assert run(fixed, str(tmp_path)) == REF[name]["stdout"]


def _db_arrays(path):
    t = ko.read_db(path, check=True)
    return t.dump8() + t.dump32()


@pytest.mark.parametrize("name", [n for n in sorted(RUNS) if n != "union"])
def test_reference_binary_runs_through_the_cli(name, tmp_path):
    """count -> .kreeq (decoded records equal the reference's), then validate from that database into
    .kwig / .bed / .bkwig: stdout and file bytes equal the reference binary's."""
    r = RUNS[name]
    k, cutoff = r["k"], r["cutoff"]
    reads = [os.path.join(GOLD, p) for p in r["reads"]]
    asm = os.path.join(GOLD, r["asm"])
    db = str(tmp_path / "db.kreeq")
    assert run(["validate", "-r"] + reads + ["-k", str(k), "-o", db], str(tmp_path)) == r["count_stdout"]
    k8, r8, k32, r32 = _db_arrays(db)
    gold = np.load(os.path.join(GOLD, "ref_runs", name + ".db.npz"))
    assert np.array_equal(k8, gold["k8"]) and np.array_equal(k32, gold["k32"])
    live = r8["cov"] != 255
    assert r8[live].tobytes() == gold["r8"][live].tobytes()
    assert r32.tobytes() == gold["r32"].tobytes()
    for ext in ("kwig", "bed", "bkwig"):
        out = str(tmp_path / ("x." + ext))
        stdout = run(["validate", "-f", asm, "-d", db, "-k", str(k), "-c", str(cutoff), "-o", out], str(tmp_path))
        assert stdout == r["validate_stdout"]
        with open(out, "rb") as f:
            assert f.read() == _gz("ref_runs/%s.%s.gz" % (name, ext)), ext
    # one-shot form (reads + assembly in one call) prints the same two blocks
    assert run(["validate", "-f", asm, "-r"] + reads + ["-k", str(k), "-c", str(cutoff)], str(tmp_path)) == r["validate_stdout"]


@pytest.fixture(scope="module")
def union_inputs(tmp_path_factory):
    d = tmp_path_factory.mktemp("union_inputs")
    with tarfile.open(os.path.join(GOLD, "ref_runs", "union_inputs.tar.gz")) as tar:
        tar.extractall(d)
    return str(d)


@pytest.mark.parametrize("tag", sorted(RUNS["union"]))
def test_union_orders_through_the_cli(union_inputs, tag, tmp_path):
    """`kreeq union` is order dependent for high-copy keys: all three orders."""
    g = RUNS["union"][tag]
    dbs = [os.path.join(union_inputs, "u%d.kreeq" % i) for i in g["order"]]
    out = str(tmp_path / "u.kreeq")
    assert run(["union", "-d"] + dbs + ["-o", out], str(tmp_path)) == g["stdout"]
    k8, r8, k32, r32 = _db_arrays(out)
    gold = np.load(os.path.join(GOLD, "ref_runs", "union%s.db.npz" % tag))
    assert np.array_equal(k8, gold["k8"]) and np.array_equal(k32, gold["k32"])
    assert r32.tobytes() == gold["r32"].tobytes()
    live = r8["cov"] != 255
    assert r8[live].tobytes() == gold["r8"][live].tobytes()
    kwig = str(tmp_path / "u.kwig")
    asm = os.path.join(GOLD, "inputs", "repeat1.fasta")
    assert run(["validate", "-f", asm, "-d", out, "-o", kwig], str(tmp_path)) == g["validate_stdout"]
    with open(kwig) as f:
        assert f.read() == g["kwig"]


def test_hist_and_errors(tmp_path):
    reads = os.path.join(GOLD, "inputs", "random1.fastq")
    hist = str(tmp_path / "x.hist")
    run(["validate", "-r", reads, "-o", hist], str(tmp_path))
    rows = [tuple(map(int, ln.split("\t"))) for ln in open(hist).read().splitlines()]
    assert rows == sorted(rows) and sum(c * n for c, n in rows) == 172 and sum(n for _, n in rows) == 96
    p = subprocess.run([EXE, "validate", "-r", reads, "-k", "33"], cwd=str(tmp_path), capture_output=True)
    assert p.returncode != 0 and b"Invalid kmer length." in p.stderr
    p = subprocess.run([EXE, "union", "-d", reads], cwd=str(tmp_path), capture_output=True)
    assert p.returncode != 0 and b"At least two databases required (-d)." in p.stderr
    p = subprocess.run([EXE, "frobnicate", "-x", "y"], cwd=str(tmp_path), capture_output=True)
    assert p.returncode != 0 and b"does not exist" in p.stderr


REF_BIN = os.path.join(ROOT, "oracle", "_ref", "kreeq")


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/kreeq not built")
@pytest.mark.parametrize("name", ["synthA_k21", "synthC_k32_c3", "random1x2_k21"])
def test_reference_binary_reads_gpu_written_database(name, tmp_path):
    """The .map.<m>.bin images are laid out by CUDA kernels (kq_db_kernels.cu); the UNMODIFIED reference must
    load them and print what it prints for its own database (checker use of oracle/_ref only)."""
    r = RUNS[name]
    k, cutoff = r["k"], r["cutoff"]
    reads = [os.path.join(GOLD, p) for p in r["reads"]]
    asm = os.path.join(GOLD, r["asm"])
    db = str(tmp_path / "db.kreeq")
    run(["validate", "-r"] + reads + ["-k", str(k), "-o", db], str(tmp_path))
    out = str(tmp_path / "x.kwig")
    p = subprocess.run([REF_BIN, "validate", "-f", asm, "-d", db, "-k", str(k), "-c", str(cutoff), "-o", out, "-j", "2"],
                       cwd=str(tmp_path), stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    assert p.returncode == 0, p.stderr.decode()
    assert p.stdout.decode() == r["validate_stdout"]
    with open(out, "rb") as f:
        assert f.read() == _gz("ref_runs/%s.kwig.gz" % name)


def test_host_and_device_codecs_agree(tmp_path):
    """KQ_DB_HOST=1 selects the host implementation of the file codec: both write databases that decode to
    the same records, and each loads what the other wrote."""
    r = RUNS["synthB_k31"]
    reads = [os.path.join(GOLD, p) for p in r["reads"]]
    asm = os.path.join(GOLD, r["asm"])
    dbs = {}
    for mode in ("0", "1"):
        env = dict(os.environ, KQ_DB_HOST=mode)
        db = str(tmp_path / ("db%s.kreeq" % mode))
        p = subprocess.run([EXE, "validate", "-r"] + reads + ["-k", "31", "-o", db], cwd=str(tmp_path), env=env, capture_output=True)
        assert p.returncode == 0 and p.stdout.decode() == r["count_stdout"]
        dbs[mode] = db
    a, b = _db_arrays(dbs["0"]), _db_arrays(dbs["1"])
    for x, y in zip(a, b):
        assert x.tobytes() == y.tobytes()
    for writer, reader in (("0", "1"), ("1", "0")):
        env = dict(os.environ, KQ_DB_HOST=reader)
        p = subprocess.run([EXE, "validate", "-f", asm, "-d", dbs[writer], "-k", "31"], cwd=str(tmp_path), env=env, capture_output=True)
        assert p.returncode == 0 and p.stdout.decode() == r["validate_stdout"]


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/kreeq not built")
def test_union_with_high_copy_repeats_at_scale(tmp_path):
    """A scaled C3: a 1 Mbp genome with 32 copies of a 5 kb repeat and a 2 kb poly-A run, four
    13x read sets counted separately, then `union -d s0 s1 s2 s3`.  Every database and the union hold k-mers
    beyond the 8-bit range (32-bit spill, order-dependent merge rules of row A4).  The same commands run
    through the unmodified reference on the box; decoded databases, stdout and .kwig must be identical."""
    import sys
    import torch
    sys.path.insert(0, ROOT)
    import bench
    from kreeq_b200 import synth
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(77)
    genome = synth.random_genome_codes(1_000_000, 30, dev)
    repeat = torch.randint(0, 4, (5000,), dtype=torch.uint8, device=dev, generator=g)
    for pos in torch.randint(0, 990_000, (32,), device=dev, generator=g).tolist():
        genome[pos:pos + 5000] = repeat
    genome[500_000:502_000] = 0                                              # poly-A
    asm = str(tmp_path / "asm.fasta")
    bench.write_fasta(asm, synth.codes_to_ascii(genome).cpu().numpy())
    L, pairs = 150, 43_500
    ours, refs = [], []
    for i in range(4):
        fq = str(tmp_path / ("s%d.fastq" % i))
        bench.write_fastq(fq, synth.paired_reads_ascii(genome, pairs, L, 0.01, 30 + i).cpu().numpy(), 2 * pairs, L)
        o, r = str(tmp_path / ("ours%d.kreeq" % i)), str(tmp_path / ("ref%d.kreeq" % i))
        so = run(["validate", "-r", fq, "-o", o], str(tmp_path))
        p = subprocess.run([REF_BIN, "validate", "-r", fq, "-o", r, "-j", "8"], cwd=str(tmp_path), capture_output=True, timeout=900)
        assert p.returncode == 0 and p.stdout.decode() == so
        ours.append(o); refs.append(r)

    def same_db(a, b):
        ta, tb = ko.read_db(a, check=False), ko.read_db(b, check=False)
        (ka8, ra8), (kb8, rb8) = ta.dump8(), tb.dump8()
        (ka32, ra32), (kb32, rb32) = ta.dump32(), tb.dump32()
        assert np.array_equal(ka8, kb8) and np.array_equal(ka32, kb32) and ka32.size > 1000
        assert ra32.tobytes() == rb32.tobytes()
        live = ra8["cov"] != 255
        assert np.array_equal(live, rb8["cov"] != 255) and ra8[live].tobytes() == rb8[live].tobytes()

    uo, ur, ux = str(tmp_path / "uo.kreeq"), str(tmp_path / "ur.kreeq"), str(tmp_path / "ux.kreeq")
    so = run(["union", "-d"] + ours + ["-o", uo], str(tmp_path))
    p = subprocess.run([REF_BIN, "union", "-d"] + refs + ["-o", ur, "-j", "8"], cwd=str(tmp_path), capture_output=True, timeout=900)
    assert p.returncode == 0 and p.stdout.decode() == so
    same_db(uo, ur)
    assert run(["union", "-d"] + refs + ["-o", ux], str(tmp_path)) == so          # our union of reference-written databases
    same_db(ux, ur)
    ko_, kr_ = str(tmp_path / "o.kwig"), str(tmp_path / "r.kwig")
    vo = run(["validate", "-f", asm, "-d", uo, "-c", "3", "-o", ko_], str(tmp_path))
    p = subprocess.run([REF_BIN, "validate", "-f", asm, "-d", ur, "-k", "21", "-c", "3", "-o", kr_, "-j", "8"], cwd=str(tmp_path), capture_output=True, timeout=900)
    assert p.returncode == 0 and p.stdout.decode() == vo
    with open(ko_, "rb") as a, open(kr_, "rb") as b:
        assert a.read() == b.read()


@pytest.mark.parametrize("slice_bytes", ["40", "3000"])
def test_parallel_fastq_slicing_is_exact(slice_bytes, fixture_dbs, tmp_path):
    """Large plain FASTQ files are cut into slices at record boundaries found by counting newlines; with the
    size thresholds lowered the reference's goldens (incl. the N-containing and two-file cases) and a file that
    ends inside a record go through that path and must print the same text."""
    env = dict(os.environ, KQ_CLI_PAR_MIN="1", KQ_CLI_SLICE=slice_bytes)
    for name in ("test.1", "test.3", "test.8"):
        args = [a.replace("testFiles/", os.path.join(GOLD, "inputs") + "/") for a in shlex.split(REF[name]["cmd"])[1:]]
        p = subprocess.run([EXE] + args, cwd=str(tmp_path), env=env, capture_output=True, timeout=300)
        assert p.returncode == 0 and p.stdout.decode() == REF[name]["stdout"], name
    # truncated inputs: cut random2.fastq inside a quality line, inside a sequence line and right after a '+' line
    with open(os.path.join(GOLD, "inputs", "random2.fastq"), "rb") as f:
        data = f.read()
    nl = [i for i, c in enumerate(data) if c == 10]
    for cut in (nl[3 * 4 + 3] - 20, nl[5 * 4 + 1] - 7, nl[7 * 4 + 2] + 1):
        fq = str(tmp_path / "cut.fastq")
        with open(fq, "wb") as f:
            f.write(data[:cut])
        base = subprocess.run([EXE, "validate", "-r", fq, "-k", "15"], cwd=str(tmp_path), capture_output=True, timeout=300)
        par = subprocess.run([EXE, "validate", "-r", fq, "-k", "15"], cwd=str(tmp_path), env=env, capture_output=True, timeout=300)
        assert base.returncode == 0 and par.returncode == 0 and base.stdout == par.stdout and b"Total kmers" in par.stdout
        if os.path.exists(REF_BIN):                       # and the reference agrees on the end-of-file rules
            ref = subprocess.run([REF_BIN, "validate", "-r", fq, "-k", "15", "-o", str(tmp_path / "r.kreeq")], cwd=str(tmp_path), capture_output=True, timeout=300)
            assert ref.stdout == par.stdout


def test_parallel_gzip_inputs_are_exact(tmp_path):
    """.fastq.gz / .fasta.gz inputs through the parallel gzip decoder (csrc/cli/pgzip.cpp): the reference's
    goldens that read gz files, with the size threshold lowered and tiny chunks so that chunk borders fall
    inside deflate blocks; then a multi-megabyte two-file data set, gz against plain, default chunking."""
    env = dict(os.environ, KQ_CLI_PGZ_MIN="1", KQ_CLI_PGZ_CHUNK="300", KQ_CLI_PAR_MIN="1", KQ_CLI_SLICE="200")
    names = [n for n in REF if ".gz" in REF[n]["cmd"]]
    assert len(names) >= 8
    for name in names:
        args = [a.replace("testFiles/", os.path.join(GOLD, "inputs") + "/") for a in shlex.split(REF[name]["cmd"])[1:]]
        p = subprocess.run([EXE] + args, cwd=str(tmp_path), env=env, capture_output=True, timeout=300)
        assert p.returncode == 0 and p.stdout.decode() == REF[name]["stdout"], name
    from helpers import random_genome
    rng = np.random.default_rng(77)
    genome = random_genome(rng, 300_000).tobytes()
    files = []
    for r in range(2):
        recs = []
        for i in range(40_000):
            s = int(rng.integers(0, len(genome) - 150))
            read = bytearray(genome[s:s + 150])
            for e in np.nonzero(rng.random(150) < 0.01)[0]:
                read[e] = b"ACGT"[int(rng.integers(0, 4))]
            recs.append(b"@r%d/%d\n" % (i, r + 1) + bytes(read) + b"\n+\n" + bytes(rng.integers(35, 74, 150, dtype=np.uint8)) + b"\n")
        fq = tmp_path / ("R%d.fastq" % (r + 1))
        fq.write_bytes(b"".join(recs))
        with open(str(fq) + ".gz", "wb") as f:
            f.write(gzip.compress(fq.read_bytes(), 6))
        files.append(str(fq))
    fa = tmp_path / "asm.fasta"
    fa.write_bytes(b">chr1\n" + genome + b"\n")
    with open(str(fa) + ".gz", "wb") as f:
        f.write(gzip.compress(fa.read_bytes(), 6))
    plain = run(["validate", "-f", str(fa), "-r"] + files + ["-o", str(tmp_path / "a.kwig")], str(tmp_path))
    env = dict(os.environ, KQ_CLI_PGZ_MIN="1")
    p = subprocess.run([EXE, "validate", "-f", str(fa) + ".gz", "-r"] + [f + ".gz" for f in files] + ["-o", str(tmp_path / "b.kwig")],
                       cwd=str(tmp_path), env=env, capture_output=True, timeout=300)
    assert p.returncode == 0 and p.stdout.decode() == plain and "Total kmers" in plain
    with open(tmp_path / "a.kwig", "rb") as a, open(tmp_path / "b.kwig", "rb") as b:
        assert a.read() == b.read()


def test_bkwig_written_on_the_gpu_path_inflates_to_the_kwig(tmp_path):
    """`-o x.bkwig` and `-o x.kwig` of the same run carry the same triples: our .bkwig reader and, where it was
    built, the reference's own kreeq-decompressor turn the binary file into the text file byte for byte."""
    asm = os.path.join(tmp_path, "asm.fasta")
    with open(asm, "wb") as f:
        f.write(_gz("ref_runs/synthA_k21.fasta.gz"))
    fq = os.path.join(tmp_path, "r.fastq")
    with open(fq, "wb") as f:
        f.write(_gz("ref_runs/synthA_k21.fastq.gz"))
    run(["validate", "-f", asm, "-r", fq, "-o", "x.kwig"], str(tmp_path))
    run(["validate", "-f", asm, "-r", fq, "-o", "x.bkwig"], str(tmp_path))
    with open(tmp_path / "x.kwig", "rb") as f:
        kwig = f.read()
    assert kwig == _gz("ref_runs/synthA_k21.kwig.gz")
    for exe in (os.path.join(ROOT, "kreeq_b200", "kreeq-decompressor"), os.path.join(ROOT, "oracle", "_ref", "kreeq-decompressor")):
        if os.path.exists(exe):
            p = subprocess.run([exe, "inflate", "-i", "x.bkwig"], cwd=str(tmp_path), capture_output=True, timeout=60)
            assert p.returncode == 0 and p.stdout == kwig, exe

#!/usr/bin/env python
"""Regenerates tests/golden/ from the reference tree (run in the authoring container only).

  python tests/golden/make_golden.py [/root/reference]

What it commits (all small; nothing here is reference SOURCE code):
  ref_golden.json      command + expected stdout of validateFiles/test.0..35.tst (the reference's own
                       golden vectors for count / stats / validate / union)
  inputs/*             the tiny FASTA/FASTQ/GFA(.gz) inputs those tests name (data, <= 6.6 kB each)
  fixture_dbs.tar.gz   testFiles/test1.kreeq + test2.kreeq as written by the reference (format pin)
  ref_runs.json + ref_runs/*   outputs of the UNMODIFIED reference binary (oracle/_ref/kreeq, built
                       by oracle/Makefile from /root/reference) on inputs/ and on seeded synthetic
                       inputs: stdout, .kwig, .bed, .bkwig, and the decoded database records (.npz)
"""
import gzip
import io
import json
import os
import shutil
import subprocess
import sys
import tarfile
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "kreeq")


def tst_cases():
    out = {}
    for i in range(36):
        with open(os.path.join(REF, "validateFiles", "test.%d.tst" % i)) as f:
            lines = f.read().split("\n")
        assert lines[1] == "embedded"
        out["test.%d" % i] = {"cmd": lines[0].strip(), "stdout": "\n".join(lines[2:])}
    return out


def copy_inputs(cases):
    os.makedirs(os.path.join(HERE, "inputs"), exist_ok=True)
    names = set()
    for c in cases.values():
        for tok in c["cmd"].split():
            if tok.startswith("testFiles/") and not tok.endswith(".kreeq"):
                names.add(tok[len("testFiles/"):])
    names |= {"repeat1.fasta", "repeat1.fastq", "decompressor1.fasta"}
    for n in sorted(names):
        shutil.copyfile(os.path.join(REF, "testFiles", n), os.path.join(HERE, "inputs", n))
    return sorted(names)


def pack_fixture_dbs():
    with tarfile.open(os.path.join(HERE, "fixture_dbs.tar.gz"), "w:gz") as tar:
        for d in ("test1.kreeq", "test2.kreeq"):
            tar.add(os.path.join(REF, "testFiles", d), arcname=d)


def run_ref(args, cwd):
    r = subprocess.run([REF_BIN] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, check=True)
    return r.stdout.decode()


def synth_case(seed, genome_len, n_reads, read_len, err, n_rate, contigs, gaps):
    """Seeded synthetic FASTA + FASTQ texts (numpy default_rng, SURVEY 8d generator scaled down)."""
    from helpers import random_genome, sample_reads
    rng = np.random.default_rng(seed)
    genome = random_genome(rng, genome_len)
    batch = sample_reads(rng, genome, n_reads, read_len, err=err, n_rate=n_rate)
    reads = [r for r in batch.split(b"N")] if n_rate == 0 else None
    # FASTQ text straight from the 'N'-joined batch when reads hold no N; otherwise re-split by length
    fq = io.BytesIO()
    step = read_len + 1
    for i in range(0, len(batch), step):
        s = batch[i:i + read_len]
        fq.write(b"@r%d\n" % (i // step) + s + b"\n+\n" + b"I" * len(s) + b"\n")
    asm = genome.copy()
    sub = rng.integers(0, asm.size, max(1, asm.size // 500))
    asm[sub] = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, sub.size)]
    asm = bytearray(asm.tobytes())
    for g in range(gaps):                       # N runs -> several segments per record
        at = int(rng.integers(100, len(asm) - 200)); ln = int(rng.integers(1, 40))
        asm[at:at + ln] = b"N" * ln
    fa = io.BytesIO()
    per = len(asm) // contigs
    for c in range(contigs):
        seq = bytes(asm[c * per:(c + 1) * per])
        if c == 1:
            seq = seq[:50].lower() + seq[50:]      # soft-masked bases count like upper case
        fa.write(b">contig%d some comment\n" % c)
        for o in range(0, len(seq), 60):
            fa.write(seq[o:o + 60] + b"\n")
    del reads
    return fa.getvalue(), fq.getvalue()


RUNS = [
    # name, asm source, reads source(s), k, cutoff
    ("random1_k21", "inputs/random1.fasta", ["inputs/random1.fastq"], 21, 0),
    ("random1x2_k21", "inputs/random1.fasta", ["inputs/random1.fastq", "inputs/random2.fastq"], 21, 0),
    ("random1_k15_c2", "inputs/random1.fasta", ["inputs/random2.fastq"], 15, 2),
    ("random3N_k21", "inputs/random1.fasta", ["inputs/random3.N.fastq"], 21, 0),
    ("repeat1_k21", "inputs/repeat1.fasta", ["inputs/repeat1.fastq"], 21, 0),
    ("synthA_k21", ("synth", 11, 6000, 900, 100, 0.01, 0.0, 3, 4), None, 21, 0),
    ("synthB_k31", ("synth", 12, 5000, 600, 120, 0.02, 0.003, 2, 3), None, 31, 0),
    ("synthC_k32_c3", ("synth", 13, 4000, 800, 90, 0.01, 0.0, 1, 2), None, 32, 3),
    ("synthD_k9", ("synth", 14, 3000, 300, 80, 0.01, 0.0, 2, 1), None, 9, 0),
]


def make_runs():
    import kreeq_oracle as ko
    out_dir = os.path.join(HERE, "ref_runs")
    shutil.rmtree(out_dir, ignore_errors=True)
    os.makedirs(out_dir)
    meta = {}
    for name, asm, reads, k, cutoff in RUNS:
        work = tempfile.mkdtemp(prefix="kq_golden_")
        try:
            if isinstance(asm, tuple):
                fa, fq = synth_case(*asm[1:])
                with open(os.path.join(out_dir, name + ".fasta.gz"), "wb") as f:
                    f.write(gzip.compress(fa, mtime=0))
                with open(os.path.join(out_dir, name + ".fastq.gz"), "wb") as f:
                    f.write(gzip.compress(fq, mtime=0))
                fa_path = os.path.join(work, "asm.fasta"); fq_paths = [os.path.join(work, "reads.fastq")]
                open(fa_path, "wb").write(fa); open(fq_paths[0], "wb").write(fq)
                entry = {"asm": "ref_runs/%s.fasta.gz" % name, "reads": ["ref_runs/%s.fastq.gz" % name]}
            else:
                fa_path = os.path.join(HERE, asm); fq_paths = [os.path.join(HERE, r) for r in reads]
                entry = {"asm": asm, "reads": reads}
            db = os.path.join(work, "db.kreeq")
            common = ["-k", str(k)] + (["-c", str(cutoff)] if cutoff else [])
            entry["count_stdout"] = run_ref(["validate", "-r"] + fq_paths + ["-k", str(k), "-o", db], work)
            for ext in ("kwig", "bed", "bkwig"):
                o = os.path.join(work, "out." + ext)
                # -k is passed explicitly: the reference constructs its DBG before loadGraph() reads the database's
                # k (src/input.cpp:88 vs :100), so without -k it would validate a k!=21 database with k=21
                stdout = run_ref(["validate", "-f", fa_path, "-d", db, "-o", o] + common, work)
                entry["validate_stdout"] = stdout
                data = open(o, "rb").read()
                with open(os.path.join(out_dir, "%s.%s.gz" % (name, ext)), "wb") as f:
                    f.write(gzip.compress(data, mtime=0))
            t = ko.read_db(db, check=True)
            k8, r8 = t.dump8(); k32, r32 = t.dump32()
            np.savez_compressed(os.path.join(out_dir, name + ".db.npz"), k8=k8, r8=r8, k32=k32, r32=r32)
            entry.update({"k": k, "cutoff": cutoff})
            meta[name] = entry
        finally:
            shutil.rmtree(work, ignore_errors=True)
    # union of two reference-built databases, both orders (row A4 is order dependent)
    work = tempfile.mkdtemp(prefix="kq_golden_")
    try:
        dbs = []
        for i, (fq, extra) in enumerate((("inputs/repeat1.fastq", "inputs/random1.fastq"), ("inputs/random2.fastq", None), ("inputs/repeat1.fastq", None))):
            d = os.path.join(work, "u%d.kreeq" % i)
            run_ref(["validate", "-r", os.path.join(HERE, fq)] + ([os.path.join(HERE, extra)] if extra else []) + ["-k", "21", "-o", d], work)
            dbs.append(d)
        with tarfile.open(os.path.join(out_dir, "union_inputs.tar.gz"), "w:gz") as tar:
            for d in dbs:
                tar.add(d, arcname=os.path.basename(d))
        meta["union"] = {}
        for order in ([0, 1, 2], [1, 0, 2], [1, 2, 0]):
            tag = "".join(map(str, order))
            o = os.path.join(work, "out%s.kreeq" % tag)
            stdout = run_ref(["union", "-d"] + [dbs[i] for i in order] + ["-o", o], work)
            t = ko.read_db(o, check=True)
            k8, r8 = t.dump8(); k32, r32 = t.dump32()
            np.savez_compressed(os.path.join(out_dir, "union%s.db.npz" % tag), k8=k8, r8=r8, k32=k32, r32=r32)
            vout = run_ref(["validate", "-f", os.path.join(HERE, "inputs/repeat1.fasta"), "-d", o, "-o", os.path.join(work, "v.kwig")], work)
            kw = open(os.path.join(work, "v.kwig"), "rb").read()
            meta["union"][tag] = {"order": order, "stdout": stdout, "validate_stdout": vout, "kwig": kw.decode()}
    finally:
        shutil.rmtree(work, ignore_errors=True)
    with open(os.path.join(HERE, "ref_runs.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    cases = tst_cases()
    with open(os.path.join(HERE, "ref_golden.json"), "w") as f:
        json.dump(cases, f, indent=1, sort_keys=True)
    print("inputs:", copy_inputs(cases))
    pack_fixture_dbs()
    if os.path.exists(REF_BIN):
        make_runs()
    else:
        print("oracle/_ref/kreeq missing: run `make -C oracle ref` first", file=sys.stderr)
        sys.exit(1)

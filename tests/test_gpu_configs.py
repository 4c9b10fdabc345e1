"""Scaled-down twins of BASELINE.json configs 3, 4, 5 through the command line at 1 and several shards, against the
unmodified reference binary (oracle/_ref/kreeq) on the same files: stdout, .kwig / .bkwig bytes, decoded databases.

(The full-size shapes are bench.py --config C3|C4|C5; here the point is parity on every input feature the configs
add: planted repeats past 255 instances in every shard of an 8-way union, 24 records with N-runs, 15 kb reads that
span staging chunks, k = 31 with 16-byte tuples.)"""
import os
import subprocess

import numpy as np
import pytest

import configs as cf
from helpers import assert_same_records

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OURS = os.path.join(ROOT, "kreeq_b200", "kreeq")
REF = os.path.join(ROOT, "oracle", "_ref", "kreeq")


def devs(n):
    import torch
    have = torch.cuda.device_count()
    return ["--devices", ",".join(str(g % have) for g in range(n))]


def run(exe, args, cwd):
    p = subprocess.run([exe] + args, cwd=cwd, capture_output=True)
    assert p.returncode == 0, (args, p.stderr.decode()[-400:])
    return p.stdout


need_ref = pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref/kreeq not built")


@need_ref
@pytest.mark.parametrize("G", [1, 4])
def test_config3_twin_union_of_8(tmp_path, G):
    import kreeq_oracle as ko
    rng = np.random.default_rng(30)
    g = cf.plant_repeats(rng, cf.genome(rng, 300_000), copies=16, rep_len=1000, polya_len=400)
    cf.write_fasta(tmp_path / "asm.fasta", [("c%d" % i, g[i * 100_000:(i + 1) * 100_000].tobytes()) for i in range(3)])
    for i in range(8):
        r = np.random.default_rng(30 + i)
        cf.write_fastq(tmp_path / ("s%d.fastq" % i), cf.reads_fixed(r, g, 26_000, 150, 0.01))
    out = {}
    for name, exe, extra in (("ref", REF, ["-j", "8"]), ("ours", OURS, devs(G))):
        w = tmp_path / name; w.mkdir()
        dbs = []
        for i in range(8):
            db = str(w / ("s%d.kreeq" % i)); dbs.append(db)
            run(exe, ["validate", "-r", str(tmp_path / ("s%d.fastq" % i)), "-k", "21", "-o", db] + extra, w)
        u = str(w / "u.kreeq")
        so_u = run(exe, ["union", "-d"] + dbs + ["-o", u] + extra, w)
        so_v = run(exe, ["validate", "-f", str(tmp_path / "asm.fasta"), "-d", u, "-k", "21", "-o", str(w / "u.kwig")] + extra, w)
        out[name] = (so_u, so_v, open(w / "u.kwig", "rb").read(), ko.read_db(u, check=(name == "ours")))
    assert out["ours"][0] == out["ref"][0] and out["ours"][1] == out["ref"][1]
    assert out["ours"][2] == out["ref"][2]
    assert_same_records(out["ours"][3].export(), out["ref"][3].export())
    k32a, r32a = out["ours"][3].dump32(); k32b, r32b = out["ref"][3].dump32()
    assert k32a.size > 0 and np.array_equal(k32a, k32b) and r32a.tobytes() == r32b.tobytes()


@need_ref
@pytest.mark.parametrize("G", [1, 8])
def test_config4_twin_records_with_gaps(tmp_path, G):
    import kreeq_oracle as ko
    rng = np.random.default_rng(4)
    g = cf.genome(rng, 1_200_000)
    cf.write_fasta(tmp_path / "asm.fasta", cf.records_with_gaps(rng, g, 24, 100, 200), width=60)
    reads = cf.reads_fixed(rng, g, 240_000, 150, 0.01)
    cf.write_fastq(tmp_path / "R1.fastq", reads[0::2]); cf.write_fastq(tmp_path / "R2.fastq", reads[1::2])
    out = {}
    for name, exe, extra in (("ref", REF, ["-j", "8"]), ("ours", OURS, devs(G))):
        w = tmp_path / name; w.mkdir()
        db = str(w / "db.kreeq")
        a = run(exe, ["validate", "-r", str(tmp_path / "R1.fastq"), str(tmp_path / "R2.fastq"), "-k", "21", "-o", db] + extra, w)
        b = run(exe, ["validate", "-f", str(tmp_path / "asm.fasta"), "-d", db, "-k", "21", "-o", str(w / "o.kwig")] + extra, w)
        out[name] = (a, b, open(w / "o.kwig", "rb").read(), ko.read_db(db, check=(name == "ours")).export())
    assert out["ours"][:3] == out["ref"][:3]
    assert_same_records(out["ours"][3], out["ref"][3])


@need_ref
@pytest.mark.parametrize("G", [1, 2])
def test_config5_twin_long_reads_k31(tmp_path, G):
    rng = np.random.default_rng(5)
    g = cf.genome(rng, 400_000)
    cf.write_fasta(tmp_path / "asm.fasta", [("c%d" % i, g[i * 40_000:(i + 1) * 40_000].tobytes()) for i in range(10)])
    cf.write_fastq(tmp_path / "hifi.fastq", cf.reads_long(rng, g, 800, 15_000, 2_000, 5_000, 25_000, 0.001))
    out = {}
    for name, exe, extra in (("ref", REF, ["-j", "8"]), ("ours", OURS, devs(G))):
        w = tmp_path / name; w.mkdir()
        a = run(exe, ["validate", "-f", str(tmp_path / "asm.fasta"), "-r", str(tmp_path / "hifi.fastq"), "-k", "31", "-o", str(w / "o.kwig")] + extra, w)
        b = run(exe, ["validate", "-f", str(tmp_path / "asm.fasta"), "-r", str(tmp_path / "hifi.fastq"), "-k", "31", "-o", str(w / "o.bkwig")] + extra, w)
        out[name] = (a, b, open(w / "o.kwig", "rb").read(), open(w / "o.bkwig", "rb").read())
    assert out["ours"] == out["ref"]

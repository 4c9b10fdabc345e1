"""kreeq_b200/kreeq-decompressor (the .bkwig reader; src/decompressor.cpp) against the
reference's two goldens for it (validateFiles/test.48.tst, test.49.tst), committed outputs of the unmodified
reference tool on reference-written .bkwig files, the live reference tool where oracle/_ref has it, and the
format property inflate(x.bkwig) == x.kwig for every committed reference run.  Host only."""
import gzip
import hashlib
import json
import os
import shlex
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
EXE = os.path.join(ROOT, "kreeq_b200", "kreeq-decompressor")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "kreeq-decompressor")

with open(os.path.join(GOLD, "ref_decompressor.json")) as f:
    REF = json.load(f)
with open(os.path.join(GOLD, "ref_runs.json")) as f:
    RUNS = json.load(f)


def unpack(run, tmp_path):
    f = tmp_path / "in.bkwig"
    with open(os.path.join(GOLD, "ref_runs", run + ".bkwig.gz"), "rb") as g:
        f.write_bytes(gzip.decompress(g.read()))
    return f


@pytest.mark.parametrize("name", sorted(REF["tst"]))
def test_reference_goldens(name, tmp_path):
    args = [a.replace("testFiles/", os.path.join(GOLD, "inputs") + "/") for a in shlex.split(REF["tst"][name]["cmd"])[1:]]
    p = subprocess.run([EXE] + args, cwd=str(tmp_path), capture_output=True, timeout=60)
    assert p.returncode == 0 and p.stdout.decode() == REF["tst"][name]["stdout"]


@pytest.mark.parametrize("name", sorted(REF["runs"]))
def test_committed_reference_tool_outputs(name, tmp_path):
    c = REF["runs"][name]
    unpack(c["bkwig"], tmp_path)
    if c["bed"]:
        (tmp_path / "in.bed").write_text(c["bed"])
    argv = [a.format(f="in.bkwig", bed="in.bed") for a in c["args"]]
    p = subprocess.run([EXE] + argv, cwd=str(tmp_path), capture_output=True, timeout=60)
    text = p.stdout.decode().replace(EXE, "kreeq-decompressor")
    assert p.returncode == c["rc"]
    assert len(text) == c["len"] and hashlib.sha256(text.encode()).hexdigest() == c["sha256"]
    if "stdout" in c:
        assert text == c["stdout"]
    if os.path.exists(REF_BIN):                      # and live, where the reference tool was built
        q = subprocess.run([REF_BIN] + argv, cwd=str(tmp_path), capture_output=True, timeout=60)
        assert q.stdout.decode().replace(REF_BIN, "kreeq-decompressor") == text


@pytest.mark.parametrize("run", sorted(r for r in RUNS if r != "union"))
def test_inflate_reproduces_the_kwig(run, tmp_path):
    """The reference writes the same (cov, F, B) triples as text (.kwig) and as binary (.bkwig); inflating the
    binary file must give the text file byte for byte."""
    f = unpack(run, tmp_path)
    with open(os.path.join(GOLD, "ref_runs", run + ".kwig.gz"), "rb") as g:
        want = gzip.decompress(g.read())
    p = subprocess.run([EXE, "inflate", "-i", str(f)], capture_output=True, timeout=60)
    assert p.returncode == 0 and p.stdout == want


def test_usage_and_errors(tmp_path):
    p = subprocess.run([EXE], capture_output=True)
    assert p.returncode == 0 and b"inflate" in p.stdout and b"lookup" in p.stdout
    p = subprocess.run([EXE, "inflate", "-i", str(tmp_path / "missing.bkwig")], capture_output=True)
    assert p.returncode == 1 and b"Error - file does not exist" in p.stdout
    f = unpack("synthA_k21", tmp_path)
    p = subprocess.run([EXE, "lookup", "-i", str(f), "nosuchpath:1-5"], capture_output=True)
    assert p.returncode != 0 and b"Could not find header (nosuchpath)" in p.stderr

"""The command line's per-base writers (kreeq_b200/csrc/cli/writers.cpp; reference: src/kreeq-output.cpp:138-399)
on the CPU: the (cov, F, B) rows of every committed reference run (read back from the reference-written .bkwig)
are laid over the run's assembly by the command line's own genome loader and written again as .kwig, .bed and
.bkwig -- each must equal the file the unmodified reference wrote, byte for byte.  Covers several records, gaps
(N runs), soft-masked bases, k = 9 ... 32 and 32-bit values; the GPU suite checks the same files end to end."""
import gzip
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
TOOL = os.path.join(ROOT, "kreeq_b200", "kreeq-feedcheck")

with open(os.path.join(GOLD, "ref_runs.json")) as f:
    RUNS = json.load(f)


def _gz(rel):
    with open(os.path.join(GOLD, rel), "rb") as f:
        return gzip.decompress(f.read())


@pytest.mark.parametrize("run", sorted(r for r in RUNS if r != "union"))
def test_writers_reproduce_reference_files(run, tmp_path):
    asm = RUNS[run]["asm"]
    fa = tmp_path / "asm.fasta"
    fa.write_bytes(_gz(asm) if asm.endswith(".gz") else open(os.path.join(GOLD, asm), "rb").read())
    (tmp_path / "in.bkwig").write_bytes(_gz("ref_runs/%s.bkwig.gz" % run))
    for ext in ("kwig", "bed", "bkwig"):
        out = tmp_path / ("out." + ext)
        p = subprocess.run([TOOL, "rewrite", str(fa), str(tmp_path / "in.bkwig"), str(out)], capture_output=True, timeout=120)
        assert p.returncode == 0, p.stderr
        assert out.read_bytes() == _gz("ref_runs/%s.%s.gz" % (run, ext)), ext


def test_writer_blocks_and_wide_values(tmp_path):
    """Block borders (64 Ki lines for tables, 1 Mi for .kwig/.bkwig) and values of every decimal width: the
    table writer's sliced columns against a direct per-line join of the .kwig the same rows give."""
    n, k = 300_000, 21
    for ext in ("kwig", "csvtable"):
        p = subprocess.run([TOOL, "writebench", str(tmp_path / ("w." + ext)), str(n), str(k)], capture_output=True, timeout=300)
        assert p.returncode == 0, p.stderr
    rows, name, start = {}, None, {}
    for ln in (tmp_path / "w.kwig").read_text().split("\n")[1:]:
        if ln.startswith("fixedStep"):
            name = ln.split()[1].split("=")[1]; rows[name] = []
        elif ln:
            rows[name].append(ln.split(","))
    assert sum(len(v) for v in rows.values()) == n and any(len(x[0]) == 10 for v in rows.values() for x in v)
    want = []
    for name, r in rows.items():
        for i in range(len(r)):
            win = [["0", "0", "0"]] * max(0, k - 1 - i) + r[max(0, i - k + 1):i + 1]
            want.append(",".join([name, str(i)] + [",".join(w[t] for w in win) for t in range(3)]))
    assert (tmp_path / "w.csvtable").read_text() == "\n".join(want) + "\n"

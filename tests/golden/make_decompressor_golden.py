#!/usr/bin/env python
"""Regenerates tests/golden/ref_decompressor.json: stdout of the UNMODIFIED reference .bkwig reader
(oracle/_ref/kreeq-decompressor, built by oracle/Makefile from /root/reference/src/decompressor.cpp) on the
committed reference-written .bkwig files, for the option combinations tests/test_decompressor.py replays.
Also copies the reference's two decompressor goldens (validateFiles/test.48.tst, test.49.tst) with their inputs.
Run in the authoring container only."""
import gzip
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
BIN = os.path.join(ROOT, "oracle", "_ref", "kreeq-decompressor")

CASES = [   # name, bkwig, arguments ({f} = the inflated .bkwig, {bed} = coordinate file written from `bed`)
    ("synthA_inflate", "synthA_k21", ["inflate", "-i", "{f}"], None),
    ("synthA_inflate_expand", "synthA_k21", ["inflate", "-i", "{f}", "--expand"], None),
    ("synthB_lookup_positional", "synthB_k31", ["lookup", "-i", "{f}", "contig0:100-160"], None),
    ("synthB_lookup_span", "synthB_k31", ["lookup", "-i", "{f}", "contig1:300-310", "-s", "5"], None),
    ("synthA_lookup_bed", "synthA_k21", ["lookup", "-i", "{f}", "-c", "{bed}"], "contig0\t5\t40\ncontig2\t100\t130\ncontig0\t900\t905\n"),
    ("synthA_lookup_expand", "synthA_k21", ["lookup", "-i", "{f}", "-c", "{bed}", "--expand"], "contig1\t200\t215\ncontig0\t500\t506\n"),
    ("synthD_lookup_to_component_end", "synthD_k9", ["lookup", "-i", "{f}", "contig0:10-100000"], None),
    ("repeat1_inflate", "repeat1_k21", ["inflate", "-i", "{f}"], None),
    ("random3N_inflate_cmd", "random3N_k21", ["inflate", "-i", "{f}", "--cmd"], None),
]


def main():
    out = {"tst": {}, "runs": {}}
    os.makedirs(os.path.join(HERE, "inputs"), exist_ok=True)
    for t in (48, 49):
        with open(os.path.join(REF, "validateFiles", "test.%d.tst" % t)) as f:
            lines = f.read().split("\n")
        out["tst"]["test.%d" % t] = {"cmd": lines[0].strip(), "stdout": "\n".join(lines[2:])}
        for tok in lines[0].split():
            if tok.startswith("testFiles/"):
                shutil.copyfile(os.path.join(REF, tok), os.path.join(HERE, "inputs", os.path.basename(tok)))
    with tempfile.TemporaryDirectory() as tmp:
        for name, run, args, bed in CASES:
            f = os.path.join(tmp, "in.bkwig")          # fixed name: --cmd echoes the path
            with open(os.path.join(HERE, "ref_runs", run + ".bkwig.gz"), "rb") as g, open(f, "wb") as o:
                o.write(gzip.decompress(g.read()))
            b = os.path.join(tmp, "in.bed")
            if bed:
                with open(b, "w") as o:
                    o.write(bed)
            argv = [a.format(f="in.bkwig", bed="in.bed") for a in args]
            p = subprocess.run([BIN] + argv, cwd=tmp, capture_output=True)
            text = p.stdout.decode().replace(BIN, "kreeq-decompressor")
            rec = {"bkwig": run, "args": args, "bed": bed, "rc": p.returncode, "len": len(text),
                   "sha256": hashlib.sha256(text.encode()).hexdigest()}
            if len(text) <= 20000:                     # long outputs are pinned by digest only
                rec["stdout"] = text
            out["runs"][name] = rec
    with open(os.path.join(HERE, "ref_decompressor.json"), "w") as f:
        json.dump(out, f, indent=1)
    print({k: (v["rc"], v["len"]) for k, v in out["runs"].items()})


if __name__ == "__main__":
    main()

"""User-level comparison on one box: the same `kreeq validate -f asm.fasta -r R1.fastq R2.fastq -k 21 -o out.kwig`
command through kreeq_b200/kreeq (GPU) and through the unmodified reference (oracle/_ref/kreeq, all host cores),
on a seeded synthetic data set (default 10 Mbp genome, 30x, 150 bp pairs, 1 % error).  Checks that stdout and the
.kwig file are byte-identical and prints one JSON line with the wall-clock times."""
import argparse
import filecmp
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (FASTQ / FASTA writers)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome-mbp", type=float, default=10.0)
    ap.add_argument("--coverage", type=float, default=30.0)
    ap.add_argument("--gz", action="store_true", help="gzip the read files (single-member .gz, as sequencers write them)")
    ap.add_argument("--verbose", action="store_true", help="pass --verbose to our binary and keep its phase log")
    a = ap.parse_args()
    import torch
    from kreeq_b200 import synth
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    L, K = bench.READ_LEN, bench.K
    gsize = int(a.genome_mbp * 1e6)
    pairs = int(gsize * a.coverage / (2 * L))
    work = tempfile.mkdtemp(prefix="kq_cli_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        genome = synth.random_genome_codes(gsize, 2, dev)
        reads = synth.paired_reads_ascii(genome, pairs, L, bench.ERR, 2).cpu().numpy().reshape(2 * pairs, L + 1)
        r1, r2 = os.path.join(work, "R1.fastq"), os.path.join(work, "R2.fastq")
        bench.write_fastq(r1, reads[0::2].reshape(-1), pairs, L)
        bench.write_fastq(r2, reads[1::2].reshape(-1), pairs, L)
        fa = os.path.join(work, "asm.fasta")
        bench.write_fasta(fa, synth.codes_to_ascii(genome).cpu().numpy())
        del reads, genome
        if a.gz:
            for f in (r1, r2):
                subprocess.check_call(["gzip", "-1", f])
            r1, r2 = r1 + ".gz", r2 + ".gz"
        out = {"genome_bp": gsize, "read_pairs": pairs, "k": K, "gz": a.gz, "fastq_bytes": os.path.getsize(r1) + os.path.getsize(r2),
               "cores": os.cpu_count()}
        res = {}
        for name, exe, extra in (("ours", os.path.join(ROOT, "kreeq_b200", "kreeq"), []),
                                 ("reference", os.path.join(ROOT, "oracle", "_ref", "kreeq"), ["-j", str(os.cpu_count() or 1)])):
            if not os.path.exists(exe):
                continue
            cwd = os.path.join(work, name); os.mkdir(cwd)
            kw = os.path.join(cwd, "out.kwig")
            t = time.perf_counter()
            if name == "ours" and a.verbose:
                extra = extra + ["--verbose"]
            p = subprocess.run([exe, "validate", "-f", fa, "-r", r1, r2, "-k", str(K), "-o", kw] + extra, cwd=cwd,
                               stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            out[name + "_s"] = time.perf_counter() - t
            if p.returncode != 0:
                raise RuntimeError(name + " failed: " + p.stderr.decode()[-500:])
            res[name] = (p.stdout, kw)
            if name == "ours" and a.gz:                  # same run with the gz files decoded by zlib, one thread per file
                t = time.perf_counter()
                q = subprocess.run([exe, "validate", "-f", fa, "-r", r1, r2, "-k", str(K), "-o", os.path.join(cwd, "zlib.kwig")], cwd=cwd,
                                   stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=dict(os.environ, KQ_CLI_PGZ_MIN="999999999999"))
                out["ours_zlib_gz_s"] = time.perf_counter() - t
                out["zlib_gz_identical"] = q.returncode == 0 and q.stdout == p.stdout and filecmp.cmp(kw, os.path.join(cwd, "zlib.kwig"), shallow=False)
            if name == "ours" and a.verbose:
                out["phases"] = p.stderr.decode().splitlines()
        if len(res) == 2:
            out["stdout_identical"] = res["ours"][0] == res["reference"][0]
            out["kwig_identical"] = filecmp.cmp(res["ours"][1], res["reference"][1], shallow=False)
            out["kwig_bytes"] = os.path.getsize(res["ours"][1])
            out["speedup"] = out["reference_s"] / out["ours_s"]
        out["stdout"] = res["ours"][0].decode()
        print(json.dumps(out))
    finally:
        shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    main()

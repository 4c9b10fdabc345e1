#!/usr/bin/env python
"""bench.py -- k-mers counted/s and assembly k-mers validated/s of the kreeq hot path on N B200s.

Workload (BASELINE.json configs[1], "C2"): synthetic 100 Mbp genome, 30x 150 bp paired reads with
1 % substitution error, k=21: count, then validate the genome.  One step = one whole count+validate
pass over that input.  At N>1 every rank extracts its own 30x read set (weak scaling), the key space
is sharded (key % 128) / (128/N) and tuples are routed with an all-to-all per round.

  value : k-mer instances counted / s, inputs (2-bit packed reads) already resident in HBM
  e2e   : the same through the C-ABI with HOST ASCII buffers (pack + H2D inside the timed region),
          validate results copied back to host memory
  --impl reference : the unmodified reference binary (oracle/_ref/kreeq) on the host cores, bounded sample
"""
import argparse
import ctypes as C
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

K = 21
READ_LEN = 150
ERR = 0.01
ALG_BYTES_COUNT = 26.25      # 0.25 in + 8 key + 9 record read + 9 record write
ALG_BYTES_VALIDATE = 30.25   # 0.25 in + 17 slot + 13 out
SECTOR_BYTES_COUNT = 64.25


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--genome-mbp", type=float, default=100.0, help="C2 genome size (default 100 Mbp)")
    ap.add_argument("--coverage", type=float, default=30.0)
    ap.add_argument("--k", type=int, default=21, help="k-mer length (default: C2's 21; 29..32 use 16-byte tuples)")
    ap.add_argument("--config", default="C2", choices=["C2", "C4", "C5"], help="BASELINE.json config (C3, the 8-way union, is tools/config3_union.py)")
    ap.add_argument("--scale", type=float, default=0, help="divide C4 / C5 by this (default: 16 = the /16 twins; 1 = full size)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cli-e2e", action="store_true", help="skip the command-line run on full-size files")
    ap.add_argument("--no-parity", action="store_true", help="skip the seeded parity problem that runs before the timed region")
    ap.add_argument("--pageable-out", action="store_true", help="e2e validate into pageable host memory (staged copy)")
    ap.add_argument("--no-union", action="store_true", help="skip the database rows (save / union / load timing)")
    ap.add_argument("--ref-sample-reads", type=int, default=1_500_000)
    a = ap.parse_args()
    global K
    if a.config == "C5" and a.k == 21:
        a.k = 31
    if not a.scale:
        a.scale = 16.0 if a.config in ("C4", "C5") else 1.0
    if a.config != "C2":                               # the side measurements below are C2's
        a.no_cpu_baseline = a.no_cli_e2e = a.no_union = a.no_parity = True
    K = a.k
    return a


def workload(args):
    """BASELINE.json configs (SURVEY 8d definitions): C2 (default, the configuration the metric is quoted on), C4 =
    3.1 Gbp assembly in 24 records with 100 N-runs vs a 30x 150 bp read set, C5 = 1 Gbp, 30x 15 kb reads, k = 31;
    --scale S divides the genome and the read set by S (C4 and C5 default to their /16 twins on one GPU)."""
    if args.config == "C5":
        genome = int(1e9 / args.scale)
        n = int(genome * 30 / 15000)
        return {"workload": "C5/%g: synthetic %.4g Mbp genome, 30x HiFi-like reads (15 kb +- 2 kb, 0.1%% error), k=%d: count then validate"
                            % (args.scale, genome / 1e6, K), "genome_bp": genome, "reads_per_gpu": n, "k": K, "scale": args.scale}
    if args.config == "C4":
        genome = int(3.1e9 / args.scale)
        pairs = int(genome * 30 / (2 * READ_LEN))
        return {"workload": "C4/%g: synthetic %.4g Mbp assembly in 24 records with 100 N-runs, 30x %d bp paired reads, 1%% error, k=%d: count then validate"
                            % (args.scale, genome / 1e6, READ_LEN, K), "genome_bp": genome, "read_pairs_per_gpu": pairs, "k": K, "scale": args.scale}
    genome = int(args.genome_mbp * 1e6)
    pairs = int(genome * args.coverage / (2 * READ_LEN))
    return {"workload": "C2: synthetic %g Mbp genome, %gx %d bp paired reads, %g%% error, k=%d: count then validate"
                        % (args.genome_mbp, args.coverage, READ_LEN, ERR * 100, K),
            "genome_bp": genome, "read_pairs_per_gpu": pairs, "k": K}


HUMAN_MBP = [248, 242, 198, 190, 182, 171, 159, 145, 138, 134, 135, 133, 114, 107, 102, 90, 83, 80, 59, 64, 47, 51, 156, 57]


def make_inputs(args, cfg, rank, dev):
    """-> (reads: pinned uint8 'N'-joined batch, segs: list of pinned uint8 ASCII segments (no N inside), k-mer instances,
    estimated distinct k-mers, push offsets (read boundaries ~97 MB apart))"""
    import numpy as np
    import torch
    from kreeq_b200 import synth
    gsize = cfg["genome_bp"]
    genome = synth.random_genome_codes(gsize, 2 + rank, dev)

    def pinned(t):
        h = torch.empty(t.numel(), dtype=torch.uint8, pin_memory=True)
        h.copy_(t)
        return h

    if args.config == "C5":
        n, per = cfg["reads_per_gpu"], 2000
        parts, lens = [], []
        for o in range(0, n, per):
            r = synth.long_reads_ascii(genome, min(per, n - o), 15000, 2000, 5000, 25000, 0.001, 5 + 1000 * rank + o)
            parts.append(r.cpu())
        reads_host = torch.empty(sum(p.numel() for p in parts), dtype=torch.uint8, pin_memory=True)
        torch.cat(parts, out=reads_host)
        del parts
        reads_np = reads_host.numpy()
        sep = np.flatnonzero(reads_np == 78)
        lens = np.diff(np.concatenate(([-1], sep))) - 1
        inst = int(np.maximum(lens - K + 1, 0).sum())
        est = int((gsize + reads_np.size * 0.001 * K * 0.85) * 1.05)
        step = 97 << 20
        cuts = [0] + [int(sep[min(sep.size - 1, np.searchsorted(sep, o))]) + 1 for o in range(step, reads_np.size, step)] + [reads_np.size]
        cuts = sorted(set(cuts))
    else:
        pairs = cfg["read_pairs_per_gpu"]
        reads_host = pinned(synth.paired_reads_ascii(genome, pairs, READ_LEN, ERR, 2 + 1000 * rank))
        reads_np = reads_host.numpy()
        inst = 2 * pairs * (READ_LEN - K + 1)
        est = int((gsize + 2 * pairs * READ_LEN * ERR * K * 0.85) * 1.05)
        B = 64 * (READ_LEN + 1) * 10000                       # ~97 MB pushes, cut at read boundaries
        cuts = list(range(0, reads_np.size, B)) + [reads_np.size]
    asm_host = pinned(synth.codes_to_ascii(genome))
    del genome
    torch.cuda.empty_cache()
    asm_np = asm_host.numpy()
    if args.config == "C4":
        # 24 records with sizes proportional to the human chromosomes; 100 runs of N (50 000 / scale bases) cut the
        # records into segments (gfalibs/src/gfa.cpp:170-203); the runs themselves never reach the device
        rng = np.random.default_rng(4 + rank)
        w = np.array(HUMAN_MBP, dtype=np.float64)
        ends = np.round(np.cumsum(w / w.sum()) * gsize).astype(np.int64)
        gaps = np.sort(rng.choice(gsize - 2000, 100, replace=False) + 1000)
        bounds = np.unique(np.concatenate(([0], ends, gaps)))
        glen = max(1, int(50000 / args.scale))
        segs = []
        for a, b in zip(bounds[:-1], bounds[1:]):
            a2 = a + (glen if a in gaps else 0)               # the run of N replaces the first bases after the cut
            if b - a2 > 0:
                segs.append(asm_np[a2:b])
    else:
        n_contigs = 10
        contig = gsize // n_contigs
        segs = [asm_np[i * contig:(i + 1) * contig] for i in range(n_contigs)]
    return reads_host, segs, asm_host, inst, est, cuts


# ---- clocks ------------------------------------------------------------------------------------------

class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx = float(r[2])
            except (ValueError, IndexError):
                continue
            for name, col in (("hw_slowdown", 4), ("hw_thermal_slowdown", 5), ("sw_thermal_slowdown", 6), ("sw_power_cap", 7)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---- reference arm (CPU) --------------------------------------------------------------------------------

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "kreeq")


def write_fastq(path, ascii_n_joined, n_reads, read_len):
    """'N'-joined uint8 array [n_reads*(L+1)] -> 4-line FASTQ (numpy, no Python loop)."""
    import numpy as np
    a = np.asarray(ascii_n_joined, dtype=np.uint8).reshape(n_reads, read_len + 1)[:, :read_len]
    row = np.empty((n_reads, 3 + read_len + 3 + read_len + 1), dtype=np.uint8)
    row[:, 0:3] = np.frombuffer(b"@r\n", dtype=np.uint8)
    row[:, 3:3 + read_len] = a
    row[:, 3 + read_len:6 + read_len] = np.frombuffer(b"\n+\n", dtype=np.uint8)
    row[:, 6 + read_len:6 + 2 * read_len] = ord("I")
    row[:, -1] = ord("\n")
    row.tofile(path)


def write_fasta(path, genome_ascii, name="chr"):
    import numpy as np
    with open(path, "wb") as f:
        f.write((">%s\n" % name).encode())
        np.asarray(genome_ascii, dtype=np.uint8).tofile(f)
        f.write(b"\n")


def reference_sample(args, workdir):
    """A bounded sample of C2 for the CPU arms: same genome generator, fewer reads, 10 Mbp of assembly."""
    import torch
    from kreeq_b200 import synth
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    gsize = int(args.genome_mbp * 1e6)
    genome = synth.random_genome_codes(gsize, 2, dev)
    n_pairs = args.ref_sample_reads // 2
    reads = synth.paired_reads_ascii(genome, n_pairs, READ_LEN, ERR, 2).cpu().numpy()
    asm = synth.codes_to_ascii(genome[:min(gsize, 10_000_000)]).cpu().numpy()
    fq = os.path.join(workdir, "sample.fastq"); fa = os.path.join(workdir, "asm.fasta")
    write_fastq(fq, reads, 2 * n_pairs, READ_LEN)
    write_fasta(fa, asm)
    return fq, fa, 2 * n_pairs * (READ_LEN - K + 1), asm.size - K + 1


def run_reference_once(fq, fa, workdir, threads):
    """count: `kreeq validate -r sample.fastq -k 21 -o db.kreeq`; validate: `-f asm.fasta -d db.kreeq -o out.kwig`
    (BASELINE.md section 3), each timed by wall clock, run in a scratch CWD."""
    db = os.path.join(workdir, "db.kreeq")
    shutil.rmtree(db, ignore_errors=True)
    t0 = time.perf_counter()
    out = subprocess.run([REF_BIN, "validate", "-r", fq, "-k", str(K), "-o", db, "-j", str(threads)],
                         cwd=workdir, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout.decode()
    t1 = time.perf_counter()
    subprocess.check_call([REF_BIN, "validate", "-f", fa, "-d", db, "-k", str(K), "-o", os.path.join(workdir, "out.kwig"), "-j", str(threads)],
                          cwd=workdir, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    t2 = time.perf_counter()
    distinct = [int(ln.split(":")[1]) for ln in out.splitlines() if ln.startswith("Distinct kmers")]
    return t1 - t0, t2 - t1, (distinct[0] if distinct else 0)


def run_reference_union(workdir, threads):
    """`kreeq union -d db.kreeq db.kreeq -o u.kreeq`: the sample database merged with itself (row A4)."""
    db = os.path.join(workdir, "db.kreeq"); u = os.path.join(workdir, "u.kreeq")
    shutil.rmtree(u, ignore_errors=True)
    t0 = time.perf_counter()
    subprocess.check_call([REF_BIN, "union", "-d", db, db, "-o", u, "-j", str(threads)],
                          cwd=workdir, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    dt = time.perf_counter() - t0
    shutil.rmtree(u, ignore_errors=True)
    return dt


def cpu_baseline(args, steps, warmup):
    """Times the unmodified reference on the host cores on a bounded sample of the workload."""
    if not os.path.exists(REF_BIN):
        return None
    threads = os.cpu_count() or 1
    base = "/dev/shm" if os.path.isdir("/dev/shm") else None
    workdir = tempfile.mkdtemp(prefix="kreeq_ref_", dir=base)
    try:
        fq, fa, inst, asm_k = reference_sample(args, workdir)
        tc, tv, distinct = [], [], 0
        for i in range(warmup + steps):
            c, v, distinct = run_reference_once(fq, fa, workdir, threads)
            if i >= warmup:
                tc.append(c); tv.append(v)
        c = sum(tc) / len(tc); v = sum(tv) / len(tv)
        union = {}
        if not getattr(args, "no_union", False) and distinct:
            us = run_reference_union(workdir, threads)
            union = {"union_value": 2 * distinct / us, "union_unit": "input records/s", "union_s": us, "union_records_in": 2 * distinct}
        return {**union, "value": inst / c, "unit": "k-mers/s", "cores": threads, "kind": "reference",
                "sample": "%d reads x %d bp of the same C2 generator (%d k-mer instances) counted by `kreeq validate -r -o db.kreeq -j %d`; "
                          "validate: first %d bp of the genome vs that DB" % (args.ref_sample_reads // 2 * 2, READ_LEN, inst, threads, asm_k + K - 1),
                "count_s": c, "validate_value": asm_k / v, "validate_s": v}
    finally:
        shutil.rmtree(workdir, ignore_errors=True)


def scratch_dir(need_bytes):
    """/dev/shm when it (and RAM) can take `need_bytes`, else the default temp dir on disk"""
    try:
        import psutil
        if os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > need_bytes * 1.3 and psutil.virtual_memory().available > need_bytes * 1.5:
            return "/dev/shm"
    except Exception:
        pass
    return None


def make_c2_files(args, workdir):
    """The FULL workload as files: R1.fastq / R2.fastq (rank 0's read set of the GPU arm: genome seed 2, read seed 2)
    and asm.fasta with the genome in 10 records.  Both arms' cli_e2e blocks run on exactly these files."""
    import torch
    from kreeq_b200 import synth
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    gsize = int(args.genome_mbp * 1e6)
    pairs = int(gsize * args.coverage / (2 * READ_LEN))
    genome = synth.random_genome_codes(gsize, 2, dev)
    r1, r2, fa = (os.path.join(workdir, n) for n in ("R1.fastq", "R2.fastq", "asm.fasta"))
    step = 1 << 21                                   # pairs per generated chunk (bounded host memory)
    # the generator draws pair chunks from ONE seeded stream, so generate in one call and write in slices
    reads = synth.paired_reads_ascii(genome, pairs, READ_LEN, ERR, 2).cpu().numpy().reshape(2 * pairs, READ_LEN + 1)
    for name, sel in ((r1, reads[0::2]), (r2, reads[1::2])):
        with open(name, "wb") as f:
            for o in range(0, pairs, step):
                n = min(step, pairs - o)
                tmp = name + ".part"
                write_fastq(tmp, np_contig(sel[o:o + n]).reshape(-1), n, READ_LEN)
                with open(tmp, "rb") as g:
                    shutil.copyfileobj(g, f, 64 << 20)
                os.remove(tmp)
    del reads
    asm = synth.codes_to_ascii(genome).cpu().numpy()
    n_contigs = 10
    contig = gsize // n_contigs
    with open(fa, "wb") as f:
        for i in range(n_contigs):
            f.write((">chr%d\n" % (i + 1)).encode())
            asm[i * contig:(i + 1) * contig].tofile(f)
            f.write(b"\n")
    inst = 2 * pairs * (READ_LEN - K + 1)
    return r1, r2, fa, inst, n_contigs * (contig - K + 1)


def np_contig(a):
    import numpy as np
    return np.ascontiguousarray(a)


def file_md5(path):
    import hashlib
    h = hashlib.md5()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(64 << 20), b""):
            h.update(blk)
    return h.hexdigest()


def run_cli_pair(exe, r1, r2, fa, workdir, extra):
    """The two commands of BASELINE.md section 3 through one binary, wall clock each:
       count    : kreeq validate -r R1.fastq R2.fastq -k K -o db.kreeq
       validate : kreeq validate -f asm.fasta -d db.kreeq -k K -o out.kwig"""
    import hashlib
    db = os.path.join(workdir, "db.kreeq"); kw = os.path.join(workdir, "out.kwig")
    shutil.rmtree(db, ignore_errors=True)
    t0 = time.perf_counter()
    a = subprocess.run([exe, "validate", "-r", r1, r2, "-k", str(K), "-o", db] + extra, cwd=workdir, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    t1 = time.perf_counter()
    if a.returncode != 0:
        raise RuntimeError("count failed: " + a.stderr.decode()[-300:])
    b = subprocess.run([exe, "validate", "-f", fa, "-d", db, "-k", str(K), "-o", kw] + extra, cwd=workdir, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    t2 = time.perf_counter()
    if b.returncode != 0:
        raise RuntimeError("validate failed: " + b.stderr.decode()[-300:])
    out = {"count_s": t1 - t0, "validate_s": t2 - t1, "kwig_md5": file_md5(kw), "kwig_bytes": os.path.getsize(kw),
           "stdout_md5": hashlib.md5(a.stdout + b.stdout).hexdigest(), "qv_rows": b.stdout.decode().splitlines()[-2:]}
    shutil.rmtree(db, ignore_errors=True); os.remove(kw)
    return out


def cli_e2e(args, which):
    """Same files, same two command lines, same scope (FASTQ parse -> database on disk; FASTA parse + database load ->
    .kwig on disk) for either binary.  which: "ours" (kreeq_b200/kreeq) or "reference" (oracle/_ref/kreeq, all cores)."""
    exe = os.path.join(ROOT, "kreeq_b200", "kreeq") if which == "ours" else REF_BIN
    if not os.path.exists(exe):
        return {"unavailable": exe + " not built"}
    gsize = int(args.genome_mbp * 1e6)
    need = gsize * args.coverage * 2.1 + (gsize * 6 * 50 if which == "ours" else gsize * 6 * 50 + gsize * args.coverage * 9)   # FASTQ + DB (+ the reference's .buf spill)
    work = tempfile.mkdtemp(prefix="kq_cli_", dir=scratch_dir(need))
    try:
        r1, r2, fa, inst, asm_k = make_c2_files(args, work)
        run = os.path.join(work, "run"); os.mkdir(run)
        extra = [] if which == "ours" else ["-j", str(os.cpu_count() or 1)]
        r = run_cli_pair(exe, r1, r2, fa, run, extra)
        r.update({"binary": os.path.relpath(exe, ROOT), "files": "R1.fastq + R2.fastq (%d bytes), asm.fasta (%d bp, 10 records)" % (os.path.getsize(r1) + os.path.getsize(r2), gsize),
                  "scratch": work.rsplit("/", 1)[0], "kmer_instances": inst, "assembly_kmers": asm_k,
                  "count_value": inst / r["count_s"], "validate_value": asm_k / r["validate_s"], "unit": "k-mers/s", "cores": os.cpu_count()})
        return r
    finally:
        shutil.rmtree(work, ignore_errors=True)


def main_reference(args):
    """The UNMODIFIED reference binary on the FULL workload files, all host cores, run once (each run is minutes of
    CPU work: the steps / warmup the driver asks for are capped to 1 / 0 and reported as such)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = workload(args)
    if not os.path.exists(REF_BIN):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/kreeq not built (run __graft_entry__.build() where /root/reference is mounted)"}))
        return
    r = cli_e2e(args, "reference")
    cb = {"value": r["count_value"], "unit": "k-mers/s", "cores": r["cores"], "kind": "reference",
          "sample": "the full workload, once: `kreeq validate -r R1.fastq R2.fastq -k %d -o db.kreeq -j %d` (%d k-mer instances), then "
                    "`kreeq validate -f asm.fasta -d db.kreeq -o out.kwig` (%d assembly k-mers); wall clock of each process" % (K, r["cores"], r["kmer_instances"], r["assembly_kmers"]),
          "count_s": r["count_s"], "validate_value": r["validate_value"], "validate_s": r["validate_s"]}
    line = {"impl": "reference", "metric": "kmers_counted_per_sec", "value": r["count_value"], "unit": "k-mers/s",
            "n_gpus": args.gpus, "steps": 1, "warmup": 0, "steps_requested": args.steps, "warmup_requested": args.warmup,
            "ms_per_step": 1e3 * (r["count_s"] + r["validate_s"]),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": dict(cfg, scope="whole command lines on files: FASTQ parse -> .kreeq database written; FASTA parse + database load -> .kwig written (compare with the GPU arm's cli_e2e block, same files and commands)"),
            "cpu_baseline": cb, "cli_e2e": r,
            "validate": {"metric": "assembly_kmers_validated_per_sec", "value": r["validate_value"], "unit": "k-mers/s"},
            "e2e": {"value": r["count_value"], "unit": "k-mers/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ---- our arm (GPU) ------------------------------------------------------------------------------------------

def main_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import kreeq_b200
    from kreeq_b200 import shard, synth
    from kreeq_b200._lib import check, lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:                                    # the ranks share the host cores (packing, file codec)
        os.environ.setdefault("KQ_THREADS", str(max(1, (os.cpu_count() or 1) // world)))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = lib()
    cfg = workload(args)
    gsize = cfg["genome_bp"]

    # -- synthetic input: weak scaling = N copies of the problem: rank r brings its own genome (seed
    #    2 + r) and the 30x read set sampled from it, and validates that genome against the shared table --
    reads_host, segs, asm_host, inst_per_rank, est_distinct, push_cuts = make_inputs(args, cfg, rank, dev)
    reads_np = reads_host.numpy()
    contig = max(s.size for s in segs)                 # largest segment: sizes the row buffers

    def pack_to_device(a):
        words = L.kq_packed_words(a.size)
        b = torch.empty(words, dtype=torch.int64, pin_memory=True)
        m = torch.empty(words, dtype=torch.int64, pin_memory=True)
        check(L.kq_pack_host(C.c_void_p(a.ctypes.data), a.size, C.c_void_p(b.data_ptr()), C.c_void_p(m.data_ptr()), 0))
        return b.to(dev), m.to(dev)

    d_rb, d_rm = pack_to_device(reads_np)
    d_asm = [pack_to_device(s) + (s.size,) for s in segs]
    asm_k = sum(max(0, s.size - K + 1) for s in segs)
    d_out = torch.empty(contig * 4, dtype=torch.int32, device=dev)          # kq_base[contig]
    d_acc = torch.zeros(3, dtype=torch.int64, device=dev)
    # e2e validate results land in page-locked host memory (the library copies rows into it directly);
    # --pageable-out measures the staged path a caller with ordinary memory gets
    if args.pageable_out:
        h_out = np.zeros(contig, dtype=kreeq_b200.BASE)
    else:
        h_out_t = torch.empty(contig * kreeq_b200.BASE.itemsize, dtype=torch.uint8, pin_memory=True)
        h_out = h_out_t.numpy().view(kreeq_b200.BASE)

    # (est_distinct: genome k-mers + ~k novel k-mers per error; every shard gets an even 1/N share of N such sets)
    g = kreeq_b200.DBG(k=K, device=local, expected_distinct=est_distinct, rank=rank, nranks=world)
    # a real (non-default) stream: kq_set_stream(NULL) would mean "the library's own stream", and the
    # CUDA events below must be recorded on the stream the kernels are launched on
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    g.set_stream(stream.cuda_stream)
    # multi-GPU count: tuples go to their owner inside the scatter kernel (peer stores over NVLink);
    # KQ_TRANSPORT=nccl selects the staged all-to-all (extract -> ncclSend/Recv -> second scatter) instead
    transport = os.environ.get("KQ_TRANSPORT", "peer") if world > 1 else None
    if transport == "peer":
        # one round per step when the slab fits (every round sweeps the whole shard through the insert kernel)
        router = shard.PeerRoutedCounter(g, rank, world, int(os.environ.get("KQ_ROUND_POS", min(reads_np.size + 1024, (1 << 32) - (1 << 24)))), dev, stream=stream)
    elif transport == "nccl":
        router = shard.RoutedCounter(g, rank, world, 512 << 20, dev)
    else:
        router = None
    # multi-GPU validate probes the owners' shards over peer memory; the peers' tables are mapped once (first
    # warm-up step, reported as peer_connect_ms) and stay mapped: the steady-state steps refill tables that
    # do not move
    vmode = os.environ.get("KQ_VALIDATE", "routed") if world > 1 else None     # "peer": probe the owners' tables over NVLink loads
    peers = shard.PeerTables(g, world, dev, keep=True) if vmode == "peer" else None
    # routed validate: up to 2^26 assembly positions per round (inboxes of nranks x that many 16-byte queries per rank)
    v_chunk = max(contig, min(sum(s.size for s in segs), 1 << 26))
    router_v = shard.RoutedValidator(g, rank, world, v_chunk, dev) if vmode == "routed" else None
    v_rounds = router_v.rounds_of(d_asm) if router_v is not None else []
    v_bufs = []
    if router_v is not None and not args.no_e2e:
        words = L.kq_packed_words(contig)
        for _ in range(2):
            v_bufs.append((torch.empty(words, dtype=torch.int64, pin_memory=True), torch.empty(words, dtype=torch.int64, pin_memory=True),
                           torch.empty(words, dtype=torch.int64, device=dev), torch.empty(words, dtype=torch.int64, device=dev), torch.cuda.Event()))
    e2e_bufs = []
    if world > 1 and not args.no_e2e:
        words = L.kq_packed_words(5 * 64 * (READ_LEN + 1) * 10000)
        for _ in range(2):
            e2e_bufs.append((torch.empty(words, dtype=torch.int64, pin_memory=True), torch.empty(words, dtype=torch.int64, pin_memory=True),
                             None, None, torch.cuda.Event()))
        wall = L.kq_packed_words(reads_np.size)
        e2e_all = (torch.zeros(wall, dtype=torch.int64, device=dev), torch.full((wall,), -1, dtype=torch.int64, device=dev))

    def count_resident():
        total = reads_np.size - K + 1
        if world > 1:
            router.count(d_rb, d_rm, reads_np.size)
            return
        # one call: the pipeline plans its own batches (one sweep over the table for the whole read set)
        check(L.kq_dev_count_packed(g.handle, C.c_void_p(d_rb.data_ptr()), C.c_void_p(d_rm.data_ptr()),
                                    reads_np.size, 0, total), g.handle)

    def validate_contigs():
        d_acc.zero_()
        for b, m, n_ in d_asm:
            check(L.kq_dev_validate_packed(g.handle, C.c_void_p(b.data_ptr()), C.c_void_p(m.data_ptr()), n_, 0,
                                           C.c_void_p(d_out.data_ptr()), C.c_void_p(d_acc.data_ptr())), g.handle)

    def validate_resident():
        if world == 1:
            validate_contigs()
            return
        if router_v is not None:
            # every rank validates its own genome: queries are routed to the owners, rows come back (two stream
            # barriers per contig); only the three QV counters are reduced
            g.finalize()
            for rnd in v_rounds:                       # several contigs per round: one pair of barriers for all of them
                router_v.round_packed(rnd)
            router_v.counters()
            return
        # KQ_VALIDATE=peer: each k-mer is probed in its owner's shard over peer memory
        with peers:
            validate_contigs()
            dist.all_reduce(d_acc)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def parity_check():
        """A seeded 2 Mbp / 30x problem (5.2e7 instances) counted and validated through the SAME table, routing
        and kernels the timed steps use, compared with constants the CPU checker produced
        (tests/golden/bench_parity.json).  Every rank feeds every world-th read; contigs are dealt to ranks."""
        from kreeq_b200 import paritycheck as pc
        reads, asm, kk = pc.problem("bench")
        want = pc.expected("bench")
        if kk != K:
            return {"ok": None, "skipped": "parity constants are for k=%d" % kk}
        g.clear()
        L.kq_set_tuning(b"direct_pct", 0)            # small batch, big table: still take the sweep (the timed path)
        mine = np.ascontiguousarray(reads.reshape(-1, 151)[rank::world]).reshape(-1)
        if world > 1:
            b, m = pack_to_device(mine)
            router.count(b, m, mine.size)
            g.finalize()
        else:
            g.traverseInReads(mine)
            g.finalize()
        L.kq_set_tuning(b"direct_pct", 25)
        keys, recs = g.export()
        got = pc.db_digest(keys, recs) + [g.instances()]
        clen = asm.size // want["contigs"]
        rows, acc = 0, kreeq_b200._lib.KqQV()
        out = np.zeros(clen, dtype=kreeq_b200.BASE)

        def vloop():
            nonlocal rows
            for i in range(rank, want["contigs"], world):
                seg = np.ascontiguousarray(asm[i * clen:(i + 1) * clen])
                check(L.kq_validate_segment(g.handle, C.c_void_p(seg.ctypes.data), clen, 0, C.c_void_p(out.ctypes.data), C.byref(acc)), g.handle)
                rows = pc.add_mod64(rows, pc.rows_digest(out, i * clen))
        if router_v is not None:
            rounds = (want["contigs"] + world - 1) // world
            for r_ in range(rounds):
                i = r_ * world + rank
                if i < want["contigs"]:
                    b, m = pack_to_device(np.ascontiguousarray(asm[i * clen:(i + 1) * clen]))
                    router_v.chunk_packed(b, m, clen)
                    check(L.kq_vroute_copy_rows(g.handle, C.c_void_p(out.ctypes.data), clen), g.handle)
                    torch.cuda.current_stream().synchronize()
                    rows = pc.add_mod64(rows, pc.rows_digest(out, i * clen))
                else:
                    router_v.chunk_packed(None, None, 0)
            qv_all = router_v.counters()
            if rank == 0:                           # counters() is already the sum over ranks
                acc.missing, acc.kcount, acc.edge_missing = qv_all
        elif world > 1:
            with peers:
                vloop()
        else:
            vloop()
        got += [rows, acc.missing, acc.kcount, acc.edge_missing]
        if world > 1:
            allv = [None] * world
            dist.all_gather_object(allv, got)
            got = [sum(v[i] for v in allv) & 0xFFFFFFFFFFFFFFFF for i in range(len(got))]
        exp = want["db"] + [want["instances"], want["rows"]] + want["qv"]
        g.clear()
        return {"ok": got == exp, "digest": got, "expected": exp,
                "what": "db [sum key*cov mod 2^64, sum edges, distinct, total], instances, validate rows digest, missing, kcount, edge_missing"}

    ev = lambda: torch.cuda.Event(enable_timing=True)

    def step_resident():
        e0, e1, e2 = ev(), ev(), ev()
        e0.record(stream)
        g.clear()
        count_resident()
        e1.record(stream)
        validate_resident()
        e2.record(stream)
        return e0, e1, e2

    def step_e2e():
        e0, e1, e2 = ev(), ev(), ev()
        e0.record(stream)
        g.clear()
        if os.environ.get("KQ_E2E_FLUSH"):                     # tuning knob: pending positions that start the pipeline
            g.set_flush_positions(int(os.environ["KQ_E2E_FLUSH"]))
        if world > 1:
            B = 64 * (READ_LEN + 1) * 10000
            # chunks of five pushes (a multiple of 64 bases, so the packed words of the chunks join up): packed on this
            # rank's share of the host cores into pinned memory and copied to their place in the device-resident packed
            # read set while the next chunk is packed; ONE routed round at the end (every round sweeps the whole shard)
            CH = 5 * B
            for i, o in enumerate(range(0, reads_np.size, CH)):
                n = min(CH, reads_np.size - o)
                hb, hm, _, _, done = e2e_bufs[i % 2]
                done.synchronize()
                check(L.kq_pack_host(C.c_void_p(reads_np[o:o + n].ctypes.data), n, C.c_void_p(hb.data_ptr()), C.c_void_p(hm.data_ptr()), 0))
                wb, wm = (n + 31) // 32, (n + 63) // 64
                e2e_all[0][o // 32:o // 32 + wb].copy_(hb[:wb], non_blocking=True)
                e2e_all[1][o // 64:o // 64 + wm].copy_(hm[:wm], non_blocking=True)
                done.record(stream)
            router.count(e2e_all[0], e2e_all[1], reads_np.size)
        else:
            for o, o2 in zip(push_cuts[:-1], push_cuts[1:]):          # ~97 MB pushes, cut at read boundaries
                g.traverseInReads(reads_np[o:o2])
        g.finalize()
        e1.record(stream)
        acc = kreeq_b200._lib.KqQV()

        def validate_host():
            for seg in segs:
                check(L.kq_validate_segment(g.handle, C.c_void_p(seg.ctypes.data), seg.size, 0,
                                            C.c_void_p(h_out.ctypes.data), C.byref(acc)), g.handle)
        def validate_host_routed():
            # pack on the host, copy in, route the queries, copy the rows of this rank's slab out
            for i, seg in enumerate(segs):
                hb, hm, db, dm, done = v_bufs[i % 2]
                done.synchronize()
                check(L.kq_pack_host(C.c_void_p(seg.ctypes.data), seg.size, C.c_void_p(hb.data_ptr()), C.c_void_p(hm.data_ptr()), 0))
                db.copy_(hb, non_blocking=True); dm.copy_(hm, non_blocking=True)
                done.record(stream)
                router_v.chunk_packed(db, dm, seg.size)
                check(L.kq_vroute_copy_rows(g.handle, C.c_void_p(h_out.ctypes.data), seg.size), g.handle)
            router_v.counters()
        if router_v is not None:
            validate_host_routed()
        elif world > 1:
            with peers:
                validate_host()
        else:
            validate_host()
        e2.record(stream)
        return e0, e1, e2

    def run(step_fn, warmup, steps):
        for _ in range(warmup):
            step_fn()
        barrier()
        c0 = g.counters()
        g.set_timing(True)
        for w in range(10):
            g.kernel_time_ms(w)
        sampler = ClockSampler(local); sampler.start()
        evs = [step_fn() for _ in range(steps)]
        barrier()
        clocks = sampler.stop()
        kt_count, kn_count = g.kernel_time_ms(2)           # insert_tuples_kernel launches
        kt_val, kn_val = g.kernel_time_ms(1)
        kt_pipe, kn_pipe = g.kernel_time_ms(0)
        kt_scat, kn_scat = g.kernel_time_ms(3)
        kt_l2, _ = g.kernel_time_ms(4); kt_s1, _ = g.kernel_time_ms(5); kt_s2, _ = g.kernel_time_ms(6)
        kt_vs, _ = g.kernel_time_ms(8); kt_va, _ = g.kernel_time_ms(9)
        g.set_timing(False)
        c1 = g.counters()
        tc = sum(a.elapsed_time(b) for a, b, _ in evs) / steps
        tv = sum(b.elapsed_time(c) for _, b, c in evs) / steps
        t = torch.tensor([tc, tv], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return {"count_ms": t[0].item(), "validate_ms": t[1].item(), "clocks": clocks,
                "launches": c1["kernel_launches"] - c0["kernel_launches"],
                "h2d": (c1["h2d_bytes"] - c0["h2d_bytes"]) / steps, "d2h": (c1["d2h_bytes"] - c0["d2h_bytes"]) / steps,
                "kt_count": kt_count, "kn_count": kn_count, "kt_val": kt_val, "kn_val": kn_val,
                "kt_pipe": kt_pipe, "kn_pipe": kn_pipe, "kt_scat": kt_scat, "kn_scat": kn_scat,
                "kt_l2": kt_l2, "kt_s1": kt_s1, "kt_s2": kt_s2, "kt_vs": kt_vs, "kt_va": kt_va}

    parity = parity_check() if not args.no_parity else {"ok": None, "skipped": "--no-parity"}
    if parity["ok"] is False:
        if rank == 0:
            print(json.dumps({"metric": "kmers_counted_per_sec", "value": None, "parity_check": parity, "error": "parity check failed"}))
        raise SystemExit(3)
    res = run(step_resident, args.warmup, args.steps)
    inst_total = inst_per_rank * world
    value = inst_total / (res["count_ms"] * 1e-3)
    peak, peak_src = measured_peaks()
    line = {"metric": "kmers_counted_per_sec", "value": value, "unit": "k-mers/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["count_ms"] + res["validate_ms"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": dict(cfg, parallelism="keyshard%d" % world, exchange=transport, kmer_instances_per_step=inst_total,
                           assembly_kmers_per_step=asm_k * world, distinct_kmers=None,
                           weak_scaling="rank r counts and validates its own C2 instance (genome seed 2+r); the table is key-sharded over all ranks",
                           l2="no flush needed: table (%.1f GB/GPU) and inputs (%.2f GB/GPU) are far larger than the 126 MB L2"
                              % (est_distinct / 0.5 * 16 / 1e9, reads_np.size * 0.375 / 1e9)),
            "count_ms": res["count_ms"], "validate_ms": res["validate_ms"],
            "validate": {"metric": "assembly_kmers_validated_per_sec", "value": asm_k * world / (res["validate_ms"] * 1e-3), "unit": "k-mers/s"},
            "clocks": res["clocks"], "gpu_launches": res["launches"], "parity_check": parity}
    # roofline of the dominant kernel of the count step, per launch, from CUDA events around each launch: every kernel of
    # the pipeline handles all k-mer instances of the batch once, so each is held against the same 26.25 algorithmic
    # bytes per k-mer; the slowest one is "the dominant kernel"
    if res["kn_count"]:
        pipeline = int(os.environ.get("KQ_PIPELINE", "3"))
        insert_name = "smem_insert_kernel" if pipeline >= 3 else "insert_tuples_kernel"
        if world > 1 and os.environ.get("KQ_ROUTE_SMEM", "0") != "1":
            insert_name = "insert_tuples_kernel"        # the routed slab is counted with L2 CAS
        scat_name = ("slice_scatter_kernel<routed>" if world > 1 else "slice_scatter_kernel") if pipeline != 4 else "scatter_fine_kernel"
        per_step = {"slice_hist_kernel (+offsets)": (res["kt_scat"] - res["kt_s1"]) / args.steps, scat_name: res["kt_s1"] / args.steps,
                    "fine_hist_kernel (+plan, offsets)": (res["kt_l2"] - res["kt_s2"]) / args.steps, "fine_scatter_kernel": res["kt_s2"] / args.steps,
                    insert_name: res["kt_count"] / args.steps}
        if world > 1:                                 # the routed steps time histogram + offsets + scatter together
            per_step = {"slice_hist_kernel<routed> + route_offsets + slice_scatter_kernel<routed>": res["kt_scat"] / args.steps,
                        "fine_hist_kernel (+offsets)": (res["kt_l2"] - res["kt_s2"]) / args.steps, "fine_scatter_kernel": res["kt_s2"] / args.steps,
                        insert_name: res["kt_count"] / args.steps}
        per_step = {k_: v for k_, v in per_step.items() if v > 0}
        dom = max(per_step, key=per_step.get)
        launches_per_step = max(1, round(res["kn_count"] / args.steps))
        per_launch_s = per_step[dom] * 1e-3 / launches_per_step
        units_per_launch = inst_per_rank / launches_per_step
        ach = ALG_BYTES_COUNT * units_per_launch / per_launch_s / 1e9
        traffic, traffic_src = None, None
        try:                                        # dram__bytes_read+write of this kernel from the committed ncu --set full capture
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                t = json.load(f)[dom.split("<")[0].split(" ")[0]]
            traffic = t["dram_bytes_per_tuple"] * units_per_launch
            traffic_src = "%s: %.1f B/tuple at %d tuples per launch, scaled to this run's tuples per launch" % (
                t["summary"], t["dram_bytes_per_tuple"], t["tuples_per_launch"])
        except Exception:
            pass
        line["roofline"] = {"bound": "hbm", "kernel": dom,
                            "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                            "traffic_source": traffic_src,
                            "peak_source": peak_src, "algorithmic_bytes_per_kmer": ALG_BYTES_COUNT,
                            "sector_bound_frac": SECTOR_BYTES_COUNT * units_per_launch / per_launch_s / 1e9 / peak,
                            "avg_launch_ms": per_launch_s * 1e3, "launches_timed": launches_per_step * args.steps,
                            "share_of_count_ms": per_step[dom] / res["count_ms"],
                            "kernels_ms_per_step": per_step,
                            "kernels_frac": {k_: ALG_BYTES_COUNT * inst_per_rank / (v * 1e-3) / 1e9 / peak for k_, v in per_step.items()},
                            "count_step_frac": ALG_BYTES_COUNT * inst_per_rank / (res["count_ms"] * 1e-3) / 1e9 / peak,
                            "pipeline_ms_per_step": res["kt_pipe"] / args.steps}
        if res["kn_val"]:
            vs = res["kt_val"] * 1e-3 / res["kn_val"]
            va = ALG_BYTES_VALIDATE * asm_k * args.steps / (res["kt_val"] * 1e-3) / 1e9      # this rank's positions / its kernels' time
            line["validate"]["roofline"] = {"bound": "hbm", "kernel": "validate_kernel" if world == 1 or vmode == "peer" else "vroute_send_kernel + vroute_answer_kernel",
                                            "achieved": va, "peak": peak, "unit": "GB/s", "frac": va / peak, "avg_launch_ms": vs * 1e3,
                                            "launches_timed": res["kn_val"]}
            if res["kt_vs"]:
                line["validate"]["roofline"]["kernels_ms_per_step"] = {"vroute_send": res["kt_vs"] / args.steps, "vroute_answer": res["kt_va"] / args.steps}
    try:
        line["config"]["distinct_kmers"] = g.distinct() if world == 1 else None
    except Exception:
        pass
    if not args.no_e2e:
        e = run(step_e2e, min(args.warmup, 1), min(args.steps, 2))
        line["e2e"] = {"value": inst_total / (e["count_ms"] * 1e-3), "unit": "k-mers/s",
                       "h2d_bytes_per_step": e["h2d"] if world == 1 else reads_np.size * 0.375,
                       "d2h_bytes_per_step": e["d2h"], "count_ms": e["count_ms"], "validate_ms": e["validate_ms"],
                       "validate_value": asm_k * world / (e["validate_ms"] * 1e-3),
                       "host_input_bytes_per_step": int(reads_np.size + sum(s.size for s in segs)), "steps": min(args.steps, 2),
                       "validate_out": "pageable host memory (staged)" if args.pageable_out else "page-locked host memory (direct)"}
    if world == 1 and not args.no_e2e:
        # the same validate pass with compact rows back (kq_validate_segment_compact: 4 bytes per base over PCIe instead
        # of 16; rows rebuilt from them are checked against the 16-byte rows of the last segment)
        try:
            from kreeq_b200._lib import BASE4, WIDE, KqQV
            h4_t = torch.empty(contig * 4, dtype=torch.uint8, pin_memory=True)
            h4 = h4_t.numpy().view(BASE4)
            wide = np.zeros(1 << 20, dtype=WIDE)

            def compact_pass():
                acc4 = KqQV()
                nw_last = 0
                for seg in segs:
                    nw = C.c_size_t(wide.size)
                    check(L.kq_validate_segment_compact(g.handle, C.c_void_p(seg.ctypes.data), seg.size, 0, C.c_void_p(h4.ctypes.data),
                                                        C.c_void_p(wide.ctypes.data), C.byref(nw), C.byref(acc4)), g.handle)
                    nw_last = nw.value
                return acc4, nw_last
            compact_pass()
            t0c = time.perf_counter(); acc4, nw_last = compact_pass(); dtc = time.perf_counter() - t0c
            full = np.zeros(segs[-1].size, dtype=kreeq_b200.BASE); accf = KqQV()
            check(L.kq_validate_segment(g.handle, C.c_void_p(segs[-1].ctypes.data), segs[-1].size, 0, C.c_void_p(full.ctypes.data), C.byref(accf)), g.handle)
            back = np.zeros(segs[-1].size, dtype=kreeq_b200.BASE)
            check(L.kq_expand_rows(C.c_void_p(h4.ctypes.data), segs[-1].size, C.c_void_p(wide.ctypes.data), nw_last, C.c_void_p(back.ctypes.data)))
            line["e2e"]["validate_compact"] = {"value": asm_k / dtc, "unit": "k-mers/s", "ms": dtc * 1e3, "d2h_bytes": int(4 * sum(s.size for s in segs)),
                                               "rows_match_16_byte_rows": bool(back.tobytes() == full.tobytes()), "wide_rows_last_segment": int(nw_last),
                                               "what": "kq_validate_segment_compact from host ASCII, {cov, fw, bw, flags} bytes into page-locked memory (host wall clock)"}
        except Exception as ex:                          # a side measurement must not cost the line
            line["e2e"]["validate_compact"] = {"error": str(ex)}
    if world == 1 and not args.no_union:
        # database rows (A4 / A8): 2 x 5.8e7-record databases of a 10 Mbp genome saved, merged, saved, loaded
        from kreeq_b200.dbbench import union_rows
        try:
            u = union_rows(10.0, 30.0, None, local)
            line["union"] = {"metric": "union_input_records_per_sec", "value": u["union_input_records_per_s"], "unit": "records/s",
                             "save_records_per_sec": u["save_records_per_s"], "load_records_per_sec": u["load_records_per_s"],
                             "records_in": u["records_in"], "records_out": u["records_out"], "db_bytes": u["db_bytes"],
                             "files_on": u["files_on"], "mode": "reference-compatible (kq_union ..., 1)"}
            if u.get("union_device_s"):
                line["union"]["device"] = {"value": u["union_device_records_per_s"], "unit": "records/s", "seconds": u["union_device_s"],
                                           "kernels": "db_decode_kernel + put_kernel + merge_ref_kernel (CUDA events)",
                                           "roofline": {"bound": "hbm", "achieved": u["union_device_gbs"], "peak": peak, "unit": "GB/s",
                                                        "frac": u["union_device_gbs"] / peak, "algorithmic_bytes_per_record": u["union_alg_bytes_per_record"]}}
        except Exception as ex:                       # never lose the main line to the side measurement
            line["union"] = {"error": str(ex)[:200]}
    if world > 1:
        line["validate"]["exchange"] = vmode
    if peers is not None:
        line["validate"]["peer_connect_ms"] = peers.connect_ms
        peers.disconnect()
    g.close()
    del d_rb, d_rm, d_asm
    torch.cuda.empty_cache()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline(args, 1, 1)
        if cb:
            cb["sample_config"] = {"reads": args.ref_sample_reads // 2 * 2, "assembly_bp": min(gsize, 10_000_000), "same_config_as_line": False}
            line["cpu_baseline"] = cb
    if rank == 0 and world == 1 and not args.no_cli_e2e:
        # like-for-like with the reference arm: our command line on the same full-size files, same two commands
        try:
            line["cli_e2e"] = cli_e2e(args, "ours")
        except Exception as ex:
            line["cli_e2e"] = {"error": str(ex)[:300]}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        main_reference(a)
    else:
        main_ours(a)

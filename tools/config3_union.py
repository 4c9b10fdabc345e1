"""BASELINE.json config 3 (SURVEY 8d "C3"): eight read sets (seeds 30..37) of the C2 genome with 64 copies of a
5 kb repeat and one 2 kb poly-A planted, counted separately to s0..s7.kreeq, `union -d s0 .. s7 -o u.kreeq`
(reference-compatible order rules, DB0 = s0), then the genome validated against the union.

  python tools/config3_union.py                 # the /16 twin: 6.25 Mbp genome, 8 x 543 750 pairs
  python tools/config3_union.py --scale 4       # 25 Mbp (needs ~25 GB of files under --dir)
  python tools/config3_union.py --scale 1       # full C3: 8 x ~20 GB of files, a ~100 GB table

Prints one JSON line: records in / out, union wall time, the device time of its decode + put / merge kernels (CUDA
events, timer 7) with the roofline of SURVEY 8d (17 (1 + R_out / R_in) bytes per input record against the measured
HBM peak), save / validate times and the QV counters.  --gpus N runs the same through the multi-GPU handle."""
import argparse
import json
import os
import shutil
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except (OSError, ValueError, KeyError):
        return 6650.0, "fallback (B200_PROFILING.md)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=16.0, help="divide C3's sizes by this (default 16: the twin)")
    ap.add_argument("--dir", default=None, help="where the databases go (default /dev/shm)")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--sets", type=int, default=8)
    a = ap.parse_args()
    import torch
    import kreeq_b200 as kb
    from kreeq_b200 import synth
    dev = torch.device("cuda", 0)
    k, L, err = 21, 150, 0.01
    gsize = int(100_000_000 / a.scale)
    pairs = int(8_700_000 / a.scale)
    genome = synth.random_genome_codes(gsize, 2, dev)
    # 64 copies of one 5 kb repeat and one 2 kb poly-A, at seeded non-overlapping places
    rng = np.random.default_rng(3)
    rep = torch.from_numpy(rng.integers(0, 4, 5000).astype(np.uint8)).to(dev)
    places = rng.choice(gsize // 7000 - 1, 65, replace=False) * 7000
    for s in places[:64]:
        genome[int(s):int(s) + 5000] = rep
    genome[int(places[64]):int(places[64]) + 2000] = 0
    asm = synth.codes_to_ascii(genome).cpu().numpy()
    base = a.dir or ("/dev/shm" if os.path.isdir("/dev/shm") else None)
    work = tempfile.mkdtemp(prefix="kq_c3_", dir=base)
    est = int((gsize + 2 * pairs * L * err * k * 0.85) * 1.05)
    out = {"config": "C3 / %g" % a.scale, "genome_bp": gsize, "pairs_per_set": pairs, "sets": a.sets, "k": k, "gpus": a.gpus,
           "files_on": base or tempfile.gettempdir()}

    def make(expected):
        return kb.MultiDBG(k=k, ngpus=a.gpus, expected_distinct=expected) if a.gpus > 1 else kb.DBG(k=k, expected_distinct=expected)
    try:
        dbs, recs, t_count, t_save = [], [], 0.0, 0.0
        for i in range(a.sets):
            reads = synth.paired_reads_ascii(genome, pairs, L, err, 30 + i).cpu().numpy()
            with make(est) as g:
                t = time.perf_counter(); g.traverseInReads(reads); g.finalize(); t_count += time.perf_counter() - t
                d = os.path.join(work, "s%d.kreeq" % i)
                t = time.perf_counter(); g.save(d); t_save += time.perf_counter() - t
                recs.append(g.stats()["distinct"]); dbs.append(d)
            del reads
        out["records_in"] = recs
        out["count_s_total"] = t_count; out["save_s_total"] = t_save
        out["db_bytes_in"] = sum(os.path.getsize(os.path.join(d, f)) for d in dbs for f in os.listdir(d))
        r_in = sum(recs)
        with make(int(1.2 * r_in / max(1, a.sets) * 2.2)) as g:
            timed = a.gpus == 1
            if timed:
                g.set_timing(True); g.kernel_time_ms(7)
            t = time.perf_counter(); g.kunion(dbs); out["union_s"] = time.perf_counter() - t
            st = g.stats()
            out["records_out"] = st["distinct"]; out["stats"] = st
            out["union_input_records_per_s"] = r_in / out["union_s"]
            if timed:
                ms, nk = g.kernel_time_ms(7); g.set_timing(False)
                peak, src = hbm_peak()
                bpr = 17.0 * (1.0 + st["distinct"] / r_in)
                gbs = bpr * r_in / (ms * 1e-3) / 1e9
                out["union_device"] = {"seconds": ms * 1e-3, "launches": nk, "records_per_s": r_in / (ms * 1e-3),
                                       "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                                                    "algorithmic_bytes_per_record": bpr, "peak_source": src}}
            d = os.path.join(work, "u.kreeq")
            t = time.perf_counter(); g.save(d); out["save_union_s"] = time.perf_counter() - t
            out["db_bytes_out"] = sum(os.path.getsize(os.path.join(d, f)) for f in os.listdir(d))
            t = time.perf_counter(); g.evaluateSegment(asm); out["validate_s"] = time.perf_counter() - t
            out["qv"] = {"missing": g.tot_missing, "kcount": g.tot_kcount, "edge_missing": g.tot_edge_missing}
    finally:
        shutil.rmtree(work, ignore_errors=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()

"""Times the database rows of the path on one GPU (SURVEY 8d "Union: input records/s"): count two read
sets of the same genome, kq_db_save each, kq_union (reference-compatible order rules) of the two, kq_db_save
of the result, kq_db_load back.  Sizes via --genome-mbp (default 10 Mbp, 30x: ~5.8e7 records per database).
Prints one JSON line."""
import argparse
import json
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome-mbp", type=float, default=10.0)
    ap.add_argument("--coverage", type=float, default=30.0)
    ap.add_argument("--dir", default=None)
    args = ap.parse_args()
    import torch
    import kreeq_b200
    from kreeq_b200 import synth
    dev = torch.device("cuda", 0)
    k, L = 21, 150
    gsize = int(args.genome_mbp * 1e6)
    pairs = int(gsize * args.coverage / (2 * L))
    genome = synth.random_genome_codes(gsize, 2, dev)
    work = tempfile.mkdtemp(prefix="kq_union_", dir=args.dir)
    out = {"genome_bp": gsize, "pairs_per_db": pairs, "k": k}
    try:
        dbs, recs = [], []
        for i in range(2):
            reads = synth.paired_reads_ascii(genome, pairs, L, 0.01, 30 + i).cpu().numpy()
            with kreeq_b200.DBG(k=k, expected_distinct=int(gsize + 2 * pairs * L * 0.01 * k * 0.9)) as g:
                t = time.perf_counter(); g.traverseInReads(reads); g.finalize(); g.sync(); out["count_s_%d" % i] = time.perf_counter() - t
                d = os.path.join(work, "s%d.kreeq" % i)
                t = time.perf_counter(); g.save(d); out["save_s_%d" % i] = time.perf_counter() - t
                recs.append(g.distinct()); dbs.append(d)
        out["records_in"] = recs
        with kreeq_b200.DBG(k=k, expected_distinct=int(1.8 * recs[0])) as g:
            t = time.perf_counter(); g.kunion(dbs); g.sync(); out["union_s"] = time.perf_counter() - t
            out["records_out"] = g.distinct()
            d = os.path.join(work, "u.kreeq")
            t = time.perf_counter(); g.save(d); out["save_union_s"] = time.perf_counter() - t
            out["db_bytes"] = sum(os.path.getsize(os.path.join(d, f)) for f in os.listdir(d))
        with kreeq_b200.DBG(k=k, expected_distinct=out["records_out"]) as g:
            t = time.perf_counter(); g.load(d); g.sync(); out["load_s"] = time.perf_counter() - t
            assert g.distinct() == out["records_out"]
        out["union_input_records_per_s"] = sum(recs) / out["union_s"]
        out["save_records_per_s"] = out["records_out"] / out["save_union_s"]
        out["load_records_per_s"] = out["records_out"] / out["load_s"]
    finally:
        shutil.rmtree(work, ignore_errors=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()

"""Timing of the database rows of the path on one GPU (SURVEY 8d "Union: input records/s"): count two read
sets of the same genome, kq_db_save each, kq_union (reference-compatible order rules) of the two, kq_db_save of
the result, kq_db_load back.  Used by bench.py (`union` object) and tools/union_bench.py."""
import os
import shutil
import tempfile
import time


def union_rows(genome_mbp=10.0, coverage=30.0, workdir=None, device=0):
    import torch
    from . import synth
    from .dbg import DBG
    dev = torch.device("cuda", device)
    k, L = 21, 150
    gsize = int(genome_mbp * 1e6)
    pairs = int(gsize * coverage / (2 * L))
    genome = synth.random_genome_codes(gsize, 2, dev)
    base = workdir or ("/dev/shm" if os.path.isdir("/dev/shm") else None)
    work = tempfile.mkdtemp(prefix="kq_union_", dir=base)
    out = {"genome_bp": gsize, "pairs_per_db": pairs, "k": k, "files_on": base or tempfile.gettempdir()}
    try:
        dbs, recs = [], []
        for i in range(2):
            reads = synth.paired_reads_ascii(genome, pairs, L, 0.01, 30 + i).cpu().numpy()
            with DBG(k=k, device=device, expected_distinct=int(gsize + 2 * pairs * L * 0.01 * k * 0.9)) as g:
                g.traverseInReads(reads); g.finalize(); g.sync()
                d = os.path.join(work, "s%d.kreeq" % i)
                t = time.perf_counter(); g.save(d); out["save_s_%d" % i] = time.perf_counter() - t
                recs.append(g.distinct()); dbs.append(d)
            del reads
        out["records_in"] = recs
        with DBG(k=k, device=device, expected_distinct=int(1.8 * recs[0])) as g:
            g.set_timing(True); g.kernel_time_ms(7)
            t = time.perf_counter(); g.kunion(dbs); g.sync(); out["union_s"] = time.perf_counter() - t
            ms, nk = g.kernel_time_ms(7); g.set_timing(False)
            out["union_device_s"] = ms * 1e-3; out["union_device_launches"] = nk
            out["records_out"] = g.distinct()
            d = os.path.join(work, "u.kreeq")
            t = time.perf_counter(); g.save(d); out["save_union_s"] = time.perf_counter() - t
            out["db_bytes"] = sum(os.path.getsize(os.path.join(d, f)) for f in os.listdir(d))
        with DBG(k=k, device=device, expected_distinct=out["records_out"]) as g:
            t = time.perf_counter(); g.load(d); g.sync(); out["load_s"] = time.perf_counter() - t
            if g.distinct() != out["records_out"]:
                raise RuntimeError("load after save lost records")
        out["union_input_records_per_s"] = sum(recs) / out["union_s"]
        if out.get("union_device_s"):
            # 17 B read per input record + 17 B written per output record, against the device time of the
            # decode + put / merge kernels (CUDA events); the rest of union_s is file I/O and the host directory pass
            out["union_device_records_per_s"] = sum(recs) / out["union_device_s"]
            out["union_alg_bytes_per_record"] = 17.0 * (1.0 + out["records_out"] / sum(recs))
            out["union_device_gbs"] = out["union_alg_bytes_per_record"] * sum(recs) / out["union_device_s"] / 1e9
        out["save_records_per_s"] = out["records_out"] / out["save_union_s"]
        out["load_records_per_s"] = out["records_out"] / out["load_s"]
    finally:
        shutil.rmtree(work, ignore_errors=True)
    return out

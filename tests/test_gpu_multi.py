"""kq_multi (the C++ multi-GPU layer the command line's --gpus uses) against the oracle and against one GPU.

SURVEY 8e invariant: for every GPU count G the decoded database, the statistics, the QV counters and the per-base
rows are those of G = 1 and of the oracle.  On a box with fewer than G GPUs every shard runs on device 0: the same
host threads, rounds, peer stores and routed queries, only the memory is local."""
import os

import numpy as np
import pytest

from helpers import assert_same_records, random_genome, sample_reads

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import kreeq_b200
    import kreeq_oracle as ko
    return kreeq_b200, ko


def devices_for(n):
    import torch
    have = torch.cuda.device_count()
    return [g % have for g in range(n)]


def problem(seed, k, genome_bp=150_000, n_batches=6, reads=2500, n_rate=0.001):
    rng = np.random.default_rng(seed)
    genome = random_genome(rng, genome_bp)
    rep = (b"A" * 400 + b"N") * 2 + (b"ACGT" * 3000 + b"N")          # keys past 127 and past 255 instances
    batches = [sample_reads(rng, genome, reads, 150, err=0.01, n_rate=n_rate) + (rep if i % 2 == 0 else b"") for i in range(n_batches)]
    return genome, batches


@pytest.mark.parametrize("k,G", [(21, 2), (21, 4), (31, 4), (21, 8)])
def test_count_validate_any_gpu_count(mods, k, G):
    kb, ko = mods
    genome, batches = problem(7000 + k + G, k)
    ref = ko.Table(k)
    n_ref = sum(ref.count_batch(b) for b in batches)
    with kb.MultiDBG(k=k, devices=devices_for(G), round_positions=300_000, validate_chunk=40_000) as g:
        for b in batches:
            g.traverseInReads(b)                      # dealt to the GPUs; rounds start while more batches arrive
        g.finalize()
        assert g.instances() == n_ref
        assert_same_records(g.export(), ref.export())
        assert g.stats() == ref.stats()
        hc, hn = g.hist(); rc_, rn = ref.hist()
        assert np.array_equal(hc, rc_) and np.array_equal(hn, rn)
        for cutoff in (0, 3):
            g.tot_missing = g.tot_kcount = g.tot_edge_missing = 0
            seg = genome.tobytes() + b"A" * 450
            out = g.evaluateSegment(seg, cov_cutoff=cutoff)         # several chunks per GPU, last round partly empty
            rc, rout, racc = ref.validate_segment(seg, cutoff)
            assert out.tobytes() == rout.tobytes()
            assert (g.tot_missing, g.tot_kcount, g.tot_edge_missing) == racc
        out = g.evaluateSegment(b"ACGTAC")                          # shorter than k
        assert not out["cov"].any()
        rk, rr = ref.export()
        rec, found = g.find(np.concatenate([rk[:300], np.arange(10**9, 10**9 + 50, dtype=np.uint64)]))
        assert found[:300].all() and rec[:300].tobytes() == rr[:300].tobytes()


@pytest.mark.parametrize("k,G", [(21, 2), (31, 4)])
def test_routed_slab_through_level2_and_shared_memory_insert(mods, k, G):
    """route_smem = 1 (off by default): the owner's slab goes through route_plan / per-task histograms /
    route_fine_offsets / level-2 scatter with private cursors / shared-memory insert; small slices so that slab bins span
    several slices and slices lie across bins"""
    kb, ko = mods
    from kreeq_b200._lib import lib
    L = lib()
    knobs = {"route_smem": (1, 0), "single_slice_bits": (8, 21), "piece_bits": (8, 12), "fine_part": (2048, 1 << 18), "direct_pct": (0, 25)}
    try:
        for n, (v, _) in knobs.items():
            assert L.kq_set_tuning(n.encode(), v) == 0
        genome, batches = problem(9100 + k + G, k)
        ref = ko.Table(k)
        n_ref = sum(ref.count_batch(b) for b in batches)
        with kb.MultiDBG(k=k, devices=devices_for(G), expected_distinct=600_000, round_positions=300_000) as g:
            for b in batches:
                g.traverseInReads(b)
            g.finalize()
            assert g.instances() == n_ref
            assert_same_records(g.export(), ref.export())
            assert g.stats() == ref.stats()
    finally:
        for n, (_, d) in knobs.items():
            L.kq_set_tuning(n.encode(), d)


def test_save_load_union_sharded(mods, tmp_path):
    """databases written by 4 shards (16 maps + one .hc.part each, merged) are the files one GPU writes: the strict
    phmap decoder reads them, one GPU loads them, 2 shards merge them in three orders like the oracle's union"""
    kb, ko = mods
    k = 21
    dirs, refs = [], []
    for i in range(3):
        genome, batches = problem(8100 + i, k, genome_bp=40_000, n_batches=2, reads=900, n_rate=0.0)
        if i == 1:
            batches = [b.replace(b"A" * 400 + b"N", b"") for b in batches]        # no poly-A in DB 1
        ref = ko.Table(k)
        for b in batches:
            ref.count_batch(b)
        ref.finalize(); refs.append(ref)
        d = str(tmp_path / ("s%d.kreeq" % i)); dirs.append(d)
        with kb.MultiDBG(k=k, devices=devices_for(4), round_positions=200_000) as g:
            for b in batches:
                g.traverseInReads(b)
            g.finalize(); g.save(d)
        assert sorted(f for f in os.listdir(d) if "part" in f) == []
        disk = ko.read_db(d, check=True)
        assert_same_records(disk.export(), ref.export())
        assert disk.stats() == ref.stats()
        with kb.DBG(k=k) as g1:
            g1.load(d)
            assert_same_records(g1.export(), ref.export())
        with kb.MultiDBG(k=k, devices=devices_for(2)) as g2:
            g2.load(d)
            assert_same_records(g2.export(), ref.export())
            assert g2.stats() == ref.stats()
    for order in ([0, 1, 2], [1, 0, 2], [2, 1, 0]):
        u = ko.union([ko.read_db(dirs[i], check=False) for i in order])
        with kb.MultiDBG(k=k, devices=devices_for(2), validate_chunk=10_000) as g:
            g.kunion([dirs[i] for i in order], ref_compatible=True)
            assert_same_records(g.export(), u.export())
            assert g.stats() == u.stats()
            seg = b"A" * 60 + b"ACGT" * 40
            out = g.evaluateSegment(seg)
            rc, rout, racc = u.validate_segment(seg)
            assert out.tobytes() == rout.tobytes()
            ud = str(tmp_path / ("u%s.kreeq" % "".join(map(str, order))))
            g.save(ud)
        disk = ko.read_db(ud, check=True)
        k8a, _ = disk.dump8(); k8b, _ = u.dump8()
        assert np.array_equal(k8a, k8b)
        k32a, r32a = disk.dump32(); k32b, r32b = u.dump32()
        assert np.array_equal(k32a, k32b) and r32a.tobytes() == r32b.tobytes()


def test_cli_gpus_flag_same_output(mods, tmp_path):
    """kreeq --devices 0,0,0,0 (4 shards) prints the stdout and writes the .kwig / database records of --gpus 1"""
    import subprocess
    kb, ko = mods
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "kreeq_b200", "kreeq")
    rng = np.random.default_rng(99)
    genome = random_genome(rng, 60_000)
    fq = tmp_path / "r.fastq"; fa = tmp_path / "a.fasta"
    batch = sample_reads(rng, genome, 6000, 150, err=0.01)
    with open(fq, "wb") as f:
        for r in batch.split(b"N")[:-1]:
            f.write(b"@r\n" + r + b"\n+\n" + b"I" * len(r) + b"\n")
    g = genome.tobytes()
    with open(fa, "wb") as f:
        f.write(b">c1\n" + g[:30_000] + b"NNNNNNNNNN" + g[30_000:45_000] + b"\n>c2\n" + g[45_000:] + b"\n")
    outs = {}
    devs = ",".join(str(d) for d in devices_for(4))
    for name, flags in (("g1", ["--gpus", "1"]), ("g4", ["--devices", devs])):
        kw = tmp_path / (name + ".kwig"); db = tmp_path / (name + ".kreeq")
        a = subprocess.run([exe, "validate", "-f", str(fa), "-r", str(fq), "-k", "21", "-o", str(kw)] + flags, capture_output=True, cwd=tmp_path)
        assert a.returncode == 0, a.stderr.decode()
        b = subprocess.run([exe, "validate", "-r", str(fq), "-k", "21", "-o", str(db)] + flags, capture_output=True, cwd=tmp_path)
        assert b.returncode == 0, b.stderr.decode()
        c = subprocess.run([exe, "validate", "-f", str(fa), "-d", str(db), "-o", str(tmp_path / (name + "b.kwig"))] + flags, capture_output=True, cwd=tmp_path)
        assert c.returncode == 0, c.stderr.decode()
        outs[name] = (a.stdout, b.stdout, c.stdout, open(kw, "rb").read(), open(tmp_path / (name + "b.kwig"), "rb").read(), ko.read_db(str(db), check=True).export())
    assert outs["g1"][:5] == outs["g4"][:5]
    assert outs["g1"][3] == outs["g1"][4]
    assert_same_records(outs["g1"][5], outs["g4"][5])


def test_routed_validate_several_segments_per_round(mods):
    """kq_vroute_send_at: the queries of several segments go out before ONE pair of barriers; rows of segment i start
    at the sum of the lengths before it.  Two shards (handles) in this process, their inboxes / row slabs connected
    by plain pointers."""
    import ctypes as C
    import torch
    kb, ko = mods
    from kreeq_b200._lib import BASE, KqQV, check, lib
    L = lib()
    k, G = 21, 2
    genome, batches = problem(4242, k, genome_bp=90_000, n_batches=3)
    ref = ko.Table(k)
    for b in batches:
        ref.count_batch(b)
    segs = [genome.tobytes()[:30_000], genome.tobytes()[30_000:52_000] + b"A" * 300, genome.tobytes()[52_000:], b"ACGT"]
    mine = {0: [0, 2], 1: [1, 3]}                                   # segments of rank 0 / rank 1 (the last is shorter than k)
    hs = []
    try:
        for r in range(G):
            g = kb.DBG(k=k, expected_distinct=400_000)
            check(L.kq_set_shard(g.handle, r, G), g.handle)
            for b in batches:
                g.traverseInReads(b)
            g.finalize()
            hs.append(g)
        chunk = 70_000
        inbox, rows = (C.c_void_p * G)(), (C.c_void_p * G)()
        for r, g in enumerate(hs):
            a, b_ = C.c_void_p(), C.c_void_p()
            check(L.kq_vroute_setup(g.handle, chunk, C.byref(a), None, C.byref(b_), None), g.handle)
            inbox[r], rows[r] = a.value, b_.value
        for g in hs:
            check(L.kq_vroute_connect(g.handle, inbox, rows), g.handle)
        keep = []
        for r, g in enumerate(hs):
            row0 = 0
            for j, i in enumerate(mine[r]):
                a = np.frombuffer(segs[i], dtype=np.uint8)
                bases, inv = kb.pack_host(a)
                db = torch.from_numpy(bases.view(np.int64)).cuda(); di = torch.from_numpy(inv.view(np.int64)).cuda()
                keep += [db, di]
                check(L.kq_vroute_send_at(g.handle, C.c_void_p(db.data_ptr()), C.c_void_p(di.data_ptr()), 0, 0, a.size,
                                          max(0, a.size - k + 1), row0, 1 if j == len(mine[r]) - 1 else 0), g.handle)
                row0 += a.size
        for g in hs:
            g.sync()
        for g in hs:
            check(L.kq_vroute_answer(g.handle, 0), g.handle)
        for g in hs:
            g.sync()
        for g in hs:
            check(L.kq_vroute_place(g.handle), g.handle)
        tot = [0, 0, 0]
        for r, g in enumerate(hs):
            n = sum(len(segs[i]) for i in mine[r])
            out = np.zeros(n, dtype=BASE)
            check(L.kq_vroute_copy_rows(g.handle, C.c_void_p(out.ctypes.data), n), g.handle)
            g.sync()
            row0 = 0
            for i in mine[r]:
                rc, rout, racc = ref.validate_segment(segs[i])
                assert out[row0:row0 + len(segs[i])].tobytes() == rout.tobytes(), (r, i)
                row0 += len(segs[i])
            q = KqQV()
            check(L.kq_vroute_counters(g.handle, C.byref(q), 1), g.handle)
            tot = [tot[0] + q.missing, tot[1] + q.kcount, tot[2] + q.edge_missing]
        want = [0, 0, 0]
        for s_ in segs:
            racc = ref.validate_segment(s_)[2]
            want = [want[j] + racc[j] for j in range(3)]
        assert tot == want
    finally:
        for g in hs:
            g.close()

"""GPU parity: the CUDA path through the C-ABI against the oracle, bit-exact (integer work)."""
import numpy as np
import pytest

from helpers import assert_same_records, random_genome, sample_reads

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import kreeq_b200
    import kreeq_oracle as ko
    return kreeq_b200, ko


@pytest.mark.parametrize("k", [21, 31, 32, 15, 4, 1])
def test_count_validate_matches_oracle(mods, k):
    kb, ko = mods
    rng = np.random.default_rng(100 + k)
    genome = random_genome(rng, 30000)
    batches = [sample_reads(rng, genome, 1500, 150, err=0.01, n_rate=0.002) for _ in range(3)]
    ref = ko.Table(k)
    with kb.DBG(k=k) as g:
        n_ref = 0
        for b in batches:
            n_ref += ref.count_batch(b)
            g.traverseInReads(b)
        g.finalize()
        assert g.instances() == n_ref
        assert_same_records(g.export(), ref.export())
        assert g.stats() == ref.stats()
        hc, hn = g.hist(); rc_, rn = ref.hist()
        assert np.array_equal(hc, rc_) and np.array_equal(hn, rn)
        for cutoff in (0, 3):
            g.tot_missing = g.tot_kcount = g.tot_edge_missing = 0
            asm = genome.copy()
            asm[rng.integers(0, asm.size, 40)] = ord("A")
            out = g.evaluateSegment(asm.tobytes(), cov_cutoff=cutoff)
            rc, rout, racc = ref.validate_segment(asm.tobytes(), cutoff)
            assert rc == 0
            assert out.tobytes() == rout.tobytes()
            assert (g.tot_missing, g.tot_kcount, g.tot_edge_missing) == racc


def test_high_copy_spill(mods):
    """>254 and >65535 instances of one k-mer: 32-bit spill record and the side table."""
    kb, ko = mods
    k = 21
    rng = np.random.default_rng(5)
    unit = random_genome(rng, 60).tobytes()
    batch = (b"A" * 500 + b"N") * 3 + (unit + b"N") * 300 + (b"ACGT" * 20000 + b"N")
    ref = ko.Table(k); ref.count_batch(batch)
    with kb.DBG(k=k) as g:
        g.traverseInReads(batch); g.finalize()
        assert_same_records(g.export(), ref.export())
        assert g.stats() == ref.stats()
        seg = b"A" * 100 + unit + b"ACGT" * 30
        out = g.evaluateSegment(seg)
        rc, rout, racc = ref.validate_segment(seg)
        assert out.tobytes() == rout.tobytes()
        assert (g.tot_missing, g.tot_kcount, g.tot_edge_missing) == racc


def test_edge_cases(mods):
    kb, ko = mods
    with kb.DBG(k=21) as g:
        g.traverseInReads(b"")                       # empty batch
        g.traverseInReads(b"ACGTN")                  # shorter than k
        g.traverseInReads(b"N" * 100)                # nothing valid
        g.traverseInReads(b"ACGTACGTACGTACGTACGTA")  # exactly one k-mer, no trailing N
        g.finalize()
        ref = ko.Table(21)
        for b in (b"", b"ACGTN", b"N" * 100, b"ACGTACGTACGTACGTACGTA"):
            ref.count_batch(b)
        assert_same_records(g.export(), ref.export())
        assert g.stats() == ref.stats()
        out = g.evaluateSegment(b"ACGT")             # segment shorter than k
        assert out.size == 4 and not out["cov"].any() and g.tot_kcount == 0
        out = g.evaluateSegment(b"ACGTACGTACGTACGTACGTAC")
        rc, rout, racc = ref.validate_segment(b"ACGTACGTACGTACGTACGTAC")
        assert out.tobytes() == rout.tobytes()


def test_large_batch_chunking_and_growth(mods):
    """One push larger than the staging chunk, table growing from its minimum size."""
    kb, ko = mods
    k = 21
    rng = np.random.default_rng(11)
    genome = random_genome(rng, 3_000_000)
    batch = sample_reads(rng, genome, 130000, 150, err=0.01)     # ~19.6 M bases > 16 Mi chunk
    ref = ko.Table(k); n_ref = ref.count_batch(batch)
    with kb.DBG(k=k) as g:
        g.traverseInReads(batch); g.finalize()
        assert g.instances() == n_ref
        assert_same_records(g.export(), ref.export())
        out = g.evaluateSegment(genome.tobytes())                # > 4 Mi-entry validate chunks? no: 3 M; still chunk edge below
        rc, rout, racc = ref.validate_segment(genome.tobytes())
        assert out.tobytes() == rout.tobytes()
        assert (g.tot_missing, g.tot_kcount, g.tot_edge_missing) == racc


def test_save_load_roundtrip_and_union(mods, tmp_path):
    kb, ko = mods
    k = 21
    rng = np.random.default_rng(21)
    genome = random_genome(rng, 20000)
    rep = (b"A" * 400 + b"N") * 2
    dirs, refs = [], []
    for i in range(3):
        batch = sample_reads(rng, genome, 800, 150, err=0.02) + (rep if i != 1 else b"") + (b"ACGT" * 100 + b"N") * (i + 1)
        ref = ko.Table(k); ref.count_batch(batch); ref.finalize(); refs.append(ref)
        d = str(tmp_path / ("s%d.kreeq" % i)); dirs.append(d)
        with kb.DBG(k=k) as g:
            g.traverseInReads(batch); g.finalize(); g.save(d)
        disk = ko.read_db(d, check=True)                          # strict phmap layout check
        assert_same_records(disk.export(), ref.export())
        assert disk.stats() == ref.stats()
        with kb.DBG(k=k) as g2:
            g2.load(d)
            assert_same_records(g2.export(), ref.export())
            assert g2.stats() == ref.stats()
    for order in ([0, 1, 2], [1, 0, 2], [2, 1, 0]):
        oracle_tabs = [ko.read_db(dirs[i], check=False) for i in order]
        u = ko.union(oracle_tabs)
        with kb.DBG(k=k) as g:
            g.kunion([dirs[i] for i in order], ref_compatible=True)
            assert_same_records(g.export(), u.export())
            assert g.stats() == u.stats()
            seg = genome.tobytes()[:5000] + b"A" * 60
            out = g.evaluateSegment(seg)
            rc, rout, racc = u.validate_segment(seg)
            assert out.tobytes() == rout.tobytes()
            ud = str(tmp_path / ("u%s.kreeq" % "".join(map(str, order))))
            g.save(ud)
        disk = ko.read_db(ud, check=True)
        k8a, r8a = disk.dump8(); k8b, r8b = u.dump8()
        assert np.array_equal(k8a, k8b)
        k32a, r32a = disk.dump32(); k32b, r32b = u.dump32()
        assert np.array_equal(k32a, k32b) and r32a.tobytes() == r32b.tobytes()
        assert disk.stats() == u.stats()


def test_sharded_handles_equal_single(mods):
    """2 and 4 shards (handles on one GPU): union of shard exports == single-handle result."""
    kb, ko = mods
    k = 21
    rng = np.random.default_rng(31)
    genome = random_genome(rng, 40000)
    batch = sample_reads(rng, genome, 3000, 150, err=0.01)
    ref = ko.Table(k); ref.count_batch(batch)
    rk, rr = ref.export()
    for nr in (2, 4):
        ks, rs, inst = [], [], 0
        outs, accs = [], np.zeros(3, dtype=np.int64)
        for r in range(nr):
            with kb.DBG(k=k, rank=r, nranks=nr) as g:
                g.traverseInReads(batch); g.finalize()
                kk, rec = g.export()
                assert np.all(((kk % np.uint64(128)) // np.uint64(128 // nr)) == r)
                ks.append(kk); rs.append(rec); inst += g.instances()
                outs.append(g.evaluateSegment(genome.tobytes()))
                accs += np.array([g.tot_missing, g.tot_kcount, g.tot_edge_missing])
        kk = np.concatenate(ks); rec = np.concatenate(rs)
        o = np.argsort(kk)
        assert np.array_equal(kk[o], rk) and rec[o].tobytes() == rr.tobytes()
        rc, rout, racc = ref.validate_segment(genome.tobytes())
        merged = np.zeros(rout.size, dtype=rout.dtype)
        for f in ("fw", "bw", "cov", "isFw"):
            merged[f] = sum(o_[f].astype(np.uint64) for o_ in outs).astype(merged[f].dtype)
        assert merged.tobytes() == rout.tobytes()
        assert tuple(int(x) for x in accs) == racc


def test_find_lookup_batch(mods):
    kb, ko = mods
    rng = np.random.default_rng(41)
    genome = random_genome(rng, 5000)
    batch = sample_reads(rng, genome, 400, 100, err=0.0)
    ref = ko.Table(21); ref.count_batch(batch)
    rk, rr = ref.export()
    with kb.DBG(k=21) as g:
        g.traverseInReads(batch); g.finalize()
        q = np.concatenate([rk[:500], rng.integers(0, 1 << 42, 500, dtype=np.uint64)])
        rec, found = g.find(q)
        assert found[:500].all()
        assert rec[:500].tobytes() == rr[:500].tobytes()
        present = np.isin(q[500:], rk)
        assert np.array_equal(found[500:], present)


@pytest.mark.parametrize("k,hint", [(21, 6_000_000), (31, 6_000_000), (21, 0)])
def test_partitioned_pipeline_multi_slice(mods, k, hint):
    """Table of several 32 MiB slices: histogram -> scatter -> L2-resident insert (8- and 16-byte tuples),
    with a size hint (one big batch) and without (planned batches + growth)."""
    kb, ko = mods
    rng = np.random.default_rng(51 + k)
    genome = random_genome(rng, 2_500_000)
    batches = [sample_reads(rng, genome, 50_000, 150, err=0.01) for _ in range(2)]
    ref = ko.Table(k); n_ref = sum(ref.count_batch(b) for b in batches)
    with kb.DBG(k=k, expected_distinct=hint) as g:
        for b in batches:
            g.traverseInReads(b)
        g.finalize()
        assert g.instances() == n_ref
        assert_same_records(g.export(), ref.export())
        assert g.stats() == ref.stats()
        out = g.evaluateSegment(genome.tobytes()[:1_000_000])
        rc, rout, racc = ref.validate_segment(genome.tobytes()[:1_000_000])
        assert out.tobytes() == rout.tobytes()
        assert (g.tot_missing, g.tot_kcount, g.tot_edge_missing) == racc


def test_wrong_size_hint_overflow_list(mods):
    """A size hint that is too small: slices fill up, the tuples that found no room are kept, the
    table is doubled and they are inserted afterwards -- nothing may be lost or counted twice."""
    kb, ko = mods
    k = 21
    rng = np.random.default_rng(61)
    genome = random_genome(rng, 2_600_000)
    batch = sample_reads(rng, genome, 70_000, 150, err=0.01)
    ref = ko.Table(k); n_ref = ref.count_batch(batch)
    with kb.DBG(k=k, expected_distinct=2_000_000) as g:           # ~4.4 M distinct arrive
        g.traverseInReads(batch); g.finalize()
        assert g.instances() == n_ref
        assert_same_records(g.export(), ref.export())
    with kb.DBG(k=k, expected_distinct=300_000) as g:             # single-slice table, fused kernel
        g.traverseInReads(batch[:151 * 20000]); g.finalize()
        ref2 = ko.Table(k); ref2.count_batch(batch[:151 * 20000])
        assert_same_records(g.export(), ref2.export())


def test_small_pushes_accumulate(mods):
    """Many small pushes join the pending stream (64-base aligned, invalid separators) and are counted
    together; a low flush threshold forces several pipeline runs."""
    kb, ko = mods
    k = 21
    rng = np.random.default_rng(71)
    genome = random_genome(rng, 50_000)
    ref = ko.Table(k)
    with kb.DBG(k=k) as g:
        g.set_flush_positions(200_000)
        n_ref = 0
        for i in range(40):
            n_reads = int(rng.integers(1, 300)); L = int(rng.integers(21, 200))
            b = sample_reads(rng, genome, n_reads, L, err=0.01, n_rate=0.001)
            if i % 7 == 0:
                b = b[:-1]                                          # batch without the trailing 'N'
            if i % 5 == 0:
                b = b[:(len(b) // 64) * 64] if len(b) >= 64 else b  # length a multiple of 64
            n_ref += ref.count_batch(b)
            g.traverseInReads(b)
        g.finalize()
        assert g.instances() == n_ref
        assert_same_records(g.export(), ref.export())
        assert g.stats() == ref.stats()


def test_routed_tuples_two_shards_one_gpu(mods):
    """kq_dev_extract_route -> (exchange stand-in) -> kq_dev_insert_tuples with two sharded handles."""
    import ctypes as C
    import torch
    kb, ko = mods
    from kreeq_b200._lib import check, lib
    L = lib()
    for k in (21, 31):
        rng = np.random.default_rng(81 + k)
        genome = random_genome(rng, 200_000)
        batch = sample_reads(rng, genome, 8000, 150, err=0.01)
        ref = ko.Table(k); ref.count_batch(batch)
        rk, rr = ref.export()
        bases, inv = kb.pack_host(batch)
        d_b = torch.from_numpy(bases.view(np.int64)).cuda(); d_i = torch.from_numpy(inv.view(np.int64)).cuda()
        npos = len(batch) - k + 1
        nr = 2
        gs = [kb.DBG(k=k, rank=r, nranks=nr) for r in range(nr)]
        tb = L.kq_tuple_bytes(gs[0].handle)
        assert tb == (8 if k <= 28 else 16)
        cap = npos
        slabs = torch.zeros((nr, cap, tb), dtype=torch.uint8, device="cuda")
        counts = torch.zeros(nr, dtype=torch.int64, device="cuda")
        check(L.kq_dev_extract_route(gs[0].handle, C.c_void_p(d_b.data_ptr()), C.c_void_p(d_i.data_ptr()), len(batch), 0, npos,
                                     C.c_void_p(slabs.data_ptr()), cap, C.c_void_p(counts.data_ptr())), gs[0].handle)
        gs[0].sync()
        cnt = counts.cpu().numpy()
        ks, rs = [], []
        for r in range(nr):
            check(L.kq_dev_insert_tuples(gs[r].handle, C.c_void_p(slabs[r].data_ptr()), int(cnt[r])), gs[r].handle)
            gs[r].finalize()
            assert gs[r].instances() == cnt[r]
            kk, rec = gs[r].export(); ks.append(kk); rs.append(rec)
            gs[r].close()
        kk = np.concatenate(ks); rec = np.concatenate(rs); o = np.argsort(kk)
        assert np.array_equal(kk[o], rk) and rec[o].tobytes() == rr.tobytes()


@pytest.mark.parametrize("k,nr,hint", [(21, 2, 0), (31, 4, 0), (21, 4, 2_000_000)])
def test_peer_routed_pipeline_shards_one_gpu(mods, k, nr, hint):
    """kq_route_hist -> (all-gather stand-in) -> kq_route_scatter into every owner's slab -> kq_route_insert,
    with nr sharded handles in one process (slab pointers passed directly instead of through CUDA IPC).
    Every "rank" routes its own read batch, in two rounds; the union of the shards equals the oracle."""
    import ctypes as C
    import torch
    kb, ko = mods
    from kreeq_b200._lib import check, lib
    L = lib()
    rng = np.random.default_rng(500 + k + nr)
    genome = random_genome(rng, 300_000)
    batches = [sample_reads(rng, genome, 6000, 150, err=0.01) for _ in range(nr)]
    ref = ko.Table(k)
    for b in batches:
        ref.count_batch(b)
    rk, rr = ref.export()
    gs = [kb.DBG(k=k, rank=r, nranks=nr, expected_distinct=hint) for r in range(nr)]
    dev = []
    for b in batches:
        bases, inv = kb.pack_host(b)
        dev.append((torch.from_numpy(bases.view(np.int64)).cuda(), torch.from_numpy(inv.view(np.int64)).cuda(), len(b)))
    slab_cap = 2 * max(len(b) for b in batches)
    ptrs = (C.c_void_p * nr)()
    for r in range(nr):
        p = C.c_void_p()
        check(L.kq_route_setup(gs[r].handle, slab_cap, 6, C.byref(p), None), gs[r].handle)
        ptrs[r] = p.value
    for r in range(nr):
        check(L.kq_route_connect(gs[r].handle, ptrs), gs[r].handle)
    P = L.kq_route_bins(gs[0].handle)
    assert P == nr << 6
    H = torch.zeros((nr, P), dtype=torch.int64, device="cuda")
    sent = 0
    for rnd in range(2):
        spans = []
        for r in range(nr):
            d_b, d_i, n = dev[r]
            total = n - k + 1
            p0 = rnd * (total // 2)
            npos = total // 2 if rnd == 0 else total - p0
            spans.append((p0, npos))
            check(L.kq_route_hist(gs[r].handle, C.c_void_p(d_b.data_ptr()), C.c_void_p(d_i.data_ptr()), n, p0, npos,
                                  C.c_void_p(H[r].data_ptr())), gs[r].handle)
        for g in gs:
            g.sync()
        nrecv = []
        for r in range(nr):
            d_b, d_i, n = dev[r]
            out = C.c_uint64()
            check(L.kq_route_scatter(gs[r].handle, C.c_void_p(d_b.data_ptr()), C.c_void_p(d_i.data_ptr()), n, spans[r][0], spans[r][1],
                                     C.c_void_p(H.data_ptr()), C.byref(out)), gs[r].handle)
            nrecv.append(out.value)
        for g in gs:
            g.sync()
        Hh = H.cpu().numpy().reshape(nr, nr, P // nr)
        for r in range(nr):                    # slab extent = real tuples + alignment padding (< 32 per sender and bin)
            real = int(Hh[:, r, :].sum())
            assert real <= nrecv[r] <= real + 32 * P
        sent += int(Hh.sum())
        for r in range(nr):
            check(L.kq_route_insert(gs[r].handle, nrecv[r]), gs[r].handle)
        for g in gs:
            g.sync()
    ks, rs = [], []
    for g in gs:
        g.finalize()
        kk, rec = g.export(); ks.append(kk); rs.append(rec)
    assert sum(g.instances() for g in gs) == sent == int(ref.stats()["total"])
    for g in gs:
        g.close()
    kk = np.concatenate(ks); rec = np.concatenate(rs); o = np.argsort(kk)
    assert np.array_equal(kk[o], rk) and rec[o].tobytes() == rr.tobytes()
    # a slab that is too small is refused on every rank before anything is written
    g2 = [kb.DBG(k=k, rank=r, nranks=2) for r in range(2)]
    p2 = (C.c_void_p * 2)()
    for r in range(2):
        p = C.c_void_p()
        check(L.kq_route_setup(g2[r].handle, 1000, 4, C.byref(p), None), g2[r].handle)
        p2[r] = p.value
    H2 = torch.zeros((2, 32), dtype=torch.int64, device="cuda")
    d_b, d_i, n = dev[0]
    for r in range(2):
        check(L.kq_route_connect(g2[r].handle, p2), g2[r].handle)
    check(L.kq_route_hist(g2[0].handle, C.c_void_p(d_b.data_ptr()), C.c_void_p(d_i.data_ptr()), n, 0, n - k + 1,
                          C.c_void_p(H2[0].data_ptr())), g2[0].handle)
    out = C.c_uint64()
    assert L.kq_route_scatter(g2[0].handle, C.c_void_p(d_b.data_ptr()), C.c_void_p(d_i.data_ptr()), n, 0, n - k + 1,
                              C.c_void_p(H2.data_ptr()), C.byref(out)) == -4
    for g in g2:
        g.close()


@pytest.mark.parametrize("hint", [0, 400_000])
def test_every_kmer_high_copy(mods, hint):
    """300x coverage: every genomic k-mer passes 127 (slot lanes move to the remainder table 63 at a time)
    and most pass 254 (the reference's 32-bit spill); counted in one large batch so the remainder table
    must have been sized before the kernels ran."""
    kb, ko = mods
    k = 21
    rng = np.random.default_rng(4242)
    genome = random_genome(rng, 60_000)
    batch = sample_reads(rng, genome, 120_000, 150, err=0.002)
    ref = ko.Table(k); ref.count_batch(batch)
    with kb.DBG(k=k, expected_distinct=hint) as g:
        g.traverseInReads(batch)
        g.finalize()
        assert_same_records(g.export(), ref.export())
        assert g.stats() == ref.stats()
        out = g.evaluateSegment(genome.tobytes())
        _, rout, racc = ref.validate_segment(genome.tobytes())
        assert out.tobytes() == rout.tobytes()
        assert (g.tot_missing, g.tot_kcount, g.tot_edge_missing) == racc


def test_peer_tables_validate_shards_one_gpu(mods):
    """kq_peer_connect_local: a sharded handle answers every position of a segment by probing the shard that
    owns each k-mer (high-copy keys included: remainder table of the owner), equal to the oracle."""
    import ctypes as C
    kb, ko = mods
    from kreeq_b200._lib import check, lib
    L = lib()
    k, nr = 21, 4
    rng = np.random.default_rng(77)
    genome = random_genome(rng, 150_000)
    batch = sample_reads(rng, genome, 20_000, 150, err=0.01) + (genome[:300].tobytes() + b"N") * 400
    ref = ko.Table(k); ref.count_batch(batch)
    gs = [kb.DBG(k=k, rank=r, nranks=nr) for r in range(nr)]
    for g in gs:
        g.traverseInReads(batch); g.finalize()
    arr = (C.c_void_p * nr)(*[g.handle for g in gs])
    seg = genome.tobytes()
    _, rout, racc = ref.validate_segment(seg, cutoff=2)
    for r in (0, 3):
        check(L.kq_peer_connect_local(gs[r].handle, arr), gs[r].handle)
        out = gs[r].evaluateSegment(seg, cov_cutoff=2)
        assert out.tobytes() == rout.tobytes()
        assert (gs[r].tot_missing, gs[r].tot_kcount, gs[r].tot_edge_missing) == racc
    with pytest.raises(kb.KreeqError):                     # an exported table must not be counted into (it could move)
        gs[1].traverseInReads(sample_reads(rng, random_genome(rng, 3_000_000), 40_000, 150, err=0.0)); gs[1].finalize()
    for g in gs:
        check(L.kq_peer_disconnect(g.handle), g.handle)
    # without peers a sharded handle only answers what it owns
    own = gs[0].evaluateSegment(seg, cov_cutoff=2)
    assert own.tobytes() != rout.tobytes()
    for g in gs:
        g.close()


def test_validate_into_page_locked_output(mods):
    """kq_validate_segment copies rows straight into page-locked caller memory and stages them for pageable
    memory; a segment of several staging chunks (4 Mi rows each) must come out identical both ways and equal
    the oracle."""
    import torch
    kb, ko = mods
    rng = np.random.default_rng(909)
    genome = random_genome(rng, 300_000)
    batch = sample_reads(rng, genome, 20_000, 150, err=0.01)
    ref = ko.Table(21); ref.count_batch(batch)
    asm = np.concatenate([genome] * 30 + [genome[:123_457]])           # 9.1 M positions: three chunks
    asm = asm.copy(); asm[rng.integers(0, asm.size, 500)] = ord("C")
    with kb.DBG(k=21) as g:
        g.traverseInReads(batch); g.finalize()
        pageable = g.evaluateSegment(asm)
        a = (g.tot_missing, g.tot_kcount, g.tot_edge_missing)
        g.tot_missing = g.tot_kcount = g.tot_edge_missing = 0
        pinned_t = torch.empty(asm.size * kb.BASE.itemsize, dtype=torch.uint8, pin_memory=True)
        pinned_t.fill_(0xAB)
        pinned = g.evaluateSegment(asm, out=pinned_t.numpy().view(kb.BASE))
        assert (g.tot_missing, g.tot_kcount, g.tot_edge_missing) == a
        assert pinned.tobytes() == pageable.tobytes()
        short = g.evaluateSegment(asm[:1000], out=pinned_t.numpy().view(kb.BASE))   # reuse, shorter segment
        assert short.tobytes() == g.evaluateSegment(asm[:1000]).tobytes() and short[980:].tobytes() == bytes(16 * 20)
    rc, rout, racc = ref.validate_segment(asm[:600_000].tobytes())
    assert rout.tobytes()[:16 * 500_000] == pageable.tobytes()[:16 * 500_000]


def test_union_of_moderately_covered_databases(mods, tmp_path):
    """Two read sets of one 150 kbp genome at ~90x each: every genomic key is below a slot lane (127) in both
    databases and passes it in the union, so ~1.5e5 keys get a remainder record in ONE merge launch (the side
    table starts at 4096 entries), and three databases take the sums past 254 (placeholders + 32-bit records).
    The databases are large enough to be decoded into table order (>= 65536 records)."""
    kb, ko = mods
    k = 21
    rng = np.random.default_rng(1234)
    genome = random_genome(rng, 150_000)
    dirs = []
    for i in range(3):
        batch = sample_reads(rng, genome, 90_000, 150, err=0.0)
        d = str(tmp_path / ("c%d.kreeq" % i)); dirs.append(d)
        with kb.DBG(k=k, expected_distinct=400_000) as g:
            g.traverseInReads(batch); g.finalize(); g.save(d)
            st = g.stats()
            assert st["distinct"] > 140_000 and 60 < st["total"] / st["distinct"] < 127
    for order in ([0, 1], [0, 1, 2], [2, 0, 1]):
        u = ko.union([ko.read_db(dirs[i], check=False) for i in order])
        with kb.DBG(k=k) as g:
            g.kunion([dirs[i] for i in order], ref_compatible=True)
            assert_same_records(g.export(), u.export())
            assert g.stats() == u.stats()
            seg = genome.tobytes()[:20000]
            out = g.evaluateSegment(seg)
            rc, rout, racc = u.validate_segment(seg)
            assert out.tobytes() == rout.tobytes()


def test_compact_validate_rows(mods):
    """kq_validate_segment_compact + kq_expand_rows == kq_validate_segment, bit for bit, including rows whose coverage
    or edge counts pass 254 (wide list), several staged chunks, and the list-too-small error"""
    import ctypes as C
    kb, ko = mods
    from kreeq_b200._lib import BASE, BASE4, WIDE, KQ_ENOMEM, KqQV, check, lib
    L = lib()
    k = 21
    rng = np.random.default_rng(77)
    genome = random_genome(rng, 60_000)
    hot = random_genome(rng, 300).tobytes()
    batch = sample_reads(rng, genome, 9000, 150, err=0.01) + (hot + b"N") * 400 + (b"A" * 300 + b"N") * 3
    seg = genome.tobytes() + hot + b"A" * 200 + genome.tobytes()[:9_000_000 // 1000]
    seg = seg * 80                                                     # ~5.5 M positions: two staged chunks
    with kb.DBG(k=k, expected_distinct=600_000) as g:
        g.traverseInReads(batch); g.finalize()
        a = np.frombuffer(seg, dtype=np.uint8)
        full = np.zeros(a.size, dtype=BASE); acc = KqQV()
        check(L.kq_validate_segment(g.handle, C.c_void_p(a.ctypes.data), a.size, 0, C.c_void_p(full.ctypes.data), C.byref(acc)), g.handle)
        assert int(full["cov"].max()) > 254
        out4 = np.zeros(a.size, dtype=BASE4); wide = np.zeros(1 << 20, dtype=WIDE); nw = C.c_size_t(wide.size); acc4 = KqQV()
        check(L.kq_validate_segment_compact(g.handle, C.c_void_p(a.ctypes.data), a.size, 0, C.c_void_p(out4.ctypes.data),
                                            C.c_void_p(wide.ctypes.data), C.byref(nw), C.byref(acc4)), g.handle)
        assert (acc4.missing, acc4.kcount, acc4.edge_missing) == (acc.missing, acc.kcount, acc.edge_missing)
        assert 0 < nw.value == int(((out4["flags"] & 2) != 0).sum())
        back = np.zeros(a.size, dtype=BASE)
        check(L.kq_expand_rows(C.c_void_p(out4.ctypes.data), a.size, C.c_void_p(wide.ctypes.data), nw.value, C.c_void_p(back.ctypes.data)))
        assert back.tobytes() == full.tobytes()
        small = C.c_size_t(3)
        rc = L.kq_validate_segment_compact(g.handle, C.c_void_p(a.ctypes.data), a.size, 0, C.c_void_p(out4.ctypes.data),
                                           C.c_void_p(wide.ctypes.data), C.byref(small), C.byref(acc4))
        assert rc == KQ_ENOMEM and small.value == nw.value

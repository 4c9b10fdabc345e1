"""GPU parity of the two-level scatter + shared-memory insert (kq_count3.cu) against the oracle.

The geometry knobs are turned down (kq_set_tuning) so that small inputs run through every part of the big-table
pipeline: many slices, several tasks per coarse bin, ring bins that wrap, slices that fill up, high-copy keys."""
import numpy as np
import pytest

from helpers import assert_same_records, random_genome, sample_reads

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import kreeq_b200
    import kreeq_oracle as ko
    return kreeq_b200, ko


@pytest.fixture()
def tuned():
    """set knobs for one test, restore the defaults afterwards"""
    from kreeq_b200._lib import lib
    L = lib()
    defaults = {"single_slice_bits": 21, "piece_bits": 12, "fine_part": 1 << 18, "smem_insert": 1, "direct_pct": 25, "warp_insert": 1, "one_pass": 0, "cta_lock": 0}

    def set_(**kw):
        for n, v in kw.items():
            assert L.kq_set_tuning(n.encode(), int(v)) == 0, n
    yield set_
    set_(**defaults)


def _run(kb, ko, k, batches, hint, genome=None):
    ref = ko.Table(k)
    n_ref = sum(ref.count_batch(b) for b in batches)
    with kb.DBG(k=k, expected_distinct=hint) as g:
        for b in batches:
            g.traverseInReads(b)
        g.finalize()
        assert g.instances() == n_ref
        assert_same_records(g.export(), ref.export())
        assert g.stats() == ref.stats()
        if genome is not None:
            out = g.evaluateSegment(genome)
            rc, rout, racc = ref.validate_segment(genome)
            assert out.tobytes() == rout.tobytes()
            assert (g.tot_missing, g.tot_kcount, g.tot_edge_missing) == racc


@pytest.mark.parametrize("k,piece,part", [(21, 8, 4096), (31, 10, 64), (15, 6, 1024), (32, 12, 1 << 18), (21, 12, 1 << 18), (21, 9, 1 << 18), (32, 9, 4096)])
def test_small_slices_match_oracle(mods, tuned, k, piece, part):
    kb, ko = mods
    tuned(single_slice_bits=8, piece_bits=piece, fine_part=part, direct_pct=0)
    rng = np.random.default_rng(900 + k + piece)
    genome = random_genome(rng, 120_000)
    batches = [sample_reads(rng, genome, 6000, 150, err=0.01, n_rate=0.001) for _ in range(2)]
    _run(kb, ko, k, batches, 700_000, genome.tobytes())


def test_small_slices_high_copy_and_runs(mods, tuned):
    """poly-A runs and satellites: the instances of one key sit next to each other in the tuple stream (warp
    combining), pass 127 (remainder table) and 65535"""
    kb, ko = mods
    tuned(single_slice_bits=8, piece_bits=8, fine_part=2048, direct_pct=0)
    rng = np.random.default_rng(5)
    unit = random_genome(rng, 60).tobytes()
    genome = random_genome(rng, 30_000)
    batch = (b"A" * 500 + b"N") * 3 + (unit + b"N") * 300 + (b"ACGT" * 20000 + b"N") + sample_reads(rng, genome, 1500, 150, err=0.01)
    _run(kb, ko, 21, [batch, batch[:200_000]], 200_000, b"A" * 100 + unit + b"ACGT" * 30)


def test_small_slices_wrong_hint(mods, tuned):
    """a size hint that is far too small: slices fill up inside the shared-memory insert, the instances that found
    no room wait in the level-1 buffer, the table grows and they are counted afterwards"""
    kb, ko = mods
    tuned(single_slice_bits=8, piece_bits=8, fine_part=4096, direct_pct=0)
    rng = np.random.default_rng(61)
    genome = random_genome(rng, 400_000)
    batch = sample_reads(rng, genome, 12_000, 150, err=0.01)
    _run(kb, ko, 21, [batch], 60_000)           # ~0.75 M distinct arrive


@pytest.mark.parametrize("k,piece", [(21, 8), (31, 9)])
def test_one_pass_level1_with_overflowing_regions(mods, tuned, k, piece):
    """one_pass = 2: level-1 scatter into fixed regions with (almost) no slack: many coarse bins overflow, their chunks
    are counted directly by the scatter kernel, region tails are filled with sentinels; plus runs of one key"""
    kb, ko = mods
    tuned(single_slice_bits=8, piece_bits=piece, fine_part=2048, direct_pct=0, one_pass=2)
    rng = np.random.default_rng(31 + k)
    genome = random_genome(rng, 150_000)
    batches = [sample_reads(rng, genome, 7000, 150, err=0.01, n_rate=0.001) + (b"A" * 3000 + b"N") * 4,
               sample_reads(rng, genome, 3000, 150, err=0.02)]
    _run(kb, ko, k, batches, 900_000, genome.tobytes())


@pytest.mark.parametrize("cta_lock", [1, 0])
def test_cta_owned_slices(mods, tuned, cta_lock):
    """warp_insert = 0: small slices counted by the CTA-owned kernels (lock bytes + barriers / shared-memory CAS)"""
    kb, ko = mods
    tuned(single_slice_bits=8, piece_bits=8, fine_part=2048, direct_pct=0, warp_insert=0, cta_lock=cta_lock)
    rng = np.random.default_rng(78)
    genome = random_genome(rng, 100_000)
    batch = (b"A" * 400 + b"N") * 3 + sample_reads(rng, genome, 5000, 150, err=0.01)
    _run(kb, ko, 21, [batch], 600_000, genome.tobytes())


def test_l2_insert_still_available(mods, tuned):
    """smem_insert = 0: level-1 scatter + L2 CAS insert on the same small-slice table"""
    kb, ko = mods
    tuned(single_slice_bits=8, piece_bits=8, smem_insert=0, direct_pct=0)
    rng = np.random.default_rng(77)
    genome = random_genome(rng, 100_000)
    _run(kb, ko, 21, [sample_reads(rng, genome, 5000, 150, err=0.01)], 600_000, genome.tobytes())


def test_property_checks_at_scale(mods):
    """size-independent properties at a size the oracle cannot reach (3e8 instances, default geometry):
    instances == reads x (L - k + 1), stats.total == instances, counting the read set twice doubles total and
    leaves distinct unchanged, host-ASCII pushes and the device-resident path agree."""
    import ctypes as C
    import torch
    kb, _ = mods
    from kreeq_b200 import synth
    from kreeq_b200._lib import check, lib
    L = lib()
    k, rl, pairs, gsize = 21, 150, 1_200_000, 12_000_000
    dev = torch.device("cuda", 0)
    genome = synth.random_genome_codes(gsize, 9, dev)
    reads = synth.paired_reads_ascii(genome, pairs, rl, 0.01, 9).cpu().numpy()
    inst = 2 * pairs * (rl - k + 1)
    with kb.DBG(k=k, expected_distinct=45_000_000) as g:
        g.traverseInReads(reads); g.finalize()
        s1 = g.stats()
        assert g.instances() == inst and s1["total"] == inst
        g.traverseInReads(reads); g.finalize()
        s2 = g.stats()
        assert s2["total"] == 2 * inst and s2["distinct"] == s1["distinct"] and g.instances() == 2 * inst
        bases, inv = kb.pack_host(reads)
        d_b = torch.from_numpy(bases.view(np.int64)).to(dev); d_i = torch.from_numpy(inv.view(np.int64)).to(dev)
        g.clear()
        check(L.kq_dev_count_packed(g.handle, C.c_void_p(d_b.data_ptr()), C.c_void_p(d_i.data_ptr()), reads.size, 0, reads.size - k + 1), g.handle)
        g.finalize()
        s3 = g.stats()
        assert s3 == s1

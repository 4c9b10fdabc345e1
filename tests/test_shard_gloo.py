"""Host-side logic of the multi-GPU path on CPU: world_size 2 over gloo.  Each rank extracts the
(key, edgeBit) tuples of ITS read batch (oracle, standing in for kq_dev_extract_route), routes them
by owner(key) = (key % 128) / (128 / nranks) with kreeq_b200.shard's all-to-all, inserts what it
receives, and the union of the two shards must equal the single-table result (SURVEY 8e invariant)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import kreeq_oracle as ko
    from helpers import random_genome, sample_reads
    from kreeq_b200 import shard
    k = 21
    rng = np.random.default_rng(77)
    genome = random_genome(rng, 20000)
    batches = [sample_reads(np.random.default_rng(100 + r), genome, 600, 120, err=0.01) + (b"A" * 300 + b"N") * 2 for r in range(world)]
    keys, edges = ko.extract(k, batches[rank])
    owner = shard.owner_of(keys, world)
    cap = int(max(np.bincount(owner, minlength=world))) + 8
    raw = shard.pack_tuples(keys, edges, k)                       # the layout kq_dev_extract_route writes
    send = torch.zeros((world, cap, raw.shape[1]), dtype=torch.uint8)
    counts = torch.zeros(world, dtype=torch.int64)
    for r in range(world):
        sel = owner == r
        n = int(sel.sum()); counts[r] = n
        send[r, :n] = torch.from_numpy(raw[sel])
    recv_counts = shard.exchange_counts(counts)
    recv = torch.zeros((int(recv_counts.sum()) + 4, raw.shape[1]), dtype=torch.uint8)
    n = shard.exchange_tuples(send, counts.numpy(), recv_counts.numpy(), recv)
    assert n == int(recv_counts.sum())
    rk, re_ = shard.unpack_tuples(recv[:n].numpy(), k)
    assert np.all(shard.owner_of(rk, world) == rank)
    t = ko.Table(k)
    t.insert_tuples(rk, re_)
    ek, er = t.export()
    # validate: every rank probes the whole assembly against its shard; counters are summed, per-base rows merged
    rc, out, acc = t.validate_segment(genome.tobytes())
    # positions whose key another shard owns must not count here: mask them (what kq_validate_segment does on a shard)
    gk, _ = ko.extract(k, genome.tobytes())
    mine = shard.owner_of(gk, world) == rank
    miss = int(((out["cov"][:gk.size] == 0) & mine).sum())
    acc_t = torch.tensor([miss, int(mine.sum())], dtype=torch.int64)
    dist.all_reduce(acc_t)
    np.savez(os.path.join(out_dir, "shard%d.npz" % rank), keys=ek, recs=er, acc=acc_t.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_routing_equals_single_table(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import kreeq_oracle as ko
    from helpers import random_genome, sample_reads
    rng = np.random.default_rng(77)
    genome = random_genome(rng, 20000)
    ref = ko.Table(21)
    for r in range(world):
        ref.count_batch(sample_reads(np.random.default_rng(100 + r), genome, 600, 120, err=0.01) + (b"A" * 300 + b"N") * 2)
    rk, rr = ref.export()
    parts = [np.load(os.path.join(tmp_path, "shard%d.npz" % r)) for r in range(world)]
    keys = np.concatenate([p["keys"] for p in parts]); recs = np.concatenate([p["recs"] for p in parts])
    o = np.argsort(keys)
    assert np.array_equal(keys[o], rk) and recs[o].tobytes() == rr.tobytes()
    rc, out, acc = ref.validate_segment(genome.tobytes())
    assert tuple(parts[0]["acc"]) == (acc[0], acc[1]) == tuple(parts[1]["acc"])


def test_cut_rounds_groups_segments_up_to_the_round_capacity():
    """host logic of the routed validate rounds (RoutedValidator.rounds_of): order kept, capacity respected"""
    import pytest
    from kreeq_b200.shard import cut_rounds
    segs = [(None, None, n) for n in (10, 20, 30, 5, 40, 40, 1)]
    rounds = cut_rounds(segs, 60)
    assert [[s[2] for s in r] for r in rounds] == [[10, 20, 30], [5, 40], [40, 1]]
    assert cut_rounds([], 60) == []
    assert [[s[2] for s in r] for r in cut_rounds(segs, 1000)] == [[10, 20, 30, 5, 40, 40, 1]]
    with pytest.raises(ValueError):
        cut_rounds([(None, None, 61)], 60)
